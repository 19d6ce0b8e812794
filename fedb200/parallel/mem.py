"""Peer memory: one VMM allocation per rank, mapped into every process of the box.

Layout of a rank's allocation (``PeerLayout``), identical on every rank so peers address each other
by offset:

    [ ctrl 4 KiB | client fp32 arena | client int64 | aux fp32 arena | aux int64 | q (u8) | scale (f32) ]

* client arena: the worker's live model - what the reference uploads as a base64 checkpoint
  (src/client.py:19-23) is simply readable here by K-AGG.
* aux arena: on the primary's rank it is the server's global model, on the backup's rank the
  replica (reference ``Primary/optimizedModel.pth`` / ``Backup/optimizedModel.pth``).
* q / scale: the fp8 block-scaled upload of the opt-in ``-c`` mode.

``PeerGroup`` does the exchange for ranks started by torchrun: every rank exports its allocation as a
POSIX fd, serves it on a unix socket (SCM_RIGHTS) and maps everyone else's.  When the fabric supports
NVLS, rank 0 also creates a multicast object over all allocations.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from dataclasses import dataclass
from typing import List, Optional

import torch

from ..ops import _lib
from .arena import Arena, Manifest

CTRL_REGION = 4096


class _CudaPtr:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap a raw device pointer."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def tensor_from_ptr(ptr: int, numel: int, dtype: torch.dtype, device: int) -> torch.Tensor:
    """A torch view (no ownership) of ``numel`` elements of ``dtype`` at device address ``ptr``."""
    nbytes = numel * torch.empty((), dtype=dtype).element_size()
    raw = torch.as_tensor(_CudaPtr(ptr, nbytes), device=torch.device("cuda", device))
    return raw.view(dtype)


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


@dataclass(frozen=True)
class PeerLayout:
    n_f32: int
    n_i64: int
    with_codec: bool = True

    @property
    def off_client_f32(self): return CTRL_REGION
    @property
    def off_client_i64(self): return _align(self.off_client_f32 + self.n_f32 * 4)
    @property
    def off_aux_f32(self): return _align(self.off_client_i64 + self.n_i64 * 8)
    @property
    def off_aux_i64(self): return _align(self.off_aux_f32 + self.n_f32 * 4)
    @property
    def off_q(self): return _align(self.off_aux_i64 + self.n_i64 * 8)
    @property
    def off_scale(self): return _align(self.off_q + (self.n_f32 if self.with_codec else 0))
    @property
    def nbytes(self): return _align(self.off_scale + (self.n_f32 // 32 * 4 if self.with_codec else 0))

    @classmethod
    def for_manifest(cls, man: Manifest, with_codec: bool = True) -> "PeerLayout":
        return cls(man.n_f32, man.n_i64, with_codec)


class PeerBuffer:
    """One rank's allocation as seen from THIS process (local or imported)."""

    def __init__(self, base_ptr: int, layout: PeerLayout, device: int, owner_rank: int):
        self.base, self.layout, self.device, self.owner_rank = base_ptr, layout, device, owner_rank

    @property
    def ctrl_ptr(self): return self.base
    def client_f32_ptr(self): return self.base + self.layout.off_client_f32
    def client_i64_ptr(self): return self.base + self.layout.off_client_i64
    def aux_f32_ptr(self): return self.base + self.layout.off_aux_f32
    def aux_i64_ptr(self): return self.base + self.layout.off_aux_i64
    def q_ptr(self): return self.base + self.layout.off_q
    def scale_ptr(self): return self.base + self.layout.off_scale

    def ctrl(self) -> torch.Tensor:
        return tensor_from_ptr(self.base, _lib.CTRL_BYTES // 4, torch.int32, self.device)

    def client_f32(self): return tensor_from_ptr(self.client_f32_ptr(), self.layout.n_f32, torch.float32, self.device)
    def client_i64(self): return tensor_from_ptr(self.client_i64_ptr(), self.layout.n_i64, torch.int64, self.device)
    def aux_f32(self): return tensor_from_ptr(self.aux_f32_ptr(), self.layout.n_f32, torch.float32, self.device)
    def aux_i64(self): return tensor_from_ptr(self.aux_i64_ptr(), self.layout.n_i64, torch.int64, self.device)
    def q(self): return tensor_from_ptr(self.q_ptr(), self.layout.n_f32, torch.uint8, self.device)
    def scale(self): return tensor_from_ptr(self.scale_ptr(), self.layout.n_f32 // 32, torch.float32, self.device)

    def client_arena(self, man: Manifest) -> Arena:
        return Arena(man, torch.device("cuda", self.device), f32=self.client_f32(), i64=self.client_i64())

    def aux_arena(self, man: Manifest) -> Arena:
        return Arena(man, torch.device("cuda", self.device), f32=self.aux_f32(), i64=self.aux_i64())


def supported(device: int = 0):
    """(vmm, posix_fd, multicast) capability flags of ``device``."""
    vmm, fd_ok, mc = C.c_int(), C.c_int(), C.c_int()
    _lib.check(_lib.lib().fedb_mem_supported(device, C.byref(vmm), C.byref(fd_ok), C.byref(mc)), "fedb_mem_supported")
    return bool(vmm.value), bool(fd_ok.value), bool(mc.value)


def fs_min_barrier(sock_dir: str, tag: str, rank: int, world: int, timeout_s: float = 120.0):
    """all-reduce(MIN) of a 0/1 flag over files in ``sock_dir`` - the rendezvous of processes that share a box but no
    torch.distributed group (gRPC-driven workers).  Returns ``fn(flag_tensor)`` with the same contract as the
    torch.distributed version used by torchrun jobs; every call is a new stage."""
    import time
    stage = [0]

    def fn(flag: torch.Tensor) -> None:
        stage[0] += 1
        mine = os.path.join(sock_dir, f"fedb200-{tag}-bar{stage[0]}-{rank}")
        with open(mine + ".tmp", "w") as f:
            f.write(str(int(flag.item())))
        os.replace(mine + ".tmp", mine)
        end = time.time() + timeout_s
        vals = {}
        while len(vals) < world:
            for r in range(world):
                if r in vals:
                    continue
                try:
                    with open(os.path.join(sock_dir, f"fedb200-{tag}-bar{stage[0]}-{r}")) as f:
                        vals[r] = int(f.read().strip() or 0)
                except (OSError, ValueError):
                    pass
            if len(vals) < world:
                if time.time() > end:
                    vals = {r: 0 for r in range(world)}       # somebody never showed up: nobody uses the feature
                    break
                time.sleep(0.01)
        flag.fill_(min(vals.values()))
    return fn


class PeerGroup:
    """All ranks' allocations mapped here.  ``buffers[r]`` is rank r's memory; ``mc_base`` is the
    multicast alias of the whole layout (0 when NVLS is unavailable or disabled)."""

    def __init__(self, rank: int, world: int, device: int, layout: PeerLayout, tag: str,
                 sock_dir: str = "/tmp", use_multicast: bool = True, timeout_ms: int = 120_000, barrier=None,
                 persistent_server: bool = False, rejoin: bool = False):
        L = _lib.lib()
        self.rank, self.world, self.device, self.layout = rank, world, device, layout
        self.local = _lib.Mem()
        _lib.check(L.fedb_mem_create(layout.nbytes, device, C.byref(self.local)), "fedb_mem_create")
        _lib.check(L.fedb_mem_export_fd(C.byref(self.local)), "fedb_mem_export_fd")
        self.alloc_bytes = int(self.local.bytes)
        self.imported: List[Optional[_lib.Mem]] = [None] * world
        self.mc = None
        self.mc_base = 0
        vmm, fd_ok, mc_ok = supported(device)
        want_mc = bool(use_multicast and mc_ok and world > 1 and not rejoin)
        if want_mc and barrier is None:
            barrier = fs_min_barrier(sock_dir, tag, rank, world)
        path = lambda r: os.path.join(sock_dir, f"fedb200-{tag}-{r}.sock")   # noqa: E731

        fds = [self.local.fd]
        mc_err = None
        if want_mc and rank == 0:
            self.mc = _lib.Mem()
            rc = L.fedb_mc_create(self.alloc_bytes, world, C.byref(self.mc))
            if rc != 0:
                mc_err = L.fedb_last_error().decode()
                self.mc = None
        # rank 0 always serves two fds (second = multicast handle or a dup of the first as "none")
        serve_fds = list(fds)
        if rank == 0:
            serve_fds.append(self.mc.fd if self.mc is not None else self.local.fd)
        arr = (C.c_int * len(serve_fds))(*serve_fds)
        err: List[Optional[str]] = [None]

        self._path = path
        self._serve_arr = arr            # keep the ctypes array alive for a persistent server

        def serve():
            # persistent mode (gRPC-driven workers): keep answering so that a restarted peer, or survivors
            # re-mapping a restarted peer, can fetch this allocation at any time
            n_clients, tmo = (1 << 30, 0x7FFFFFFF) if persistent_server else (world - 1, timeout_ms)
            rc = L.fedb_fd_serve(path(rank).encode(), arr, len(serve_fds), n_clients, tmo)
            if rc != 0:
                err[0] = L.fedb_last_error().decode()

        th = None
        if world > 1:
            th = threading.Thread(target=serve, daemon=True)
            th.start()
        mc_fd = -1
        self.missing: List[int] = []
        for r in range(world):
            if r == rank:
                continue
            n = 2 if r == 0 else 1
            got = (C.c_int * n)()
            rc = L.fedb_fd_fetch(path(r).encode(), got, n, 5000 if rejoin else timeout_ms)
            if rc != 0:
                if rejoin:               # a peer that is down right now: mapped later through remap()
                    self.missing.append(r)
                    continue
                _lib.check(rc, f"fd fetch from rank {r}")
            m = _lib.Mem()
            _lib.check(L.fedb_mem_import_fd(got[0], self.alloc_bytes, device, C.byref(m)), f"import rank {r}")
            os.close(got[0])
            self.imported[r] = m
            if r == 0:
                mc_fd = got[1]
        if th is not None and not persistent_server:
            th.join()
            if err[0]:
                raise _lib.FedbError(err[0])
        self.buffers = [PeerBuffer(int(self.local.ptr if r == rank else (self.imported[r].ptr if self.imported[r] else 0)),
                                   layout, device, r) for r in range(world)]

        # ---- multicast: every rank adds its device, barrier, every rank binds + maps
        self.multicast_error = mc_err
        if want_mc:
            flag = torch.tensor([1 if (rank != 0 or self.mc is not None) else 0], device=f"cuda:{device}")
            if barrier is not None:
                barrier(flag)            # all-reduce(min): did rank 0 manage to create the object?
            ok = bool(flag.item())
            if ok and rank != 0:
                self.mc = _lib.Mem()
                ok = L.fedb_mc_import_fd(mc_fd, self.alloc_bytes, C.byref(self.mc)) == 0
            if ok:
                ok = L.fedb_mc_add_device(C.byref(self.mc), device) == 0
            flag.fill_(1 if ok else 0)
            if barrier is not None:
                barrier(flag)
            if bool(flag.item()):
                rc = L.fedb_mc_bind_and_map(C.byref(self.mc), C.byref(self.local), device)
                flag.fill_(1 if rc == 0 else 0)
                if rc != 0:
                    self.multicast_error = L.fedb_last_error().decode()
                if barrier is not None:
                    barrier(flag)
                if bool(flag.item()):
                    self.mc_base = int(self.mc.ptr)
            elif self.multicast_error is None:
                self.multicast_error = L.fedb_last_error().decode()
        if mc_fd >= 0:
            try:
                os.close(mc_fd)
            except OSError:
                pass

    @property
    def me(self) -> PeerBuffer:
        return self.buffers[self.rank]

    def remap(self, r: int, timeout_ms: int = 30_000) -> None:
        """Peer ``r`` was restarted (new process, new allocation): fetch and map its memory again."""
        L = _lib.lib()
        n = 2 if r == 0 else 1
        got = (C.c_int * n)()
        _lib.check(L.fedb_fd_fetch(self._path(r).encode(), got, n, timeout_ms), f"fd fetch from restarted rank {r}")
        m = _lib.Mem()
        _lib.check(L.fedb_mem_import_fd(got[0], self.alloc_bytes, self.device, C.byref(m)), f"import rank {r}")
        for fd in got:
            try:
                os.close(fd)
            except OSError:
                pass
        old = self.imported[r]
        self.imported[r] = m
        self.buffers[r] = PeerBuffer(int(m.ptr), self.layout, self.device, r)
        if old is not None:
            L.fedb_mem_release(C.byref(old))
        if r in self.missing:
            self.missing.remove(r)

    def mc_client_f32_ptr(self) -> int:
        return self.mc_base + self.layout.off_client_f32 if self.mc_base else 0

    def close(self):
        L = _lib.lib()
        for m in self.imported:
            if m is not None:
                L.fedb_mem_release(C.byref(m))
        L.fedb_mem_release(C.byref(self.local))
