"""Python face of the sm_100a kernels: thin, explicit wrappers over the C ABI.

Every function launches on the CURRENT torch stream (so ``torch.cuda.graph`` capture records it) and
bumps ``_lib.launches`` - the ``gpu_launches`` figure bench.py reports.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import AggDesc, LeNetDesc, check, current_stream, lib, ptr

DEFAULT_TIMEOUT_NS = 20_000_000_000   # bounded spins: 20 s, then status != 0 instead of a hang


# ------------------------------------------------------------------------------------ control blocks
def new_ctrl(device) -> torch.Tensor:
    """A zeroed ``fedb_ctrl_t`` in ordinary device memory (single-process use)."""
    return torch.zeros(_lib.CTRL_BYTES // 4, dtype=torch.int32, device=device)


def ctrl_field(ctrl: torch.Tensor, name: str) -> int:
    off = _lib.CTRL_OFF[name] // 4
    return int(ctrl[off].item()) & 0xFFFFFFFF


# ------------------------------------------------------------------------------------ K-AGG / K-BCAST
def fedavg_raw(src_f32: Sequence[int], dst_f32: Sequence[int], begin: int, end: int, *, ctrl_local: int,
               src_i64: Sequence[int] = (), dst_i64: Sequence[int] = (), n_i64: int = 0,
               weights: Optional[Sequence[float]] = None, dst_bf16: Sequence[int] = (),
               src_q: Sequence[int] = (), src_scale: Sequence[int] = (), base: int = 0,
               mc_src: int = 0, mc_dst: int = 0, sync_mode: int = 0, rank: int = 0, world: int = 1,
               ctrl_peer: Sequence[int] = (), timeout_ns: int = DEFAULT_TIMEOUT_NS, version: int = 0,
               dst_ctrl: Sequence[int] = (), max_blocks: int = 0, stream: Optional[int] = None, gate: int = 0) -> None:
    """Pointer-level entry (addresses as ints): used by the peer plane where sources and destinations
    are mappings of other processes' memory."""
    d = AggDesc()
    n_src = len(src_q) if src_q else len(src_f32)
    if not 1 <= n_src <= _lib.MAX_PEERS:
        raise ValueError(f"fedavg over {n_src} sources (1..{_lib.MAX_PEERS})")
    if len(dst_f32) > _lib.MAX_DST:
        raise ValueError(f"too many destinations ({len(dst_f32)} > {_lib.MAX_DST})")
    for i, p in enumerate(src_f32):
        d.src[i] = p
    for i, p in enumerate(src_i64):
        d.src_i64[i] = p
    for i, p in enumerate(dst_f32):
        d.dst[i] = p
    for i, p in enumerate(dst_i64):
        d.dst_i64[i] = p
    for i, p in enumerate(dst_bf16):
        d.dst_bf16[i] = p
    for i, p in enumerate(src_q):
        d.src_q[i] = p
    for i, p in enumerate(src_scale):
        d.src_scale[i] = p
    for i, p in enumerate(ctrl_peer):
        d.ctrl_peer[i] = p
    for i, p in enumerate(dst_ctrl):
        d.dst_ctrl[i] = p
    uniform = weights is None or len(set(float(w) for w in weights)) <= 1
    if not uniform:
        for i, w in enumerate(weights):
            d.weight[i] = float(w)
    d.n_src, d.n_dst, d.n_dst_i64, d.uniform = n_src, len(dst_f32), len(dst_i64), int(uniform)
    d.do_i64 = int(n_i64 > 0 and len(src_i64) == n_src)
    d.begin, d.end, d.n_i64 = begin, end, n_i64
    d.base, d.mc_src, d.mc_dst = base, mc_src, mc_dst
    d.sync_mode, d.rank, d.world = sync_mode, rank, world
    d.ctrl_local, d.timeout_ns, d.version = ctrl_local, timeout_ns, version
    d.gate = gate
    check(lib().fedb_fedavg(C.byref(d), max_blocks, current_stream() if stream is None else stream), "fedb_fedavg")
    _lib.count_launch()


_local_ctrl = {}


def _ctrl_for(device) -> torch.Tensor:
    key = torch.device(device).index or 0
    if key not in _local_ctrl:
        _local_ctrl[key] = new_ctrl(device)
    return _local_ctrl[key]


def fedavg(out, clients, weights=None, extra_dst=(), stream=None, bf16_shadows=None) -> None:
    """out <- FedAvg(clients) for ``Arena`` objects that live on one GPU (or are mapped into this
    process).  The result is ALSO stored into every arena of ``extra_dst`` (broadcast / replica)."""
    man = out.manifest
    dsts = [out] + list(extra_dst)
    fedavg_raw([c.f32.data_ptr() for c in clients], [a.f32.data_ptr() for a in dsts], 0, man.n_f32,
               ctrl_local=_ctrl_for(out.f32.device).data_ptr(),
               src_i64=[c.i64.data_ptr() for c in clients], dst_i64=[a.i64.data_ptr() for a in dsts],
               n_i64=man.n_i64, weights=weights,
               dst_bf16=[ptr(b) for b in bf16_shadows] if bf16_shadows else (), stream=stream)


def broadcast(src, dsts, bf16_shadows=None, stream=None) -> None:
    """K-BCAST alone: one source arena to many (re-admission, promotion)."""
    man = src.manifest
    fedavg_raw([src.f32.data_ptr()], [a.f32.data_ptr() for a in dsts], 0, man.n_f32,
               ctrl_local=_ctrl_for(src.f32.device).data_ptr(), src_i64=[src.i64.data_ptr()],
               dst_i64=[a.i64.data_ptr() for a in dsts], n_i64=man.n_i64,
               dst_bf16=[ptr(b) for b in bf16_shadows] if bf16_shadows else (), stream=stream)


def replicate(dst, src, dst_ctrl: torch.Tensor, version: int, stream: Optional[int] = None) -> None:
    """K-REPL: versioned arena copy (primary -> backup)."""
    man = src.manifest
    check(lib().fedb_replicate(dst.f32.data_ptr(), src.f32.data_ptr(), man.n_f32, dst.i64.data_ptr(),
                               src.i64.data_ptr(), man.n_i64, dst_ctrl.data_ptr(), version,
                               current_stream() if stream is None else stream), "fedb_replicate")
    _lib.count_launch()


def push_copy(src_ptr: int, dst_ptrs: Sequence[int], nbytes: int, mc_dst: int = 0, stream: Optional[int] = None) -> None:
    """K-PUSH: bit-exact copy of ``nbytes`` at ``src_ptr`` into every peer pointer (or through the multicast alias)."""
    arr = (C.c_void_p * max(len(dst_ptrs), 1))(*dst_ptrs)
    check(lib().fedb_push_copy(src_ptr, arr, len(dst_ptrs), mc_dst or None, nbytes,
                               current_stream() if stream is None else stream), "fedb_push_copy")
    _lib.count_launch()


# ------------------------------------------------------------------------------------ K-CODEC
def codec_encode(cur: torch.Tensor, base: torch.Tensor, residual: torch.Tensor, q: torch.Tensor,
                 scale: torch.Tensor) -> None:
    n = cur.numel()
    assert q.numel() >= n and scale.numel() >= (n + 31) // 32 and residual.numel() >= n
    check(lib().fedb_codec_encode(cur.data_ptr(), base.data_ptr(), residual.data_ptr(), q.data_ptr(),
                                  scale.data_ptr(), n, current_stream()), "fedb_codec_encode")
    _lib.count_launch()


def codec_decode(out: torch.Tensor, base: torch.Tensor, q: torch.Tensor, scale: torch.Tensor) -> None:
    check(lib().fedb_codec_decode(out.data_ptr(), base.data_ptr(), q.data_ptr(), scale.data_ptr(), out.numel(),
                                  current_stream()), "fedb_codec_decode")
    _lib.count_launch()


# ------------------------------------------------------------------------------------ K10
def sgd_step(param: torch.Tensor, grad: torch.Tensor, momentum: torch.Tensor, lr: float, mu: float,
             weight_decay: float, shadow_bf16: Optional[torch.Tensor] = None) -> None:
    check(lib().fedb_sgd_step(param.data_ptr(), grad.data_ptr(), momentum.data_ptr(), ptr(shadow_bf16),
                              param.numel(), lr, mu, weight_decay, current_stream()), "fedb_sgd_step")
    _lib.count_launch()


def sgd_step_end(param: torch.Tensor, grad: torch.Tensor, momentum: torch.Tensor, lr: float, mu: float,
                 weight_decay: float, counters: Optional[torch.Tensor] = None) -> None:
    """SGD step that also clears ``grad`` and advances the arena's int64 counters (BatchNorm steps) by one."""
    nc = 0 if counters is None else counters.numel()
    check(lib().fedb_sgd_step_end(param.data_ptr(), grad.data_ptr(), momentum.data_ptr(), param.numel(), lr, mu,
                                  weight_decay, ptr(counters) if nc else 0, nc, current_stream()), "fedb_sgd_step_end")
    _lib.count_launch()


def zero_(t: torch.Tensor) -> torch.Tensor:
    """Zero-fill of a contiguous tensor as a memset on the current stream (a memset node under graph capture)."""
    assert t.is_contiguous() or (t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last))
    check(lib().fedb_zero_async(t.data_ptr(), t.numel() * t.element_size(), current_stream()), "fedb_zero_async")
    return t


def adam_step(param, grad, m, v, step_dev: torch.Tensor, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8,
              weight_decay=0.0) -> None:
    check(lib().fedb_adam_step(param.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), param.numel(),
                               lr, beta1, beta2, eps, weight_decay, step_dev.data_ptr(), current_stream()),
          "fedb_adam_step")
    _lib.count_launch()


# ------------------------------------------------------------------------------------ K9 / K13
def cross_entropy_raw(logits: torch.Tensor, target: torch.Tensor, dlogits: Optional[torch.Tensor],
                      loss_out: Optional[torch.Tensor], stats: Optional[torch.Tensor]) -> None:
    B, Cn = logits.shape
    check(lib().fedb_cross_entropy(logits.data_ptr(), target.data_ptr(), ptr(dlogits), ptr(loss_out), ptr(stats),
                                   B, Cn, current_stream()), "fedb_cross_entropy")
    _lib.count_launch(2)


def cross_entropy_static_raw(logits: torch.Tensor, target: torch.Tensor, dlogits: Optional[torch.Tensor],
                             loss_out: Optional[torch.Tensor], stats: Optional[torch.Tensor]) -> None:
    """``cross_entropy_raw`` for hosts that cannot allocate device memory at call time (a GPU already
    fully owned by the caller's pool, or a latency-critical loop where cudaMalloc would synchronise the
    device): the library's row scratch is a fixed 128 KiB array, so B is limited to 16384 rows and calls
    that may overlap on one GPU must be ordered by the caller.  Make the first call on a device outside
    stream capture (it looks up the array's address once per device)."""
    B, Cn = logits.shape
    fn = _lib.bind("fedb_cross_entropy_static", C.c_int,
                   [C.c_void_p] * 5 + [C.c_int, C.c_int, C.c_void_p])
    check(fn(logits.data_ptr(), target.data_ptr(), ptr(dlogits), ptr(loss_out), ptr(stats), B, Cn,
             current_stream()), "fedb_cross_entropy_static")
    _lib.count_launch(2)


class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        logits = logits.contiguous()
        dlogits = torch.empty_like(logits)
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        cross_entropy_raw(logits, target, dlogits, loss, None)
        ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        return dlogits * g, None


def cross_entropy(logits, target):
    return _CrossEntropy.apply(logits, target)


# ------------------------------------------------------------------------------------ K14
def prepare_batch(images_u8: torch.Tensor, nhwc: bool = False, augment: bool = False, seed: int = 0,
                  seed_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``seed_dev``: int64 device word mixed into the seed at run time (stays live across CUDA-graph replays)."""
    B = images_u8.shape[0]
    out = torch.empty((B, 32, 32, 3) if nhwc else (B, 3, 32, 32), dtype=torch.float32, device=images_u8.device)
    check(lib().fedb_prepare_batch_dev(images_u8.data_ptr(), out.data_ptr(), B, int(nhwc), int(augment), seed,
                                       ptr(seed_dev), current_stream()), "fedb_prepare_batch_dev")
    _lib.count_launch()
    return out.permute(0, 3, 1, 2) if nhwc else out


# ------------------------------------------------------------------------------------ LeNet fused step
LENET_KEYS = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "fc1.weight", "fc1.bias",
              "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias")


class LeNetPlan:
    """Binds a LeNet arena to the fused kernels of csrc/lenet.cu and owns their workspace."""

    def __init__(self, arena, lr=0.1, momentum=0.9, weight_decay=5e-4, augment=False, max_batch=128):
        man = arena.manifest
        if [e.key for e in man.entries] != list(LENET_KEYS):
            raise ValueError("LeNetPlan needs a LeNet arena")
        self.arena = arena
        self.desc = LeNetDesc()
        self.desc.param = arena.f32.data_ptr()
        self.desc.momentum = ptr(arena.momentum)
        for i, k in enumerate(LENET_KEYS):
            self.desc.off[i] = man.by_key[k].offset
        self.desc.lr, self.desc.mu, self.desc.weight_decay = lr, momentum, weight_decay
        self.desc.augment, self.desc.seed = int(augment), 0
        self.max_batch = 0
        self.workspace = None          # training workspace: its address is baked into captured graphs
        self.eval_batch_cap = 0
        self.eval_workspace = None     # separate, so growing it never invalidates a training graph
        self._aux_cap, self._aux_workspace = 0, None
        self.stats = torch.zeros(4, dtype=torch.float32, device=arena.f32.device)   # loss, correct, samples, steps
        self.reserve(max_batch)

    def reserve(self, B: int) -> None:
        if B > self.max_batch:
            nbytes = lib().fedb_lenet_workspace_bytes(B)
            self.workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.arena.f32.device)
            self.max_batch = B

    def train_step(self, images_u8: torch.Tensor, labels: torch.Tensor, step_seed: int = 0) -> None:
        B = images_u8.shape[0]
        assert B <= self.max_batch
        check(lib().fedb_lenet_train_step(C.byref(self.desc), images_u8.data_ptr(), labels.data_ptr(), B,
                                          self.workspace.data_ptr(), self.stats.data_ptr(), step_seed,
                                          current_stream()), "fedb_lenet_train_step")
        _lib.count_launch(2)

    def _eval_ws(self, B: int) -> torch.Tensor:
        """Evaluation only needs the two per-sample rows (loss, correct)."""
        if B > self.eval_batch_cap:
            self.eval_workspace = torch.empty(lib().fedb_lenet_eval_bytes(B), dtype=torch.uint8,
                                              device=self.arena.f32.device)
            self.eval_batch_cap = B
        return self.eval_workspace

    def _aux_ws(self, B: int) -> torch.Tensor:
        """Full per-sample workspace for forward() / grad(): separate from the training one, whose address is baked
        into captured graphs."""
        if B > self._aux_cap:
            self._aux_workspace = torch.empty(lib().fedb_lenet_workspace_bytes(B), dtype=torch.uint8,
                                              device=self.arena.f32.device)
            self._aux_cap = B
        return self._aux_workspace

    def bind_eval_params(self, f32: torch.Tensor) -> None:
        """Evaluation reads its weights from ``f32`` (same layout as the arena) instead of the live arena: the snapshot
        of a global model that is scored on a second stream while the next local pass already trains the arena."""
        d = LeNetDesc()
        C.memmove(C.byref(d), C.byref(self.desc), C.sizeof(LeNetDesc))
        d.param = f32.data_ptr()
        self._eval_desc, self._eval_params = d, f32

    def eval_batch(self, images_u8: torch.Tensor, labels: torch.Tensor, stats: Optional[torch.Tensor] = None,
                   snapshot: bool = False) -> None:
        B = images_u8.shape[0]
        check(lib().fedb_lenet_eval(C.byref(self._eval_desc if snapshot else self.desc), images_u8.data_ptr(), labels.data_ptr(), B,
                                    self._eval_ws(B).data_ptr(), ptr(self.stats if stats is None else stats),
                                    current_stream()), "fedb_lenet_eval")
        _lib.count_launch(2)

    def eval_queue(self, images_u8: torch.Tensor, labels: torch.Tensor, counter: torch.Tensor, ctas: int,
                   snapshot: bool = False) -> None:
        """One launch of the work-queue evaluation (csrc/lenet.cu, lenet_eval_kernel<true>): ``ctas`` persistent CTAs score
        chunks of the batch until ``counter`` (int32 device word, zeroed by the caller once per evaluation) runs out."""
        B = images_u8.shape[0]
        check(lib().fedb_lenet_eval_queue(C.byref(self._eval_desc if snapshot else self.desc), images_u8.data_ptr(),
                                          labels.data_ptr(), B, self._eval_ws(B).data_ptr(), counter.data_ptr(), ctas,
                                          current_stream()), "fedb_lenet_eval_queue")
        _lib.count_launch()

    def eval_queue_stats(self, B: int, stats: torch.Tensor) -> None:
        check(lib().fedb_lenet_eval_stats(self._eval_ws(B).data_ptr(), B, stats.data_ptr(), current_stream()),
              "fedb_lenet_eval_stats")
        _lib.count_launch()

    def forward(self, images: torch.Tensor) -> torch.Tensor:
        B = images.shape[0]
        logits = torch.empty(B, 10, dtype=torch.float32, device=images.device)
        u8 = images.dtype == torch.uint8
        check(lib().fedb_lenet_forward(C.byref(self.desc), images.data_ptr() if u8 else 0,
                                       0 if u8 else images.contiguous().data_ptr(), B, self._aux_ws(B).data_ptr(),
                                       logits.data_ptr(), current_stream()), "fedb_lenet_forward")
        _lib.count_launch(1)
        return logits

    def grad(self, images_f32: torch.Tensor, labels: torch.Tensor):
        B = images_f32.shape[0]
        dparam = torch.zeros(self.arena.manifest.n_param, dtype=torch.float32, device=images_f32.device)
        loss = torch.zeros(4, dtype=torch.float32, device=images_f32.device)
        check(lib().fedb_lenet_grad(C.byref(self.desc), images_f32.contiguous().data_ptr(), labels.data_ptr(), B,
                                    self._aux_ws(B).data_ptr(), dparam.data_ptr(), loss.data_ptr(),
                                    current_stream()), "fedb_lenet_grad")
        _lib.count_launch(2)
        return dparam, loss[0]


# ------------------------------------------------------------------------------------ generic model ops
def __getattr__(name):
    # conv2d / batch_norm / pool2d / ... of the wider zoo live in fedb200.ops.cuda_nn (csrc/nn_*.cu)
    from . import cuda_nn
    try:
        return getattr(cuda_nn, name)
    except AttributeError:
        raise AttributeError(f"fedb200.ops.cuda has no kernel named {name!r}") from None
