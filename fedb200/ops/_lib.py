"""ctypes binding of ``libfedb200.so`` (C ABI declared in include/fedb200.h).

The library is built in-tree by ``make`` / ``__graft_entry__.build()`` and loaded from
``fedb200/lib``.  There is no fallback: if a CUDA device is present and the library is missing
or fails to load, ``lib()`` raises - a GPU box must run the hand-written kernels, not PyTorch's.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libfedb200.so")

MAX_PEERS = 16
MAX_DST = 20

u8p, f32p, i64p, u16p, u32p = (C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_int64),
                               C.POINTER(C.c_uint16), C.POINTER(C.c_uint32))
vp = C.c_void_p


class Ctrl(C.Structure):
    _fields_ = [("arrive", C.c_uint32 * MAX_PEERS), ("done", C.c_uint32 * MAX_PEERS),
                ("epoch", C.c_uint32), ("status", C.c_uint32), ("blocks_done", C.c_uint32),
                ("abort_flag", C.c_uint32), ("version", C.c_uint32), ("reserved", C.c_uint32 * 27)]


CTRL_BYTES = C.sizeof(Ctrl)
CTRL_OFF = {name: getattr(Ctrl, name).offset for name, _ in Ctrl._fields_}


class AggDesc(C.Structure):
    _fields_ = [
        ("src", vp * MAX_PEERS), ("src_i64", vp * MAX_PEERS),
        ("dst", vp * MAX_DST), ("dst_i64", vp * MAX_DST), ("dst_bf16", vp * MAX_DST),
        ("weight", C.c_float * MAX_PEERS),
        ("n_src", C.c_int32), ("n_dst", C.c_int32), ("n_dst_i64", C.c_int32), ("uniform", C.c_int32),
        ("do_i64", C.c_int32),
        ("begin", C.c_int64), ("end", C.c_int64), ("n_i64", C.c_int64),
        ("src_q", vp * MAX_PEERS), ("src_scale", vp * MAX_PEERS), ("base", vp),
        ("mc_src", vp), ("mc_dst", vp),
        ("sync_mode", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
        ("ctrl_local", vp), ("ctrl_peer", vp * MAX_PEERS),
        ("timeout_ns", C.c_uint64), ("version", C.c_uint32), ("dst_ctrl", vp * MAX_DST),
        ("gate", vp),
    ]


class LeNetDesc(C.Structure):
    _fields_ = [("param", vp), ("momentum", vp), ("off", C.c_int64 * 10),
                ("lr", C.c_float), ("mu", C.c_float), ("weight_decay", C.c_float),
                ("augment", C.c_int32), ("seed", C.c_uint64), ("seed_dev", vp)]


class Mem(C.Structure):
    _fields_ = [("handle", C.c_uint64), ("ptr", C.c_uint64), ("bytes", C.c_uint64),
                ("device", C.c_int32), ("fd", C.c_int32)]


_SIGS = {
    "fedb_last_error": (C.c_char_p, []),
    "fedb_version": (C.c_int, []),
    "fedb_fedavg": (C.c_int, [C.POINTER(AggDesc), C.c_int, vp]),
    "fedb_replicate": (C.c_int, [vp, vp, C.c_int64, vp, vp, C.c_int64, vp, C.c_uint32, vp]),
    "fedb_flag_barrier": (C.c_int, [vp, C.POINTER(vp), C.c_int, C.c_int, C.c_uint64, vp]),
    "fedb_codec_encode": (C.c_int, [vp, vp, vp, vp, vp, C.c_int64, vp]),
    "fedb_codec_decode": (C.c_int, [vp, vp, vp, vp, C.c_int64, vp]),
    "fedb_sgd_step": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_float, C.c_float, C.c_float, vp]),
    "fedb_sgd_step_end": (C.c_int, [vp, vp, vp, C.c_int64, C.c_float, C.c_float, C.c_float, vp, C.c_int64, vp]),
    "fedb_adam_step": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float,
                                 C.c_float, vp, vp]),
    "fedb_cross_entropy": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int, vp]),
    "fedb_prepare_batch": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_uint64, vp]),
    "fedb_prepare_batch_dev": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_uint64, vp, vp]),
    "fedb_lenet_workspace_bytes": (C.c_size_t, [C.c_int]),
    "fedb_lenet_eval_bytes": (C.c_size_t, [C.c_int]),
    "fedb_lenet_train_step": (C.c_int, [C.POINTER(LeNetDesc), vp, vp, C.c_int, vp, vp, C.c_uint64, vp]),
    "fedb_lenet_eval": (C.c_int, [C.POINTER(LeNetDesc), vp, vp, C.c_int, vp, vp, vp]),
    "fedb_lenet_eval_queue": (C.c_int, [C.POINTER(LeNetDesc), vp, vp, C.c_int, vp, vp, C.c_int, vp]),
    "fedb_lenet_eval_stats": (C.c_int, [vp, C.c_int, vp, vp]),
    "fedb_lenet_forward": (C.c_int, [C.POINTER(LeNetDesc), vp, vp, C.c_int, vp, vp, vp]),
    "fedb_lenet_grad": (C.c_int, [C.POINTER(LeNetDesc), vp, vp, C.c_int, vp, vp, vp, vp]),
    "fedb_mem_supported": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "fedb_mem_create": (C.c_int, [C.c_uint64, C.c_int, C.POINTER(Mem)]),
    "fedb_mem_export_fd": (C.c_int, [C.POINTER(Mem)]),
    "fedb_mem_import_fd": (C.c_int, [C.c_int, C.c_uint64, C.c_int, C.POINTER(Mem)]),
    "fedb_mem_release": (C.c_int, [C.POINTER(Mem)]),
    "fedb_mc_create": (C.c_int, [C.c_uint64, C.c_int, C.POINTER(Mem)]),
    "fedb_mc_import_fd": (C.c_int, [C.c_int, C.c_uint64, C.POINTER(Mem)]),
    "fedb_mc_add_device": (C.c_int, [C.POINTER(Mem), C.c_int]),
    "fedb_mc_bind_and_map": (C.c_int, [C.POINTER(Mem), C.POINTER(Mem), C.c_int]),
    "fedb_fd_serve": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int]),
    "fedb_fd_fetch": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.c_int]),
    "fedb_copy_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "fedb_zero_async": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    "fedb_push_copy": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_uint64, C.c_void_p]),
}

_lib = None
_lock = threading.Lock()
launches = 0          # number of fedb200 kernel-launching calls made by this process (bench evidence)


class FedbError(RuntimeError):
    pass


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise FedbError(f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()) first. "
                                    "fedb200 has no eager-PyTorch fallback on CUDA tensors.")
                handle = C.CDLL(LIB_PATH)
                for name, (res, args) in _SIGS.items():
                    fn = getattr(handle, name, None)
                    if fn is None:
                        continue     # optional entry points are bound by their own modules
                    fn.restype, fn.argtypes = res, args
                _lib = handle
    return _lib


def bind(name: str, restype, argtypes):
    """Declare an entry point that is not in the table above (extension modules)."""
    fn = getattr(lib(), name)
    fn.restype, fn.argtypes = restype, argtypes
    return fn


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().fedb_last_error().decode(errors="replace")
        raise FedbError(f"{what or 'libfedb200'} failed with code {rc}: {msg}")


def count_launch(n: int = 1) -> None:
    global launches
    launches += n


def ptr(t) -> int:
    """Device address of a tensor (None -> NULL)."""
    return 0 if t is None else t.data_ptr()


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
