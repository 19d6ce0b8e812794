"""Autograd bindings of the generic NHWC kernels (csrc/nn_generic.cu) and the tcgen05 GEMM/conv
paths (csrc/gemm_tc.cu) - the op set behind ``fedb200.ops.functional`` on CUDA.

Layout contract: activations are 4-D tensors in ``channels_last`` memory format (NHWC bytes, NCHW
shape), fp32.  Every function makes its inputs channels_last-contiguous (a no-op when they already
are) and returns channels_last outputs.  Weights keep the checkpoint's OIHW layout.
"""
from __future__ import annotations

import ctypes as C
from typing import List

import torch

from . import _lib
from ._lib import check, current_stream, lib

CL = torch.channels_last
ACT = {"none": 0, "relu": 1, "swish": 2, "sigmoid": 3}
_BN_CHUNKS = 296          # upper bound of row chunks per BatchNorm reduction (scratch = 2 * C * chunks floats): 2 per SM
                          # (more chunks make the second, per-channel pass longer than the streaming pass it serves)

_vp, _i, _ll, _f = C.c_void_p, C.c_int, C.c_longlong, C.c_float
_BOUND = False


def _bind():
    global _BOUND
    if _BOUND:
        return
    ip = C.POINTER(C.c_int)
    sigs = {
        "fedb_nn_weight_transpose": [_vp, _vp, _vp, _i, _i, _i, _i, _vp],
        "fedb_nn_conv_fwd": [_vp, _vp, _vp, _vp, ip, _vp],
        "fedb_nn_conv_dgrad": [_vp, _vp, _vp, ip, _vp],
        "fedb_nn_conv_wgrad": [_vp, _vp, _vp, _vp, ip, _vp],
        "fedb_nn_bn_fwd16": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _f, _f, _i, _i, _vp, _ll, _vp, _vp],
        "fedb_nn_bn_bwd16": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _i, _vp, _ll, _vp, _vp],
        "fedb_nn_act_fwd": [_vp, _vp, _ll, _i, _vp],
        "fedb_nn_act_bwd": [_vp, _vp, _vp, _ll, _i, _vp],
        "fedb_nn_add_act16": [_vp, _vp, _vp, _ll, _i, _vp, _vp],
        "fedb_nn_scale": [_vp, _vp, _vp, _ll, _ll, _i, _i, _vp],
        "fedb_nn_scale_bwd_gate": [_vp, _vp, _vp, _i, _i, _i, _vp],
        "fedb_nn_bernoulli_mask": [_vp, _ll, _f, C.c_uint64, _vp, _vp],
        "fedb_nn_pool_fwd": [_vp, _vp, _vp, ip, _i, _vp],
        "fedb_nn_pool_bwd": [_vp, _vp, _vp, ip, _i, _vp],
        "fedb_nn_channel_copy": [_vp, _vp, _ll, _i, _i, _i, _i, _i, _i, _i, _vp],
        "fedb_nn_linear_fwd": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
        "fedb_nn_linear_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
        "fedb_dw_conv_fwd_k": [_vp, _vp, _vp, _vp, ip, _i, _vp],
        "fedb_dw_conv_dgrad_k": [_vp, _vp, _vp, ip, _i, _vp],
        "fedb_dw_conv_wgrad_k": [_vp, _vp, _vp, _vp, ip, _i, _vp],
    }
    for name, args in sigs.items():
        _lib.bind(name, C.c_int, args)
    _lib.bind("fedb_dw_conv_wgrad_scratch_k", C.c_longlong, [ip, _i])
    _BOUND = True


def _call(name, *args, launches=1):
    _bind()
    check(getattr(lib(), name)(*args), name)
    _lib.count_launch(launches)


def _cl(x: torch.Tensor) -> torch.Tensor:
    return x.contiguous(memory_format=CL) if x.dim() == 4 else x.contiguous()


def _p(t):
    return 0 if t is None else t.data_ptr()


def _ints(*v):
    return (C.c_int * len(v))(*v)


def _train_scope():
    """The training-step scope of cuda_gemm when one is open (engine's step), else None."""
    from . import cuda_gemm
    sc = cuda_gemm.scope()
    return sc if (sc is not None and sc.train) else None


def _direct_grad(p):
    """Inside a training scope parameter gradients are written straight into ``p.grad`` - the parameter's slot of the
    gradient arena, zeroed at the start of the step - and autograd is handed None: no temporary, no accumulation
    kernel.  Outside (plain ``loss.backward()`` of user code) the gradient is returned as usual."""
    if p is None or _train_scope() is None:
        return None
    g = p.grad
    return g if (g is not None and g.is_contiguous() and g.dtype == torch.float32) else None


def _shadow(t: torch.Tensor):
    """bf16 buffer for the shadow copy a producer kernel writes next to its fp32 result (training scope only; the
    flat size must allow 8-byte stores)."""
    if _train_scope() is None or t.dim() != 4 or t.shape[1] % 8 != 0:
        return None
    return torch.empty(t.numel(), dtype=torch.bfloat16, device=t.device)


# ------------------------------------------------------------------------------------ convolution
def _conv_shape(x, w, stride, padding, groups):
    N, Cc, H, W = x.shape
    K, Cg, R, S = w.shape
    P = (H + 2 * padding[0] - R) // stride[0] + 1
    Q = (W + 2 * padding[1] - S) // stride[1] + 1
    return (N, H, W, Cc, K, R, S, P, Q, stride[0], stride[1], padding[0], padding[1], groups)


def _is_depthwise(shp) -> bool:
    """3x3 / 5x5 / 7x7 depthwise (groups == Cin == Cout), channels a multiple of 4: the streaming kernels of
    csrc/depthwise.cu."""
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    return groups == Cc == K and R == S and R in (3, 5, 7) and Cc % 4 == 0


def _dw_shape(shp):
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    return _ints(N, H, W, Cc, P, Q, sh, sw, ph, pw)


class _Conv2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, stride, padding, groups):
        x = _cl(x)
        shp = _conv_shape(x, w, stride, padding, groups)
        N, H, W, Cc, K, R, S, P, Q = shp[:9]
        y = torch.empty((N, K, P, Q), dtype=torch.float32, device=x.device, memory_format=CL)
        from . import cuda_gemm as cuda_tc
        ctx.dw = _is_depthwise(shp)
        ctx.gw, ctx.gb = _direct_grad(w), _direct_grad(bias)
        if ctx.dw:
            ctx.tc = False
            _call("fedb_dw_conv_fwd_k", x.data_ptr(), w.data_ptr(), _p(bias), y.data_ptr(), _dw_shape(shp), R, current_stream())
        elif cuda_tc.conv_forward(x, w, bias, y, shp):
            ctx.tc = True
        else:
            ctx.tc = False
            wt = torch.empty(w.numel(), dtype=torch.float32, device=x.device)
            _call("fedb_nn_weight_transpose", w.data_ptr(), wt.data_ptr(), 0, K, Cc // groups, R, S, current_stream())
            _call("fedb_nn_conv_fwd", x.data_ptr(), wt.data_ptr(), _p(bias), y.data_ptr(), _ints(*shp), current_stream())
        ctx.save_for_backward(x, w)
        ctx.shp, ctx.has_bias = shp, bias is not None
        from .cuda_gemm import bf16_of
        ctx.x16 = bf16_of(x)                 # shadow left by the producer of x: kept for the weight gradient
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        if ctx.x16 is not None:
            x._fedb_bf16 = ctx.x16
        shp = ctx.shp
        N, H, W, Cc, K, R, S, P, Q = shp[:9]
        groups = shp[13]
        dy = _cl(dy)
        dx = dw = db = None
        from . import cuda_gemm as cuda_tc
        direct = ctx.gw is not None and (not ctx.has_bias or ctx.gb is not None)
        if ctx.dw:
            if ctx.needs_input_grad[0]:
                dx = torch.empty_like(x, memory_format=CL)
                _call("fedb_dw_conv_dgrad_k", dy.data_ptr(), w.data_ptr(), dx.data_ptr(), _dw_shape(shp), R, current_stream())
            if ctx.needs_input_grad[1]:
                dw = ctx.gw if direct else torch.zeros_like(w)
                ds = _dw_shape(shp)
                _bind()
                scratch = torch.empty(int(lib().fedb_dw_conv_wgrad_scratch_k(ds, R)), dtype=torch.float32, device=x.device)
                _call("fedb_dw_conv_wgrad_k", x.data_ptr(), dy.data_ptr(), dw.data_ptr(), scratch.data_ptr(), ds, R,
                      current_stream(), launches=2)
                db = dy.sum(dim=(0, 2, 3)) if ctx.has_bias else None
                if direct:
                    if db is not None:
                        ctx.gb += db
                    dw = db = None
            return dx, dw, db, None, None, None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x, memory_format=CL)
            if not cuda_tc.conv_dgrad(dy, w, dx, shp):
                wt2 = torch.empty(w.numel(), dtype=torch.float32, device=x.device)
                _call("fedb_nn_weight_transpose", w.data_ptr(), 0, wt2.data_ptr(), K, Cc // groups, R, S, current_stream())
                _call("fedb_nn_conv_dgrad", dy.data_ptr(), wt2.data_ptr(), dx.data_ptr(), _ints(*shp), current_stream())
        if ctx.needs_input_grad[1]:
            dw = ctx.gw if direct else torch.zeros_like(w)
            db = (ctx.gb if direct else torch.zeros(K, dtype=torch.float32, device=x.device)) if ctx.has_bias else None
            if not cuda_tc.conv_wgrad(x, dy, dw, db, shp, direct=direct):
                _call("fedb_nn_conv_wgrad", x.data_ptr(), dy.data_ptr(), dw.data_ptr(), _p(db), _ints(*shp), current_stream())
            if direct:
                dw = db = None
        return dx, dw, db, None, None, None


def conv2d(x, w, bias, stride, padding, groups):
    return _Conv2d.apply(x, w, bias, tuple(stride), tuple(padding), groups)


# ------------------------------------------------------------------------------------ batch norm
class _BatchNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, run_mean, run_var, training, momentum, eps, act):
        x = _cl(x)
        Cc = x.shape[1]
        rows = x.numel() // Cc
        y = torch.empty_like(x, memory_format=CL)
        mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
        invstd = torch.empty(Cc, dtype=torch.float32, device=x.device)
        scratch = torch.empty(2 * Cc * _BN_CHUNKS, dtype=torch.float32, device=x.device) if training else None
        y16 = _shadow(y) if training else None
        _call("fedb_nn_bn_fwd16", x.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(), run_mean.data_ptr(),
              run_var.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rows, Cc, eps, momentum, int(training), act,
              _p(scratch), 2 * Cc * _BN_CHUNKS, _p(y16), current_stream(), launches=3 if training else 2)
        ctx.save_for_backward(x, gamma, beta, mean, invstd)
        ctx.act, ctx.rows, ctx.training = act, rows, bool(training)
        ctx.gg, ctx.gbeta = _direct_grad(gamma), _direct_grad(beta)
        if y16 is not None:
            y._fedb_bf16 = y16
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, beta, mean, invstd = ctx.saved_tensors
        dy = _cl(dy)
        Cc = x.shape[1]
        dx = torch.empty_like(x, memory_format=CL)
        # each BatchNorm is used once per step, so writing its two gradients into the (zeroed) arena slots IS accumulating
        direct = ctx.gg is not None and ctx.gbeta is not None
        dgamma = ctx.gg if direct else torch.empty(Cc, dtype=torch.float32, device=x.device)
        dbeta = ctx.gbeta if direct else torch.empty(Cc, dtype=torch.float32, device=x.device)
        scratch = torch.empty(2 * Cc * _BN_CHUNKS, dtype=torch.float32, device=x.device)
        dx16 = _shadow(dx)
        _call("fedb_nn_bn_bwd16", x.data_ptr(), dy.data_ptr(), dx.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
              gamma.data_ptr(), beta.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), ctx.rows, Cc, ctx.act,
              int(ctx.training), scratch.data_ptr(), 2 * Cc * _BN_CHUNKS, _p(dx16), current_stream(), launches=3)
        if dx16 is not None:
            dx._fedb_bf16 = dx16
        if direct:
            return dx, None, None, None, None, None, None, None, None
        return dx, dgamma, dbeta, None, None, None, None, None, None


def batch_norm(x, bn, act="none"):
    """BatchNorm2d with PyTorch's buffer semantics (running stats momentum 0.1 with the unbiased
    variance, ``num_batches_tracked += 1`` in train mode) and an optional fused activation."""
    y = _BatchNorm.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.training,
                         bn.momentum if bn.momentum is not None else 0.1, bn.eps, ACT[act])
    if bn.training:
        sc = _train_scope()
        if sc is None or not sc.fused_counters:      # inside the engine's step the end-of-step SGD kernel counts for all
            bn.num_batches_tracked += 1
    return y


# ------------------------------------------------------------------------------------ elementwise
class _Act(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, act):
        x = _cl(x)
        y = torch.empty_like(x)
        _call("fedb_nn_act_fwd", x.data_ptr(), y.data_ptr(), x.numel(), act, current_stream())
        ctx.save_for_backward(x)
        ctx.act = act
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dy = _match(dy, x)
        dx = torch.empty_like(x)
        _call("fedb_nn_act_bwd", x.data_ptr(), dy.data_ptr(), dx.data_ptr(), x.numel(), ctx.act, current_stream())
        return dx, None


def _match(t, like):
    """Same memory layout as ``like`` (both views of the same logical shape)."""
    if like.dim() == 4 and like.is_contiguous(memory_format=CL):
        return t.contiguous(memory_format=CL)
    return t.contiguous()


def relu(x):
    return _Act.apply(x, 1)


def swish(x):
    return _Act.apply(x, 2)


def sigmoid(x):
    return _Act.apply(x, 3)


class _AddAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, act):
        a = _cl(a)
        b = _match(b, a)
        y = torch.empty_like(a)
        y16 = _shadow(y)
        _call("fedb_nn_add_act16", a.data_ptr(), b.data_ptr(), y.data_ptr(), a.numel(), act, _p(y16), current_stream())
        if y16 is not None:
            y._fedb_bf16 = y16
        if act:
            ctx.save_for_backward(y)
        ctx.act = act
        return y

    @staticmethod
    def backward(ctx, dy):
        if ctx.act:
            (y,) = ctx.saved_tensors
            dy = _match(dy, y)
            d = torch.empty_like(y)
            # relu'(z) as a function of the output: y > 0
            _call("fedb_nn_act_bwd", y.data_ptr(), dy.data_ptr(), d.data_ptr(), y.numel(), 1, current_stream())
            return d, d, None
        return dy, dy, None


def add(a, b, act="none"):
    if act not in ("none", "relu"):
        raise ValueError("add supports act none|relu")
    return _AddAct.apply(a, b, ACT[act])


class _ChannelScale(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, s):
        x = _cl(x)
        N, Cc, H, W = x.shape
        s2 = s.reshape(N, Cc).contiguous()
        y = torch.empty_like(x, memory_format=CL)
        _call("fedb_nn_scale", x.data_ptr(), s2.data_ptr(), y.data_ptr(), x.numel(), H * W * Cc, Cc, 0, current_stream())
        ctx.save_for_backward(x, s2)
        ctx.s_shape = s.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        x, s2 = ctx.saved_tensors
        dy = _cl(dy)
        N, Cc, H, W = x.shape
        dx = torch.empty_like(x, memory_format=CL)
        _call("fedb_nn_scale", dy.data_ptr(), s2.data_ptr(), dx.data_ptr(), x.numel(), H * W * Cc, Cc, 0, current_stream())
        ds = torch.empty(N, Cc, dtype=torch.float32, device=x.device)
        _call("fedb_nn_scale_bwd_gate", x.data_ptr(), dy.data_ptr(), ds.data_ptr(), H * W, Cc, N, current_stream())
        return dx, ds.reshape(ctx.s_shape)


def channel_scale(x, s):
    return _ChannelScale.apply(x, s)


_mask_ctr = {}


def _mask(n, keep, device):
    """Bernoulli(keep)/keep mask; the draw depends on a device counter so graph replays differ."""
    key = torch.device(device).index or 0
    if key not in _mask_ctr:
        _mask_ctr[key] = torch.zeros(1, dtype=torch.int64, device=device)
    ctr = _mask_ctr[key]
    m = torch.empty(n, dtype=torch.float32, device=device)
    _call("fedb_nn_bernoulli_mask", m.data_ptr(), n, keep, 0x5EED, ctr.data_ptr(), current_stream())
    ctr.add_(1)
    return m


class _SampleScale(torch.autograd.Function):
    """x * m with m per sample (mode 1) or per element (plain product on a flat view)."""

    @staticmethod
    def forward(ctx, x, m, per_sample):
        x = x.contiguous(memory_format=CL) if x.dim() == 4 else x.contiguous()
        y = torch.empty_like(x)
        _call("fedb_nn_scale", x.data_ptr(), m.data_ptr(), y.data_ptr(), x.numel(), per_sample, 1, 1, current_stream())
        ctx.save_for_backward(m)
        ctx.per_sample = per_sample
        return y

    @staticmethod
    def backward(ctx, dy):
        (m,) = ctx.saved_tensors
        dy = dy.contiguous(memory_format=CL) if dy.dim() == 4 else dy.contiguous()
        dx = torch.empty_like(dy)
        _call("fedb_nn_scale", dy.data_ptr(), m.data_ptr(), dx.data_ptr(), dy.numel(), ctx.per_sample, 1, 1, current_stream())
        return dx, None, None


def drop_connect(x, keep):
    return _SampleScale.apply(x, _mask(x.shape[0], keep, x.device), x.numel() // x.shape[0])


def dropout(x, p):
    return _SampleScale.apply(x, _mask(x.numel(), 1.0 - p, x.device), 1)


# ------------------------------------------------------------------------------------ pooling
class _Pool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, is_max, k, s, p):
        x = _cl(x)
        N, Cc, H, W = x.shape
        P = (H + 2 * p[0] - k[0]) // s[0] + 1
        Q = (W + 2 * p[1] - k[1]) // s[1] + 1
        y = torch.empty((N, Cc, P, Q), dtype=torch.float32, device=x.device, memory_format=CL)
        arg = torch.empty(y.numel(), dtype=torch.int32, device=x.device) if is_max else None
        shp = (N, H, W, Cc, P, Q, k[0], k[1], s[0], s[1], p[0], p[1])
        _call("fedb_nn_pool_fwd", x.data_ptr(), y.data_ptr(), _p(arg), _ints(*shp), int(is_max), current_stream())
        ctx.shp, ctx.is_max, ctx.xshape = shp, is_max, x.shape
        if is_max:
            ctx.save_for_backward(arg)
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = _cl(dy)
        arg = ctx.saved_tensors[0] if ctx.is_max else None
        dx = torch.empty(ctx.xshape, dtype=torch.float32, device=dy.device, memory_format=CL)
        _call("fedb_nn_pool_bwd", dy.data_ptr(), _p(arg), dx.data_ptr(), _ints(*ctx.shp), int(ctx.is_max), current_stream())
        return dx, None, None, None, None


def pool2d(x, kind, k, s, p):
    return _Pool.apply(x, kind == "max", tuple(k), tuple(s), tuple(p))


# ------------------------------------------------------------------------------------ channel moves
def _chcopy(src, dst, src_off, dst_off, width, shuffle=1, accumulate=False):
    rows = src.numel() // src.shape[1]
    _call("fedb_nn_channel_copy", src.data_ptr(), dst.data_ptr(), rows, src.shape[1], dst.shape[1], src_off, dst_off,
          width, shuffle, int(accumulate), current_stream())


class _Cat(torch.autograd.Function):
    @staticmethod
    def forward(ctx, *xs):
        xs = [_cl(x) for x in xs]
        N, _, H, W = xs[0].shape
        widths = [x.shape[1] for x in xs]
        out = torch.empty((N, sum(widths), H, W), dtype=torch.float32, device=xs[0].device, memory_format=CL)
        off = 0
        for x, wd in zip(xs, widths):
            _chcopy(x, out, 0, off, wd)
            off += wd
        ctx.widths = widths
        return out

    @staticmethod
    def backward(ctx, dy):
        dy = _cl(dy)
        N, _, H, W = dy.shape
        outs, off = [], 0
        for wd in ctx.widths:
            d = torch.empty((N, wd, H, W), dtype=torch.float32, device=dy.device, memory_format=CL)
            _chcopy(dy, d, off, 0, wd)
            outs.append(d)
            off += wd
        return tuple(outs)


def cat_channels(xs: List[torch.Tensor]):
    return _Cat.apply(*xs)


class _Slice(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, lo, hi):
        x = _cl(x)
        N, Cc, H, W = x.shape
        out = torch.empty((N, hi - lo, H, W), dtype=torch.float32, device=x.device, memory_format=CL)
        _chcopy(x, out, lo, 0, hi - lo)
        ctx.lo, ctx.hi, ctx.C = lo, hi, Cc
        return out

    @staticmethod
    def backward(ctx, dy):
        dy = _cl(dy)
        N, _, H, W = dy.shape
        dx = torch.empty((N, ctx.C, H, W), dtype=torch.float32, device=dy.device, memory_format=CL)
        from .cuda import zero_
        zero_(dx)
        _chcopy(dy, dx, 0, ctx.lo, ctx.hi - ctx.lo)
        return dx, None, None


def channel_slice(x, lo, hi):
    return _Slice.apply(x, lo, hi)


class _Shuffle(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, groups):
        x = _cl(x)
        out = torch.empty_like(x, memory_format=CL)
        _chcopy(x, out, 0, 0, x.shape[1], shuffle=groups)
        ctx.groups = groups
        return out

    @staticmethod
    def backward(ctx, dy):
        dy = _cl(dy)
        dx = torch.empty_like(dy, memory_format=CL)
        # the inverse of a (g, C/g) transpose is the (C/g, g) transpose
        _chcopy(dy, dx, 0, 0, dy.shape[1], shuffle=dy.shape[1] // ctx.groups)
        return dx, None


def channel_shuffle(x, groups):
    return _Shuffle.apply(x, groups)


def dual_path_merge(x, y, d):
    a = add(channel_slice(x, 0, d), channel_slice(y, 0, d))
    return relu(cat_channels([a, channel_slice(x, d, x.shape[1]), channel_slice(y, d, y.shape[1])]))


# ------------------------------------------------------------------------------------ linear
class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, act):
        x = x.contiguous()
        B, Kd = x.shape
        J = w.shape[0]
        y = torch.empty((B, J), dtype=torch.float32, device=x.device)
        # classifier heads of the zoo are [batch, <=512] x [10, <=512]: far below a tensor-core tile, CUDA-core kernels
        _call("fedb_nn_linear_fwd", x.data_ptr(), w.data_ptr(), _p(bias), y.data_ptr(), B, Kd, J, act, current_stream())
        ctx.save_for_backward(x, w, y if act else None)
        ctx.act, ctx.has_bias = act, bias is not None
        ctx.gw, ctx.gb = _direct_grad(w), _direct_grad(bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        dy = dy.contiguous()
        if ctx.act:                      # relu only on linears in the zoo
            d = torch.empty_like(dy)
            _call("fedb_nn_act_bwd", y.data_ptr(), dy.data_ptr(), d.data_ptr(), dy.numel(), 1, current_stream())
            dy = d
        B, Kd = x.shape
        J = w.shape[0]
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        direct = ctx.gw is not None and (not ctx.has_bias or ctx.gb is not None) and ctx.needs_input_grad[1]
        dw = (ctx.gw if direct else torch.zeros_like(w)) if ctx.needs_input_grad[1] else None
        db = (ctx.gb if direct else torch.zeros(J, dtype=torch.float32, device=x.device)) if (ctx.has_bias and dw is not None) else None
        _call("fedb_nn_linear_bwd", dy.data_ptr(), x.data_ptr(), w.data_ptr(), _p(dx), _p(dw), _p(db), B, Kd, J,
              current_stream(), launches=2)
        if direct:
            dw = db = None
        return dx, dw, db, None


def linear(x, w, bias, act="none"):
    if act not in ("none", "relu"):
        raise ValueError("linear supports act none|relu")
    return _Linear.apply(x, w, bias, ACT[act])


def conv_relu_pool(x, w, bias):
    """Generic composition (the fused single-kernel version lives in csrc/lenet.cu and is what the
    LeNet engine runs; this path serves model.forward() on CUDA tensors outside the engine)."""
    y = relu(conv2d(x, w, bias, (1, 1), (0, 0), 1))
    return pool2d(y, "max", (2, 2), (2, 2), (0, 0))


def cross_entropy(logits, target):
    from .cuda import cross_entropy as ce
    return ce(logits, target)
