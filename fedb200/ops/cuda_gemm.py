"""Bindings of the tcgen05/TMEM/TMA GEMM (csrc/gemm_tc.cu) and the conv lowering onto it.

A 1x1 / stride-1 / ungrouped convolution over NHWC activations IS a GEMM over the flat
[pixels, channels] matrices:
  forward  Y[pix,K]  = X[pix,C] W[K,C]^T          TF32, operands exactly as stored
  dgrad    dX[pix,C] = dY[pix,K] Wt[C,K]^T        TF32, with a transposed copy of the (small) weight
  wgrad    dW[K,C]   = dY[pix,K]^T X[pix,C]       BF16 copies of dY and X, MN-major (as stored), split-K
Shapes outside the covered set return False and the caller runs the generic fp32 kernel.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check, current_stream

import os

_fns = {}
MIN_PIXELS = 128          # below one M tile the generic kernel is as good
WGRAD_TILES = int(os.environ.get("FEDB200_WGRAD_TILES", "148"))   # split-K weight gradients aim at this many tiles: the GEMM is
                          # persistent (one CTA per SM walks the tiles), so ONE wave is ideal - more splits only add
                          # red.global.add traffic to L2 (7 M atomics for a 128-channel 3x3 layer at 444 tiles: 50 -> ~30 us)
ENABLED = os.environ.get("FEDB200_TC", "1") != "0"      # 0: every convolution takes the generic fp32 kernels (debugging)


class PackJob(C.Structure):
    """One record of the device table walked by fedb_pack_multi (csrc/gemm_tc.cu: struct PackJob, 56 bytes)."""
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("total", C.c_int64), ("kind", C.c_int32), ("K", C.c_int32),
                ("C", C.c_int32), ("T", C.c_int32), ("Rp", C.c_int32), ("groups", C.c_int32), ("chunk0", C.c_int32),
                ("pad_", C.c_int32)]


PACK_CHUNK = 9216         # FEDB_PACK_CHUNK: elements per CTA of fedb_pack_multi


_SCOPE = None             # the StepScope of the training step / evaluation pass that is running (None: no caching)


def scope():
    return _SCOPE


class StepScope:
    """Everything about a model's convolutions that is the same in every step, kept across steps.

    Outside a scope every convolution call packs its filter for the tensor-core kernel, casts its operands to bf16 for
    the weight gradient, zero-fills a packed gradient and un-packs it again - about eight small launches per layer and
    step.  Inside a scope (the engine opens one around every training step and around every evaluation pass):
      * packed filters live in persistent buffers; ``begin()`` refreshes ALL of them with one launch
        (fedb_pack_multi) after the weights changed - a layer met for the first time is packed on the spot and joins
        the table;
      * weight gradients are accumulated by the split-K kernels into persistent packed buffers and
        ``end_backward()`` adds ALL of them into the gradient arena (and clears them) with one launch;
      * gradients go straight into the arena views (``param.grad``) instead of through autograd's accumulation;
      * BatchNorm / residual-add kernels leave a bf16 shadow of their output next to it (``tensor._fedb_bf16``), the
        operand format of the weight-gradient GEMM, so no cast pass re-reads the activation.
    The addresses in the job tables never change, so a captured CUDA graph replays them.  Weights must not change
    between ``begin()`` and the last use of the step (they do not: SGD runs after ``end_backward()``)."""

    def __init__(self, device, train: bool):
        self.device, self.train = torch.device(device), train
        self.pack, self.pack_jobs = {}, []
        self.unpack, self.unpack_jobs = {}, []
        self._tables = {}
        self._retired = []          # tables a captured graph may still read
        self._prev = None
        self.fused_counters = False  # True: the step's SGD kernel advances the BatchNorm counters (engine), not each layer

    def __enter__(self):
        global _SCOPE
        self._prev, _SCOPE = _SCOPE, self
        return self

    def __exit__(self, *exc):
        global _SCOPE
        _SCOPE = self._prev
        return False

    # ---- tables
    def _table(self, which, jobs):
        tab = self._tables.get(which)
        if tab is None or tab[1] != len(jobs):
            chunks = 0
            for j in jobs:                               # prefix sum: where each job's CTAs start in the flat grid
                j.chunk0 = chunks
                chunks += (j.total + PACK_CHUNK - 1) // PACK_CHUNK
            raw = bytes((PackJob * len(jobs))(*jobs))
            dev = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.device)
            if tab is not None:
                self._retired.append(tab[0])
            tab = (dev, len(jobs), chunks)
            self._tables[which] = tab
        return tab

    def sync_tables(self) -> None:
        """Uploads the job tables now (a plain H2D copy): call after a warm-up pass and BEFORE capturing a graph."""
        if self.pack_jobs:
            self._table("pack", self.pack_jobs)
        if self.unpack_jobs:
            self._table("unpack", self.unpack_jobs)

    def _run(self, which, jobs):
        if not jobs:
            return
        dev, n, chunks = self._table(which, jobs)
        check(_f("fedb_pack_multi")(dev.data_ptr(), n, chunks, current_stream()), "fedb_pack_multi")
        _lib.count_launch()

    def begin(self) -> None:
        """The weights may have changed since the last step / pass: refresh every packed filter (one launch)."""
        self._run("pack", self.pack_jobs)

    def end_backward(self) -> None:
        """Packed weight gradients -> += gradient arena, packed buffers cleared (one launch)."""
        self._run("unpack", self.unpack_jobs)

    # ---- registration
    def packed(self, w, kind: int, K: int, Cc: int, T: int, Rp: int, groups: int, n_out: int) -> torch.Tensor:
        key = (w.data_ptr(), kind, K, Cc, T, Rp, groups)
        t = self.pack.get(key)
        if t is None:
            if kind in (0, 1) and T == 9 and groups == 1 and K % 32 == 0 and Cc % 32 == 0:
                kind += 5                                # coalesced 32 x 32 x 9 tiles (no padding in either layout)
            t = torch.empty(n_out, dtype=torch.float32, device=w.device)
            job = PackJob(w.data_ptr(), t.data_ptr(), n_out, kind, K, Cc, T, Rp, groups, 0, 0)
            self.pack[key] = t
            self.pack_jobs.append(job)
            first = PackJob(w.data_ptr(), t.data_ptr(), n_out, kind, K, Cc, T, Rp, groups, 0, 0)   # a table of one, now
            one = torch.frombuffer(bytearray(bytes(first)), dtype=torch.uint8).to(self.device)
            check(_f("fedb_pack_multi")(one.data_ptr(), 1, (n_out + PACK_CHUNK - 1) // PACK_CHUNK, current_stream()),
                  "fedb_pack_multi")
            _lib.count_launch()
            self._retired.append(one)
        return t

    def wgrad_buffer(self, grad: torch.Tensor, K: int, Cc: int, T: int) -> torch.Tensor:
        """Persistent zeroed [K][T][Cc] accumulator whose content ``end_backward()`` adds into ``grad`` (OIHW)."""
        key = (grad.data_ptr(), K, Cc, T)
        t = self.unpack.get(key)
        if t is None:
            t = torch.zeros(K * T * Cc, dtype=torch.float32, device=grad.device)
            self.unpack[key] = t
            kind = 7 if (T == 9 and K % 32 == 0 and Cc % 32 == 0) else 2
            self.unpack_jobs.append(PackJob(t.data_ptr(), grad.data_ptr(), K * T * Cc, kind, K, Cc, T, 0, 1, 0, 0))
        return t


def _wgrad_tiles(cin: int) -> int:
    """Tile target of a split-K weight gradient: one wave.  Inputs narrower than 128 channels run 64-column tiles, two CTAs per
    SM (csrc/gemm_tc.cu, FEDB_GEMM_CTAS_SMALL), so their wave is twice as many tiles (measured: the 64-channel layers of
    ResNet-18 need 65 us at 144 tiles and 52 us at 296)."""
    return WGRAD_TILES * (2 if cin < 128 else 1)


def _zero(t: torch.Tensor) -> None:
    from . import cuda as K
    K.zero_(t)


def bf16_of(t: torch.Tensor):
    """The bf16 shadow a producer kernel left next to ``t`` (None when there is none)."""
    s16 = getattr(t, "_fedb_bf16", None)
    if s16 is not None and s16.numel() == t.numel():
        return s16
    return None


def _f(name):
    if name not in _fns:
        vp, i, ll = C.c_void_p, C.c_int, C.c_longlong
        sig = {
            "fedb_gemm_tf32": [vp, vp, vp, vp, i, i, i, ll, ll, ll, i, i, i, vp],
            "fedb_gemm_bf16_tn": [vp, vp, vp, i, i, i, ll, ll, ll, i, i, vp],
            "fedb_cast_bf16": [vp, vp, ll, vp],
            "fedb_cast_bf16_pad": [vp, vp, ll, i, i, vp],
            "fedb_cast_bf16_pad_p": [vp, vp, ll, i, i, ll, vp],
            "fedb_transpose": [vp, vp, i, i, vp],
            "fedb_conv3x3_tf32": [vp, vp, vp, vp, i, i, i, i, i, i, vp],
            "fedb_pack3x3": [vp, vp, i, i, i, vp],
            "fedb_conv3x3_wgrad_bf16": [vp, vp, vp, i, i, i, i, i, i, vp],
            "fedb_conv_tf32": [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, vp],
            "fedb_conv_wgrad_bf16": [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp],
            "fedb_pack_filter": [vp, vp, i, i, i, i, vp],
            "fedb_pack_multi": [vp, i, i, vp],
            "fedb_pack_dgrad_s2": [vp, vp, i, i, vp],
            "fedb_im2col_small": [vp, vp, vp, i, i, i, i, i, i, i, vp],
            "fedb_conv_tf32_g": [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, ll, ll, vp],
            "fedb_conv_tf32_grouped": [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, i, vp],
            "fedb_conv_dgrad_s2_tf32_grouped": [vp, vp, vp, i, i, i, i, i, i, i, vp],
            "fedb_conv_wgrad_bf16_grouped": [vp, vp, vp, i, i, i, i, i, i, i, i, i, i, vp],
            "fedb_pack_filter_grouped": [vp, vp, i, i, i, i, i, vp],
            "fedb_pack_dgrad_s2_grouped": [vp, vp, i, i, i, vp],
            "fedb_conv_wgrad_bf16_g": [vp, vp, vp, i, i, i, i, i, i, i, i, i, ll, ll, vp],
            "fedb_conv_dgrad_s2_tf32_g": [vp, vp, vp, i, i, i, i, i, i, ll, ll, vp],
            "fedb_nn_channel_copy": [vp, vp, ll, i, i, i, i, i, i, i, vp],
            "fedb_conv_dgrad_s2_tf32": [vp, vp, vp, i, i, i, i, i, i, vp],
        }[name]
        _fns[name] = _lib.bind(name, C.c_int, sig)
    return _fns[name]


def gemm_tf32(A, B, D, M, N, K, lda, ldb, ldd, bias=None, relu=False, splits=1, accumulate=False):
    """D[M,N] (+)= A[M,K] B[N,K]^T on the tensor cores (TF32 inputs, fp32 accumulation in TMEM)."""
    check(_f("fedb_gemm_tf32")(A.data_ptr(), B.data_ptr(), D.data_ptr(), 0 if bias is None else bias.data_ptr(), M, N, K,
                               lda, ldb, ldd, int(relu), splits, int(accumulate), current_stream()), "fedb_gemm_tf32")
    _lib.count_launch()


def gemm_bf16_tn(At, Bt, D, M, N, K, lda, ldb, ldd, splits=1, accumulate=False):
    """D[M,N] (+)= At[K,M]^T Bt[K,N] with bf16 operands (reduction dimension outermost), fp32 out."""
    check(_f("fedb_gemm_bf16_tn")(At.data_ptr(), Bt.data_ptr(), D.data_ptr(), M, N, K, lda, ldb, ldd, splits,
                                  int(accumulate), current_stream()), "fedb_gemm_bf16_tn")
    _lib.count_launch()


def cast_bf16(x: torch.Tensor) -> torch.Tensor:
    out = torch.empty(x.numel(), dtype=torch.bfloat16, device=x.device)
    check(_f("fedb_cast_bf16")(x.data_ptr(), out.data_ptr(), x.numel(), current_stream()), "fedb_cast_bf16")
    _lib.count_launch()
    return out


def cast_bf16_rows(x: torch.Tensor, rows: int, C: int):
    """bf16 copy of an fp32 [rows][C] matrix (an NHWC activation flattened over pixels) and its row pitch in elements:
    C itself when that already is a 16-byte pitch, else C rounded up to 8 with zero columns.  A shadow left by the
    kernel that produced ``x`` is used as it is."""
    if C % 8 == 0:
        s16 = bf16_of(x)
        return (s16 if s16 is not None else cast_bf16(x)), C
    Cp = (C + 7) // 8 * 8
    out = torch.empty(rows * Cp, dtype=torch.bfloat16, device=x.device)
    check(_f("fedb_cast_bf16_pad")(x.data_ptr(), out.data_ptr(), rows, C, Cp, current_stream()), "fedb_cast_bf16_pad")
    _lib.count_launch()
    return out, Cp


def transpose(w2d: torch.Tensor, rows: int, cols: int) -> torch.Tensor:
    if _SCOPE is not None:
        return _SCOPE.packed(w2d, 4, rows, cols, 1, 0, 1, rows * cols)
    out = torch.empty(rows * cols, dtype=torch.float32, device=w2d.device)
    check(_f("fedb_transpose")(w2d.data_ptr(), out.data_ptr(), rows, cols, current_stream()), "fedb_transpose")
    _lib.count_launch()
    return out


def matmul_nt(a: torch.Tensor, b: torch.Tensor, bias=None, relu=False) -> torch.Tensor:
    """a[M,K] @ b[N,K]^T (both row-major) -> [M,N]."""
    M, K = a.shape
    N = b.shape[0]
    d = torch.empty(M, N, dtype=torch.float32, device=a.device)
    gemm_tf32(a, b, d, M, N, K, a.stride(0), b.stride(0), N, bias=bias, relu=relu)
    return d


def _is_pointwise(shp, cin_mult: int = 4, cout_mult: int = 1) -> bool:
    """1x1 / stride 1 / ungrouped.  Channel counts only need the 16-byte row pitch TMA asks for (fp32: % 4, bf16: % 8):
    tails of the 32/64-wide K tile and of the N tile are zero-filled by the TMA unit and masked by the epilogue."""
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    return (R == 1 and S == 1 and sh == 1 and sw == 1 and ph == 0 and pw == 0 and groups == 1
            and Cc % cin_mult == 0 and K % cout_mult == 0 and N * H * W >= MIN_PIXELS)


def _pow2(v: int) -> bool:
    return v >= 1 and (v & (v - 1)) == 0


def _is_3x3(shp, cin, cout) -> bool:
    """3x3 / stride 1 / pad 1 / ungrouped, power-of-two maps; `cin` (the GEMM's reduction channels) only needs the 16-byte
    pixel pitch: every tap is padded to whole 32-channel K tiles (zeros from TMA on the activation side, explicit zeros
    in the packed filter)."""
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    return (R == 3 and S == 3 and sh == 1 and sw == 1 and ph == 1 and pw == 1 and groups == 1 and cin % 4 == 0
            and cout % 4 == 0 and _pow2(H) and _pow2(W) and 2 <= W <= 32 and N * H * W >= MIN_PIXELS
            and (128 // W <= H or (128 // (W * H)) >= 1))


def _strided(shp):
    """(fs, stride, pad) when the layer is a stride-2 3x3/pad-1 or 1x1/pad-0 ungrouped convolution whose OUTPUT maps
    are powers of two (the down-sampling layers of the ResNet / DLA / RegNet families), else None."""
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    if groups != 1 or R != S or sh != 2 or sw != 2 or H % 2 or W % 2:
        return None
    if not ((R == 3 and ph == 1 and pw == 1) or (R == 1 and ph == 0 and pw == 0)):
        return None
    if not (_pow2(P) and _pow2(Q) and 2 <= Q <= 32 and P * 2 == H and Q * 2 == W and N * P * Q >= MIN_PIXELS):
        return None
    return R, 2, ph


def _pad32(v: int) -> int:
    return (v + 31) // 32 * 32


def pack_filter(w, mode: int, taps: int) -> torch.Tensor:
    """mode 0: OIHW -> [K][tap][C padded to 32] (forward); mode 1: -> [C][flipped tap][K padded to 32] (data gradient)."""
    K, Cc = w.shape[0], w.shape[1]
    n = K * taps * _pad32(Cc) if mode == 0 else Cc * 9 * _pad32(K)
    if _SCOPE is not None and (mode == 0 or taps == 9):
        return _SCOPE.packed(w, mode, K, Cc, taps, _pad32(Cc) if mode == 0 else _pad32(K), 1, n)
    out = torch.empty(n, dtype=torch.float32, device=w.device)
    check(_f("fedb_pack_filter")(w.data_ptr(), out.data_ptr(), K, Cc, mode, taps, current_stream()), "fedb_pack_filter")
    _lib.count_launch()
    return out


def pack3x3(w, mode: int) -> torch.Tensor:
    return pack_filter(w, mode, 9)


def conv3x3_tf32(x, wk, y, bias, N, H, W, cin, cout, relu=False):
    check(_f("fedb_conv3x3_tf32")(x.data_ptr(), wk.data_ptr(), y.data_ptr(), 0 if bias is None else bias.data_ptr(),
                                  N, H, W, cin, cout, int(relu), current_stream()), "fedb_conv3x3_tf32")
    _lib.count_launch()


def _small_reduction(shp) -> bool:
    """Cin * R * S <= 32 and ungrouped (the 3-channel stem): explicit im2col to a [pixels, 32] matrix, then a GEMM.
    A 1x1 convolution whose input already has a 16-byte pixel pitch is a GEMM as it stands and is not routed here."""
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    if R == 1 and Cc % 4 == 0:
        return False
    return groups == 1 and R == S and sh == sw and ph == pw and Cc * R * S <= 32 and N * P * Q >= MIN_PIXELS


def _channel_copy(src, dst, rows, cin, cout, width, accumulate=False):
    check(_f("fedb_nn_channel_copy")(src.data_ptr(), dst.data_ptr(), rows, cin, cout, 0, 0, width, 1, int(accumulate),
                                     current_stream()), "fedb_nn_channel_copy")
    _lib.count_launch()


def _im2col(x, shp, f32: bool, bf16: bool):
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    pix = N * P * Q
    o32 = torch.empty(pix * 32, dtype=torch.float32, device=x.device) if f32 else None
    o16 = torch.empty(pix * 64, dtype=torch.bfloat16, device=x.device) if bf16 else None
    check(_f("fedb_im2col_small")(x.data_ptr(), 0 if o32 is None else o32.data_ptr(), 0 if o16 is None else o16.data_ptr(),
                                  N, H, W, Cc, R, sh, ph, current_stream()), "fedb_im2col_small")
    _lib.count_launch()
    return o32, o16


def _group_geometry(shp, direction: str = "any"):
    """(Cg, Kg) when a grouped 3x3 / 1x1 convolution (stride 1 or 2) can run on the implicit-GEMM kernels: ALL groups in
    one launch, the group index shifts the channel coordinates of the TMA boxes (ResNeXt, RegNet, DPN, ShuffleNet).
    Only the REDUCTION side needs 16-byte aligned group offsets (it is the TMA operand): the input channels per group for
    the forward pass, the output channels per group for the data gradient; the other side is written by the epilogue,
    which falls back to scalar stores.  ``direction``: "fwd", "dgrad" or "any" (either one qualifies)."""
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    if groups < 2 or Cc % groups or K % groups:
        return None
    Cg, Kg = Cc // groups, K // groups
    ok = {"fwd": Cg % 4 == 0, "dgrad": Kg % 4 == 0, "wgrad": True}
    ok["any"] = ok["fwd"] or ok["dgrad"]
    if not ok[direction] or R != S or R not in (1, 3) or sh != sw or sh not in (1, 2) or ph != (R - 1) // 2 or pw != ph:
        return None
    if not (_pow2(P) and _pow2(Q) and 2 <= Q <= 32 and P * sh == H and Q * sw == W and N * P * Q >= MIN_PIXELS):
        return None
    return Cg, Kg


def _pack_grouped(w, mode: int, taps: int, groups: int) -> torch.Tensor:
    K, Cg = w.shape[0], w.shape[1]
    Kg = K // groups
    n = K * taps * _pad32(Cg) if mode == 0 else groups * Cg * taps * (Kg if taps == 1 else _pad32(Kg))
    if _SCOPE is not None and mode == 1:
        return _SCOPE.packed(w, 1, K, Cg, taps, Kg if taps == 1 else _pad32(Kg), groups, n)
    out = torch.empty(n, dtype=torch.float32, device=w.device)
    check(_f("fedb_pack_filter_grouped")(w.data_ptr(), out.data_ptr(), K, Cg, groups, mode, taps, current_stream()),
          "fedb_pack_filter_grouped")
    _lib.count_launch()
    return out


def _grouped_forward(x, w, bias, y, shp, Cg, Kg) -> bool:
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    wk = pack_filter(w, 0, 9) if R == 3 else w          # rows only know their own group's inputs: the plain packing
    check(_f("fedb_conv_tf32_grouped")(x.data_ptr(), wk.data_ptr(), y.data_ptr(), 0 if bias is None else bias.data_ptr(),
                                       N, H, W, Cc, K, groups, R, sh, ph, 0, current_stream()), "fedb_conv_tf32_grouped")
    _lib.count_launch()
    return True


def _grouped_dgrad(dy, w, dx, shp, Cg, Kg) -> bool:
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    if sh == 1:
        wk = _pack_grouped(w, 1, R * R, groups)         # [C][flipped tap][Kg padded] (1x1: the per-group transposes)
        check(_f("fedb_conv_tf32_grouped")(dy.data_ptr(), wk.data_ptr(), dx.data_ptr(), 0, N, H, W, K, Cc, groups, R, 1, ph, 0,
                                           current_stream()), "fedb_conv_tf32_grouped")
        _lib.count_launch()
        return True
    if R == 3 and _SCOPE is not None:
        wk = _SCOPE.packed(w, 3, Kg, Cg, 9, _pad32(Kg), groups, 9 * Cc * _pad32(Kg))
    elif R == 3:
        wk = torch.empty(9 * Cc * _pad32(Kg), dtype=torch.float32, device=w.device)
        check(_f("fedb_pack_dgrad_s2_grouped")(w.data_ptr(), wk.data_ptr(), K, Cg, groups, current_stream()),
              "fedb_pack_dgrad_s2_grouped")
        _lib.count_launch()
    else:
        wk = _pack_grouped(w, 1, 1, groups)
        _zero(dx)
    check(_f("fedb_conv_dgrad_s2_tf32_grouped")(dy.data_ptr(), wk.data_ptr(), dx.data_ptr(), N, P, Q, K, Cc, groups, R,
                                                current_stream()), "fedb_conv_dgrad_s2_tf32_grouped")
    _lib.count_launch(4 if R == 3 else 1)
    return True


MAX_SPLIT_GROUPS = 8       # odd per-group channel counts: weight gradient group by group, up to this many groups


def _grouped_wgrad_by_group(x, dy, dw, db, shp, Cg, Kg) -> bool:
    """Per-group channel counts that give no 16-byte bf16 offset inside the whole tensor (ShuffleNetG: 25, 100): every
    group's channels are extracted into dense zero-padded bf16 matrices and run through the ungrouped kernel."""
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    taps, pix, pix_in = R * R, N * P * Q, N * H * W
    kp, cp = (Kg + 7) // 8 * 8, (Cg + 7) // 8 * 8
    tiles = ((Kg + 127) // 128) * ((Cg + 127) // 128) * taps
    splits = max(1, min(pix // 256, WGRAD_TILES // max(tiles, 1)))
    cast = _f("fedb_cast_bf16_pad_p")
    for g in range(groups):
        dy16 = torch.empty(pix * kp, dtype=torch.bfloat16, device=x.device)
        x16 = torch.empty(pix_in * cp, dtype=torch.bfloat16, device=x.device)
        check(cast(dy.data_ptr() + 4 * g * Kg, dy16.data_ptr(), pix, Kg, kp, K, current_stream()), "fedb_cast_bf16_pad_p")
        check(cast(x.data_ptr() + 4 * g * Cg, x16.data_ptr(), pix_in, Cg, cp, Cc, current_stream()), "fedb_cast_bf16_pad_p")
        _lib.count_launch(2)
        dwg = dw[g * Kg:(g + 1) * Kg]
        dwk = torch.zeros(Kg * taps * Cg, dtype=torch.float32, device=x.device) if taps > 1 else dwg
        check(_f("fedb_conv_wgrad_bf16_g")(dy16.data_ptr(), x16.data_ptr(), dwk.data_ptr(), N, H, W, Cg, Kg, R, sh, ph, splits,
                                           kp, cp, current_stream()), "fedb_conv_wgrad_bf16_g")
        _lib.count_launch()
        if taps > 1:
            check(_f("fedb_pack_filter")(dwk.data_ptr(), dwg.data_ptr(), Kg, Cg, 2, taps, current_stream()), "fedb_pack_filter")
            _lib.count_launch()
    if db is not None:
        db += dy.sum(dim=(0, 2, 3))
    return True


def _grouped_wgrad(x, dy, dw, db, shp, Cg, Kg, direct: bool = False) -> bool:
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    if Cg % 8 or Kg % 8:                                 # no 16-byte bf16 group offsets inside the whole tensors
        return _grouped_wgrad_by_group(x, dy, dw, db, shp, Cg, Kg) if groups <= MAX_SPLIT_GROUPS else False
    taps, pix = R * R, N * P * Q
    tiles = ((Kg + 127) // 128) * ((Cg + 127) // 128) * taps * groups
    splits = max(1, min(pix // 256, WGRAD_TILES // max(tiles, 1)))
    dy16, x16 = bf16_of(dy), bf16_of(x)                  # shadows left by the producers (BatchNorm / residual kernels)
    dy16 = cast_bf16(dy) if dy16 is None else dy16
    x16 = cast_bf16(x) if x16 is None else x16
    kept = _SCOPE is not None and _SCOPE.train and direct and taps > 1
    dwk = dw if taps == 1 else (_SCOPE.wgrad_buffer(dw, K, Cg, taps) if kept else
                                torch.zeros(K * taps * Cg, dtype=torch.float32, device=x.device))
    check(_f("fedb_conv_wgrad_bf16_grouped")(dy16.data_ptr(), x16.data_ptr(), dwk.data_ptr(), N, H, W, Cc, K, groups, R, sh, ph,
                                             splits, current_stream()), "fedb_conv_wgrad_bf16_grouped")
    _lib.count_launch()
    if taps > 1 and not kept:
        check(_f("fedb_pack_filter")(dwk.data_ptr(), dw.data_ptr(), K, Cg, 2, taps, current_stream()), "fedb_pack_filter")
        _lib.count_launch()
    if db is not None:
        db += dy.sum(dim=(0, 2, 3))
    return True


def conv_forward(x, w, bias, y, shp) -> bool:
    if not ENABLED:
        return False
    gg = _group_geometry(shp, "fwd")
    if gg is not None:
        return _grouped_forward(x, w, bias, y, shp, *gg)
    if shp[13] > 1:
        return False
    if _small_reduction(shp):
        N, H, W, Cc, K, R, S, P, Q = shp[:9]
        xcol, _ = _im2col(x, shp, True, False)
        wp = torch.zeros(K * 32, dtype=torch.float32, device=x.device)
        _channel_copy(w, wp, K, Cc * R * S, 32, Cc * R * S)              # [K][C*R*S] -> [K][32], zero padded
        gemm_tf32(xcol, wp, y, N * P * Q, K, 32, 32, 32, K, bias=bias)
        return True
    if _is_3x3(shp, shp[3], shp[4]):
        N, H, W, Cc, K = shp[:5]
        conv3x3_tf32(x, pack3x3(w, 0), y, bias, N, H, W, Cc, K)      # implicit GEMM, shifted TMA boxes per tap
        return True
    st = _strided(shp)
    if st is not None and shp[3] % 4 == 0 and shp[4] % 4 == 0:
        N, H, W, Cc, K = shp[:5]
        fs, stride, pad = st
        wk = pack_filter(w, 0, 9) if fs == 3 else w                  # a 1x1 filter already is [Cout][Cin]
        check(_f("fedb_conv_tf32")(x.data_ptr(), wk.data_ptr(), y.data_ptr(), 0 if bias is None else bias.data_ptr(),
                                   N, H, W, Cc, K, fs, stride, pad, 0, current_stream()), "fedb_conv_tf32")
        _lib.count_launch()
        return True
    if not _is_pointwise(shp, 4, 1):
        return False
    N, H, W, Cc, K = shp[:5]
    gemm_tf32(x, w, y, N * H * W, K, Cc, Cc, Cc, K, bias=bias)          # Y[pix,K] = X[pix,C] W[K,C]^T
    return True


def conv_dgrad(dy, w, dx, shp) -> bool:
    if not ENABLED:
        return False
    gg = _group_geometry(shp, "dgrad")
    if gg is not None:
        return _grouped_dgrad(dy, w, dx, shp, *gg)
    if shp[13] > 1:
        return False
    if _is_3x3(shp, shp[4], shp[3]):
        N, H, W, Cc, K = shp[:5]
        conv3x3_tf32(dy, pack3x3(w, 1), dx, None, N, H, W, K, Cc)    # dX = conv3x3(dY, flipped + transposed filter)
        return True
    st = _strided(shp)
    if st is not None and shp[4] % 4 == 0 and shp[3] % 4 == 0:
        # stride 2: four parity classes of dX, each a dense stride-1 implicit GEMM over dY, scattered by the epilogue
        N, H, W, Cc, K, _, _, P, Q = shp[:9]
        fs = st[0]
        if fs == 3 and _SCOPE is not None:
            wc = _SCOPE.packed(w, 3, K, Cc, 9, _pad32(K), 1, 9 * _pad32(K) * Cc)
        elif fs == 3:
            wc = torch.empty(9 * _pad32(K) * Cc, dtype=torch.float32, device=w.device)
            check(_f("fedb_pack_dgrad_s2")(w.data_ptr(), wc.data_ptr(), K, Cc, current_stream()), "fedb_pack_dgrad_s2")
            _lib.count_launch()
        else:
            wc = transpose(w, K, Cc)
            _zero(dx)                                    # only the even/even pixels of dX receive gradient
        check(_f("fedb_conv_dgrad_s2_tf32")(dy.data_ptr(), wc.data_ptr(), dx.data_ptr(), N, P, Q, K, Cc, fs,
                                            current_stream()), "fedb_conv_dgrad_s2_tf32")
        _lib.count_launch(4 if fs == 3 else 1)
        return True
    if not _is_pointwise(shp, 1, 4):
        return False
    N, H, W, Cc, K = shp[:5]
    wt = transpose(w, K, Cc)                                             # Wt[C,K]
    gemm_tf32(dy, wt, dx, N * H * W, Cc, K, K, K, Cc)                    # dX[pix,C] = dY[pix,K] Wt[C,K]^T
    return True


def conv_wgrad(x, dy, dw, db, shp, direct: bool = False) -> bool:
    """``direct``: ``dw`` is the parameter's slot in the gradient arena itself (stable address), so inside a training
    scope the packed accumulator can be kept and un-packed by the step's single ``end_backward()`` launch."""
    if not ENABLED:
        return False
    gg = _group_geometry(shp, "wgrad")
    if gg is not None:
        return _grouped_wgrad(x, dy, dw, db, shp, *gg, direct=direct)
    if shp[13] > 1:
        return False
    if _small_reduction(shp):
        N, H, W, Cc, K, R, S, P, Q = shp[:9]
        pix = N * P * Q
        _, xcol16 = _im2col(x, shp, False, True)
        dy16, kp = cast_bf16_rows(dy, pix, K)
        dwp = torch.zeros(K * 64, dtype=torch.float32, device=x.device)
        splits = max(1, min(pix // 256, WGRAD_TILES))
        gemm_bf16_tn(dy16, xcol16, dwp, K, 64, pix, kp, 64, 64, splits=splits, accumulate=True)  # dWp[K][64] = dY^T Xcol
        _channel_copy(dwp, dw, K, 64, Cc * R * S, Cc * R * S, accumulate=True)                     # += into OIHW
        if db is not None:
            db += dy.sum(dim=(0, 2, 3))
        return True
    if _is_3x3(shp, shp[3], shp[4]):
        N, H, W, Cc, K = shp[:5]
        pix = N * H * W
        tiles = ((K + 127) // 128) * ((Cc + 127) // 128) * 9
        splits = max(1, min(pix // 256, _wgrad_tiles(Cc) // max(tiles, 1)))
        kept = _SCOPE is not None and _SCOPE.train and direct
        dwk = _SCOPE.wgrad_buffer(dw, K, Cc, 9) if kept else torch.zeros(K * 9 * Cc, dtype=torch.float32, device=x.device)
        # keep both alive: two temporaries would share one allocation; channel counts that are not multiples of 8 get
        # zero columns so that the bf16 rows keep a 16-byte pitch
        (dy16, kp), (x16, cp) = cast_bf16_rows(dy, pix, K), cast_bf16_rows(x, pix, Cc)
        check(_f("fedb_conv_wgrad_bf16_g")(dy16.data_ptr(), x16.data_ptr(), dwk.data_ptr(), N, H, W, Cc, K, 3, 1, 1,
                                           splits, kp, cp, current_stream()), "fedb_conv_wgrad_bf16_g")
        _lib.count_launch()
        if not kept:
            check(_f("fedb_pack3x3")(dwk.data_ptr(), dw.data_ptr(), K, Cc, 2, current_stream()), "fedb_pack3x3")   # += into OIHW
            _lib.count_launch()
        if db is not None:
            db += dy.sum(dim=(0, 2, 3))
        return True
    st = _strided(shp)
    if st is not None:
        N, H, W, Cc, K = shp[:5]
        fs, stride, pad = st
        taps, pix = fs * fs, N * shp[7] * shp[8]
        tiles = ((K + 127) // 128) * ((Cc + 127) // 128) * taps
        splits = max(1, min(pix // 256, _wgrad_tiles(Cc) // max(tiles, 1)))
        (dy16, kp), (x16, cp) = cast_bf16_rows(dy, pix, K), cast_bf16_rows(x, N * H * W, Cc)
        kept = _SCOPE is not None and _SCOPE.train and direct and taps > 1
        dwk = dw if taps == 1 else (_SCOPE.wgrad_buffer(dw, K, Cc, taps) if kept else
                                    torch.zeros(K * taps * Cc, dtype=torch.float32, device=x.device))
        check(_f("fedb_conv_wgrad_bf16_g")(dy16.data_ptr(), x16.data_ptr(), dwk.data_ptr(), N, H, W, Cc, K, fs, stride, pad,
                                           splits, kp, cp, current_stream()), "fedb_conv_wgrad_bf16_g")
        _lib.count_launch()
        if taps > 1 and not kept:
            check(_f("fedb_pack_filter")(dwk.data_ptr(), dw.data_ptr(), K, Cc, 2, taps, current_stream()), "fedb_pack_filter")
            _lib.count_launch()
        if db is not None:
            db += dy.sum(dim=(0, 2, 3))
        return True
    if not _is_pointwise(shp, 1, 1):
        return False
    N, H, W, Cc, K = shp[:5]
    pix = N * H * W
    # dW[K,C] = dY[pix,K]^T X[pix,C]: reduction over pixels, split across CTAs so that small layers
    # (few output tiles) still fill the machine; partial tiles are reduced with red.global.add
    tiles = ((K + 127) // 128) * ((Cc + 127) // 128)
    # 1x1 layers: half a wave of tiles (measured on MobileNet: 74 -> 2.13 ms, 148 -> 2.21 ms per step; their K loops are short, the
    # atomics of the partial tiles are what costs)
    splits = max(1, min(pix // 256, (WGRAD_TILES // 2) // max(tiles, 1)))
    (dy16, kp), (x16, cp) = cast_bf16_rows(dy, pix, K), cast_bf16_rows(x, pix, Cc)
    gemm_bf16_tn(dy16, x16, dw, K, Cc, pix, kp, cp, Cc, splits=splits, accumulate=True)
    if db is not None:
        db += dy.sum(dim=(0, 2, 3))       # bias on a pointwise conv that reaches a GEMM tile: rare in the zoo
    return True
