"""tcgen05 GEMM (csrc/gemm_tc.cu) against fp32 PyTorch.  Inputs are rounded to TF32 by the tensor
cores (10-bit mantissa), so tolerances are TF32-level: ~1e-3 relative to the row/column norms."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def G():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fedb200.ops import cuda_gemm
    return cuda_gemm


def _tol(K):
    return dict(rtol=2e-3, atol=2e-3 * (K ** 0.5))


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (256, 64, 96), (1000, 32, 64), (4096, 512, 512), (300, 192, 160),
                                   (2048, 1024, 512), (300, 24, 40), (512, 144, 24), (130, 6, 16)])   # K / N tails
def test_gemm_nt(G, M, N, K):
    torch.manual_seed(0)
    a, b = torch.randn(M, K, device=DEV), torch.randn(N, K, device=DEV)
    bias = torch.randn(N, device=DEV)
    got = G.matmul_nt(a, b, bias=bias, relu=True)
    want = torch.relu(a.double() @ b.double().t() + bias.double()).float()
    torch.testing.assert_close(got, want, **_tol(K))


@pytest.mark.parametrize("M,N,K,splits", [(128, 128, 256, 1), (256, 64, 4096, 8), (512, 512, 2048, 4), (64, 64, 8192, 16),
                                          (1024, 512, 512, 2), (24, 144, 1000, 3), (96, 16, 520, 1)])     # M / N / K tails
def test_gemm_bf16_tn_split_k(G, M, N, K, splits):
    """wgrad form: D[M,N] = A^T B with A stored [K][M], B stored [K][N] (bf16), reduction split over CTAs."""
    torch.manual_seed(2)
    a, b = torch.randn(K, M, device=DEV), torch.randn(K, N, device=DEV)
    a16, b16 = G.cast_bf16(a), G.cast_bf16(b)
    assert torch.equal(a16.view(K, M), a.to(torch.bfloat16))
    d = torch.zeros(M, N, device=DEV)
    G.gemm_bf16_tn(a16, b16, d, M, N, K, M, N, N, splits=splits, accumulate=True)
    want = (a.to(torch.bfloat16).double().t() @ b.to(torch.bfloat16).double()).float()
    torch.testing.assert_close(d, want, rtol=1e-3, atol=1e-3 * (K ** 0.5))


def test_transpose(G):
    w = torch.randn(96, 160, device=DEV)
    assert torch.equal(G.transpose(w, 96, 160).view(160, 96), w.t().contiguous())


def test_pointwise_conv_uses_tensor_cores_and_matches(G):
    import torch.nn.functional as TF
    from fedb200.ops import _lib, functional as F
    torch.manual_seed(3)
    x = torch.randn(8, 64, 8, 8)
    w = torch.randn(128, 64, 1, 1) * 0.1
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    xb, wb = x.clone().to(DEV).requires_grad_(True), w.clone().to(DEV).requires_grad_(True)
    ya = TF.conv2d(xa, wa)
    before = _lib.launches
    yb = F.conv2d(xb, wb)
    assert _lib.launches - before == 1            # one GEMM launch, no weight transpose / generic kernel
    torch.testing.assert_close(yb.cpu(), ya.detach(), rtol=2e-3, atol=2e-2)
    g = torch.randn_like(ya)
    ya.backward(g), yb.backward(g.to(DEV))
    torch.testing.assert_close(xb.grad.cpu(), xa.grad, rtol=2e-3, atol=2e-2)
    # the weight gradient reads bf16 copies of dY and X: per-term error 2^-8, ~0.1 absolute after 512 terms of N(0,1)
    torch.testing.assert_close(wb.grad.cpu(), wa.grad, rtol=1e-2, atol=0.4)


@pytest.mark.parametrize("B,C,K,HW", [(8, 116, 116, 16), (4, 24, 144, 32), (16, 12, 36, 8), (4, 16, 96, 32)])
def test_pointwise_conv_any_16_byte_pitched_channel_count(G, B, C, K, HW):
    """1x1 convolutions whose channel counts are multiples of 4 but not of 8 / 32 (ShuffleNetV2's 116, DenseNet's 12): K / N tails
    of the GEMM tiles come from TMA zero fill, the weight gradient reads bf16 rows zero-padded to a 16-byte pitch."""
    import torch.nn.functional as TF
    from fedb200.ops import functional as F
    torch.manual_seed(9)
    x = torch.randn(B, C, HW, HW)
    w = torch.randn(K, C, 1, 1) * 0.1
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    xb, wb = x.clone().to(DEV).requires_grad_(True), w.clone().to(DEV).requires_grad_(True)
    ya = TF.conv2d(xa, wa)
    yb = F.conv2d(xb, wb)
    torch.testing.assert_close(yb.cpu(), ya.detach(), rtol=3e-3, atol=3e-3 * float(ya.detach().abs().max()))
    g = torch.randn_like(ya)
    ya.backward(g), yb.backward(g.to(DEV))
    torch.testing.assert_close(xb.grad.cpu(), xa.grad, rtol=3e-3, atol=3e-3 * float(xa.grad.abs().max()))
    torch.testing.assert_close(wb.grad.cpu(), wa.grad, rtol=1e-2, atol=1e-2 * float(wa.grad.abs().max()))


@pytest.mark.parametrize("B,C,K,HW", [(4, 64, 64, 32), (8, 64, 128, 16), (16, 128, 256, 8), (32, 256, 512, 4), (64, 512, 512, 2),
                                      (5, 32, 32, 8), (4, 48, 24, 16), (8, 96, 208, 8), (4, 16, 40, 32), (8, 128, 32, 16),
                                      (4, 48, 12, 32), (8, 116, 116, 8), (4, 60, 20, 16)])     # wgrad through zero-padded bf16 rows
def test_conv3x3_implicit_gemm_fwd_and_dgrad(G, B, C, K, HW):
    """3x3 / stride 1 / pad 1 convolution as an implicit GEMM: every filter tap is a shifted 4-D TMA box of the NHWC
    activation (the zero padding comes from TMA's out-of-bounds fill)."""
    import torch.nn.functional as TF
    from fedb200.ops import functional as F
    torch.manual_seed(4)
    x = torch.randn(B, C, HW, HW)
    w = torch.randn(K, C, 3, 3) * 0.05
    bias = torch.randn(K)
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    xb, wb = x.clone().to(DEV).requires_grad_(True), w.clone().to(DEV).requires_grad_(True)
    ya = TF.conv2d(xa, wa, bias, padding=1)
    yb = F.conv2d(xb, wb, bias.to(DEV), 1, 1, 1)
    scale = float(ya.detach().abs().max())
    torch.testing.assert_close(yb.cpu(), ya.detach(), rtol=3e-3, atol=3e-3 * scale)
    g = torch.randn_like(ya)
    ya.backward(g), yb.backward(g.to(DEV))
    torch.testing.assert_close(xb.grad.cpu(), xa.grad, rtol=3e-3, atol=3e-3 * float(xa.grad.abs().max()))
    torch.testing.assert_close(wb.grad.cpu(), wa.grad, rtol=1e-2, atol=1e-2 * float(wa.grad.abs().max()))


@pytest.mark.parametrize("B,C,K,HW,fs", [(4, 64, 128, 32, 3), (8, 128, 256, 16, 3), (16, 256, 512, 8, 3), (4, 64, 128, 32, 1),
                                         (16, 256, 512, 8, 1), (64, 64, 64, 4, 3), (8, 48, 88, 16, 3), (8, 24, 56, 32, 1)])
def test_strided_conv_on_tensor_cores(G, B, C, K, HW, fs):
    """Stride-2 3x3 / 1x1 convolutions: the TMA tensor map skips every other input pixel (element strides), so the
    forward and the weight gradient reuse the implicit-GEMM kernel; the data gradient is decomposed by the parity of the
    dX pixel into four dense stride-1 implicit GEMMs over dY whose epilogues write the interleaved pixels."""
    import torch.nn.functional as TF
    from fedb200.ops import _lib, functional as F
    torch.manual_seed(5)
    x = torch.randn(B, C, HW, HW)
    w = torch.randn(K, C, fs, fs) * 0.05
    pad = fs // 2
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    xb, wb = x.clone().to(DEV).requires_grad_(True), w.clone().to(DEV).requires_grad_(True)
    ya = TF.conv2d(xa, wa, None, 2, pad)
    yb = F.conv2d(xb, wb, None, 2, pad, 1)
    torch.testing.assert_close(yb.cpu(), ya.detach(), rtol=3e-3, atol=3e-3 * float(ya.detach().abs().max()))
    g = torch.randn_like(ya)
    ya.backward(g), yb.backward(g.to(DEV))
    # the data gradient runs as four parity-class implicit GEMMs in TF32
    torch.testing.assert_close(xb.grad.cpu(), xa.grad, rtol=3e-3, atol=3e-3 * float(xa.grad.abs().max()))
    torch.testing.assert_close(wb.grad.cpu(), wa.grad, rtol=1e-2, atol=1e-2 * float(wa.grad.abs().max()))


@pytest.mark.parametrize("B,K,HW,bias", [(8, 64, 32, False), (4, 32, 32, True), (16, 128, 16, False), (4, 44, 32, False),
                                         (8, 24, 16, False)])
def test_stem_conv_through_im2col_gemm(G, B, K, HW, bias):
    """3-channel stems: Cin*9 = 27 is padded to one 32-wide K tile by an explicit im2col, then the GEMM kernel runs the
    forward (TF32) and the weight gradient (bf16 with dY rows zero-padded to a 16-byte pitch, split over pixels)."""
    import torch.nn.functional as TF
    from fedb200.ops import functional as F
    torch.manual_seed(6)
    x = torch.randn(B, 3, HW, HW)
    w = torch.randn(K, 3, 3, 3) * 0.2
    b = torch.randn(K) if bias else None
    wa = w.clone().requires_grad_(True)
    wb = w.clone().to(DEV).requires_grad_(True)
    ba = b.clone().requires_grad_(True) if bias else None
    bb = b.clone().to(DEV).requires_grad_(True) if bias else None
    ya = TF.conv2d(x, wa, ba, 1, 1)
    yb = F.conv2d(x.to(DEV), wb, bb, 1, 1, 1)
    torch.testing.assert_close(yb.cpu(), ya.detach(), rtol=3e-3, atol=3e-3 * float(ya.detach().abs().max()))
    g = torch.randn_like(ya)
    ya.backward(g), yb.backward(g.to(DEV))
    torch.testing.assert_close(wb.grad.cpu(), wa.grad, rtol=1e-2, atol=1e-2 * float(wa.grad.abs().max()))
    if bias:
        torch.testing.assert_close(bb.grad.cpu(), ba.grad, rtol=1e-3, atol=1e-3 * float(ba.grad.abs().max()))


@pytest.mark.parametrize("B,C,K,HW,fs,stride,groups", [(4, 128, 128, 16, 3, 1, 2), (4, 256, 256, 16, 3, 2, 4), (8, 64, 128, 8, 1, 1, 2),
                                                     (4, 128, 256, 32, 1, 2, 8), (4, 192, 96, 8, 3, 1, 3), (4, 128, 128, 8, 3, 1, 32),
                                                     (8, 152, 152, 8, 3, 2, 19), (4, 368, 368, 4, 3, 1, 46), (4, 96, 192, 16, 3, 2, 8),
                                                     (4, 200, 50, 16, 1, 1, 2), (4, 50, 176, 16, 1, 1, 2), (8, 60, 240, 8, 1, 1, 3)])
def test_grouped_conv_on_tensor_cores(G, B, C, K, HW, fs, stride, groups):
    """Grouped convolutions (ResNeXt, RegNet, DPN): all groups in one implicit-GEMM launch; the group index shifts the channel
    coordinate of the TMA boxes, spill-over into the next group meets zeros of the packed filter or the epilogue masks."""
    import torch.nn.functional as TF
    from fedb200.ops import functional as F
    torch.manual_seed(8)
    x = torch.randn(B, C, HW, HW)
    w = torch.randn(K, C // groups, fs, fs) * 0.05
    pad = fs // 2
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    xb, wb = x.clone().to(DEV).requires_grad_(True), w.clone().to(DEV).requires_grad_(True)
    ya = TF.conv2d(xa, wa, None, stride, pad, 1, groups)
    yb = F.conv2d(xb, wb, None, stride, pad, groups)
    torch.testing.assert_close(yb.cpu(), ya.detach(), rtol=3e-3, atol=3e-3 * float(ya.detach().abs().max()))
    g = torch.randn_like(ya)
    ya.backward(g), yb.backward(g.to(DEV))
    torch.testing.assert_close(xb.grad.cpu(), xa.grad, rtol=3e-3, atol=3e-3 * float(xa.grad.abs().max()))
    torch.testing.assert_close(wb.grad.cpu(), wa.grad, rtol=1e-2, atol=1e-2 * float(wa.grad.abs().max()))
