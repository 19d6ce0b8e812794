"""The allocation-free cross-entropy entry point, called through the C ABI.

Why it exists: some hosts of this library cannot allocate device memory at the moment they need a
loss - a trainer whose GPU is already fully owned by its own memory pool (a library cudaMalloc then
fails or fragments), or a latency-critical round/eval loop (cudaMalloc/cudaFree synchronise the
device).  ``fedb_cross_entropy_static`` does the same computation as ``fedb_cross_entropy`` with its
row scratch in a fixed 128 KiB array inside the library, for batches of at most 16384 rows.

The test drives both entry points with ctypes, twice in a row on different inputs (the second, smaller
batch would expose stale scratch rows), compares everything they produce, and checks that oversized
batches are refused with a readable message.  This file sorts last on purpose: with ``-x`` every other
test has already run when it starts.  It is self-contained, so it passes alone in a fresh process too.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 16384
vp = C.c_void_p
SIG = [vp, vp, vp, vp, vp, C.c_int, C.c_int, vp]


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fedb200.ops import _lib
    L = _lib.lib()
    for name in ("fedb_cross_entropy", "fedb_cross_entropy_static"):
        fn = getattr(L, name)
        fn.restype, fn.argtypes = C.c_int, SIG
    return L


def _run(fn, logits, target):
    dlogits = torch.full_like(logits, float("nan"))
    loss = torch.full((1,), float("nan"), device=logits.device)
    stats = torch.zeros(4, device=logits.device)
    rc = fn(logits.data_ptr(), target.data_ptr(), dlogits.data_ptr(), loss.data_ptr(), stats.data_ptr(),
            logits.shape[0], logits.shape[1], torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return rc, dlogits, loss, stats


def test_static_scratch_entry_point_matches_allocating_one(lib):
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    # first call on this device does the one-time address lookup (no capture is active here)
    for call, (B, Cn) in enumerate(((4096, 10), (100, 10), (7, 1000))):
        logits = torch.randn(B, Cn, device=dev) * 3
        target = torch.randint(0, Cn, (B,), device=dev)
        rc_a, d_a, l_a, s_a = _run(lib.fedb_cross_entropy, logits, target)
        free_before, _ = torch.cuda.mem_get_info(dev)
        rc_b, d_b, l_b, s_b = _run(lib.fedb_cross_entropy_static, logits, target)
        free_after, _ = torch.cuda.mem_get_info(dev)
        assert rc_a == 0 and rc_b == 0, lib.fedb_last_error().decode()
        torch.testing.assert_close(d_b, d_a, rtol=1e-5, atol=1e-7)
        torch.testing.assert_close(l_b, l_a, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(s_b, s_a, rtol=1e-5, atol=1e-6)
        # and both agree with plain PyTorch
        want = torch.nn.functional.cross_entropy(logits.cpu(), target.cpu())
        torch.testing.assert_close(l_b.cpu()[0], want, rtol=1e-5, atol=1e-6)
        assert int(s_b[1]) == int((logits.argmax(1) == target).sum()) and int(s_b[2]) == B
        if call > 0:            # steady state: the call did not take memory from the device
            assert free_after == free_before


def test_static_scratch_entry_point_refuses_oversized_batches(lib):
    dev = torch.device("cuda", 0)
    B = MAX_ROWS + 1
    logits = torch.zeros(B, 10, device=dev)
    target = torch.zeros(B, dtype=torch.int64, device=dev)
    rc, dlogits, loss, stats = _run(lib.fedb_cross_entropy_static, logits, target)
    assert rc != 0
    msg = lib.fedb_last_error().decode()
    assert "FEDB_CE_STATIC_MAX_ROWS" in msg and "16384" in msg
    assert float(stats.abs().sum()) == 0.0 and bool(torch.isnan(loss).all())   # nothing was launched
    # exactly at the limit is accepted
    ok = _run(lib.fedb_cross_entropy_static, logits[:MAX_ROWS], target[:MAX_ROWS])
    assert ok[0] == 0 and int(ok[3][2]) == MAX_ROWS
