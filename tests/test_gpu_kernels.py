"""Numerics of the hand-written CUDA kernels against plain PyTorch fp32 references (run with
``-m gpu`` on a B200).  Aggregation is checked BIT-EXACT against the CPU oracle."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from fedb200 import models
from fedb200.parallel.arena import Arena
from fedb200.parallel.fedavg import aggregate_arenas, fedavg_state_dicts
from fedb200.utils import data as D

DEV = "cuda:0"


@pytest.fixture(scope="module")
def K():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fedb200.ops import cuda
    return cuda


def _client_models(n, name):
    ms = []
    for i in range(n):
        torch.manual_seed(10 + i)
        m = models.build(name)
        for k, v in m.state_dict().items():
            if k.endswith("num_batches_tracked"):
                v.fill_(190 + 3 * i)
            elif "running" in k:
                v.copy_(torch.randn_like(v).abs())
        ms.append(m)
    return ms


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
def test_fedavg_bit_exact_vs_cpu_oracle(K, n):
    ms = _client_models(n, "MobileNet")
    sds = [{k: v.clone() for k, v in m.state_dict().items()} for m in ms]
    arenas = [Arena.adopt(m.to(DEV), DEV) for m in ms]
    out, extra = Arena(arenas[0].manifest, DEV), Arena(arenas[0].manifest, DEV)
    K.fedavg(out, arenas, extra_dst=[extra])
    torch.cuda.synchronize()
    ref = models.MobileNet()
    ref.load_state_dict(fedavg_state_dicts(sds))          # N3 truncation happens in this load
    for k, v in ref.state_dict().items():
        assert torch.equal(out.view(k).cpu(), v), k
        assert torch.equal(extra.view(k).cpu(), v), k      # broadcast destination got the same bits


def test_fedavg_weighted_and_in_place_destinations(K):
    ms = _client_models(3, "LeNet")
    sds = [{k: v.clone() for k, v in m.state_dict().items()} for m in ms]
    arenas = [Arena.adopt(m.to(DEV), DEV) for m in ms]
    w = [6272, 6144, 6272]
    out = Arena(arenas[0].manifest, DEV)
    K.fedavg(out, arenas, weights=w, extra_dst=arenas)     # clients are sources AND destinations
    torch.cuda.synchronize()
    want = fedavg_state_dicts(sds, weights=w)
    for k, v in want.items():
        assert torch.equal(out.view(k).cpu(), v), k
        for a in arenas:
            assert torch.equal(a.view(k).cpu(), v), k


def test_fedavg_bf16_shadow_and_aggregate_arenas_dispatch(K):
    ms = _client_models(2, "LeNet")
    arenas = [Arena.adopt(m.to(DEV), DEV) for m in ms]
    out = Arena(arenas[0].manifest, DEV)
    shadow = torch.zeros(out.manifest.n_f32, dtype=torch.bfloat16, device=DEV)
    K.fedavg(out, arenas, bf16_shadows=[shadow])
    torch.cuda.synchronize()
    assert torch.equal(shadow, out.f32.to(torch.bfloat16))
    out2 = Arena(arenas[0].manifest, DEV)
    aggregate_arenas(arenas, out2)                          # product path picks the kernel on CUDA
    assert torch.equal(out2.f32, out.f32)


@pytest.mark.parametrize("how", ["timeout", "abort"])
def test_fedavg_failed_flag_wait_leaves_every_arena_untouched(K, how):
    """A peer that never arrives (dead worker) or an abort: the launch must neither reduce nor fan out (SURVEY B3 must
    not come back through the failure path), must record the reason and must not advance the epoch."""
    from fedb200.ops import _lib
    ms = _client_models(2, "LeNet")
    arenas = [Arena.adopt(m.to(DEV), DEV) for m in ms]
    before = [a.f32.clone() for a in arenas]
    before_i = [a.i64.clone() for a in arenas]
    ctrl = [K.new_ctrl(DEV), K.new_ctrl(DEV)]           # rank 0 launches, "rank 1" never does
    if how == "abort":
        ctrl[0][_lib.CTRL_OFF["abort_flag"] // 4] = 1
    n = arenas[0].manifest.n_f32
    K.fedavg_raw([a.f32.data_ptr() for a in arenas], [a.f32.data_ptr() for a in arenas], 0, n,
                 ctrl_local=ctrl[0].data_ptr(), src_i64=[a.i64.data_ptr() for a in arenas],
                 dst_i64=[a.i64.data_ptr() for a in arenas], n_i64=arenas[0].manifest.n_i64,
                 sync_mode=1, rank=0, world=2, ctrl_peer=[c.data_ptr() for c in ctrl],
                 timeout_ns=2_000_000 if how == "timeout" else 5_000_000_000)
    torch.cuda.synchronize()
    assert K.ctrl_field(ctrl[0], "status") == (1 if how == "timeout" else 3)
    assert K.ctrl_field(ctrl[0], "epoch") == 0 and K.ctrl_field(ctrl[0], "blocks_done") == 0
    for a, b, bi in zip(arenas, before, before_i):
        assert torch.equal(a.f32, b) and torch.equal(a.i64, bi)
    # the announcement was withdrawn from the peer's block: a late peer cannot finish the round alone
    assert int(ctrl[1][0]) == 0 and int(ctrl[1][16]) == 0          # arrive[0], done[0] as seen by rank 1


def test_fedavg_flag_protocol_two_ranks_on_one_gpu(K):
    """Both ranks of a 2-rank K-AGG launched on one GPU (two streams): bit-exact mean in both arenas, epoch advanced."""
    ms = _client_models(2, "LeNet")
    sds = [{k: v.clone() for k, v in m.state_dict().items()} for m in ms]
    arenas = [Arena.adopt(m.to(DEV), DEV) for m in ms]
    ctrl = [K.new_ctrl(DEV), K.new_ctrl(DEV)]
    n = arenas[0].manifest.n_f32
    half = (n // 2) // 4 * 4
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    for r, (lo, hi) in enumerate([(0, half), (half, n)]):
        with torch.cuda.stream(streams[r]):
            K.fedavg_raw([a.f32.data_ptr() for a in arenas], [a.f32.data_ptr() for a in arenas], lo, hi,
                         ctrl_local=ctrl[r].data_ptr(), src_i64=[a.i64.data_ptr() for a in arenas],
                         dst_i64=[a.i64.data_ptr() for a in arenas], n_i64=arenas[0].manifest.n_i64 if r == 0 else 0,
                         sync_mode=1, rank=r, world=2, ctrl_peer=[c.data_ptr() for c in ctrl],
                         timeout_ns=5_000_000_000, max_blocks=64)
    torch.cuda.synchronize()
    want = fedavg_state_dicts(sds)
    for c in ctrl:
        assert K.ctrl_field(c, "status") == 0 and K.ctrl_field(c, "epoch") == 1
    for a in arenas:
        for k, v in want.items():
            assert torch.equal(a.view(k).cpu(), v), k


def test_replicate_stamps_version(K):
    m = models.MobileNet().to(DEV)
    src = Arena.adopt(m, DEV)
    dst = Arena(src.manifest, DEV)
    ctrl = K.new_ctrl(DEV)
    K.replicate(dst, src, ctrl, version=7)
    torch.cuda.synchronize()
    assert torch.equal(dst.f32, src.f32) and torch.equal(dst.i64, src.i64)
    assert K.ctrl_field(ctrl, "version") == 7 and K.ctrl_field(ctrl, "blocks_done") == 0


def test_codec_error_feedback_and_fused_decode(K):
    torch.manual_seed(0)
    n = 4096 * 3
    base = torch.randn(n, device=DEV)
    curs = [base + 0.01 * torch.randn(n, device=DEV) for _ in range(2)]
    qs, scales, res = [], [], []
    for cur in curs:
        q = torch.zeros(n, dtype=torch.uint8, device=DEV)
        s = torch.zeros(n // 32, device=DEV)
        r = torch.zeros(n, device=DEV)
        K.codec_encode(cur, base, r, q, s)
        qs.append(q), scales.append(s), res.append(r)
    dec = [torch.empty(n, device=DEV) for _ in curs]
    for d, q, s in zip(dec, qs, scales):
        K.codec_decode(d, base, q, s)
    torch.cuda.synchronize()
    for cur, d, r in zip(curs, dec, res):
        delta = cur - base
        # e4m3 has 3 mantissa bits: relative error <= 2^-4 of the block amax
        amax = delta.view(-1, 32).abs().amax(1, keepdim=True).expand(-1, 32).reshape(-1)
        assert float(((d - cur).abs() / amax).max()) <= 2 ** -4 + 1e-6
        # error feedback: decoded + residual reproduces the input delta
        torch.testing.assert_close((d - base) + r, delta, rtol=0, atol=1e-6)
    # fused decode inside K-AGG == mean of the separately decoded models
    ctrl = K.new_ctrl(DEV)
    out = torch.empty(n, device=DEV)
    K.fedavg_raw([], [out.data_ptr()], 0, n, ctrl_local=ctrl.data_ptr(), src_q=[q.data_ptr() for q in qs],
                 src_scale=[s.data_ptr() for s in scales], base=base.data_ptr())
    torch.cuda.synchronize()
    want = base + ((dec[0] - base) + (dec[1] - base)) / 2
    torch.testing.assert_close(out, want, rtol=0, atol=1e-6)
    # second round: the residual is consumed (bias does not accumulate)
    q2 = torch.zeros(n, dtype=torch.uint8, device=DEV)
    s2 = torch.zeros(n // 32, device=DEV)
    r_before = res[0].clone()
    K.codec_encode(curs[0], base, res[0], q2, s2)
    d2 = torch.empty(n, device=DEV)
    K.codec_decode(d2, base, q2, s2)
    torch.cuda.synchronize()
    two_round_mean = ((dec[0] - base) + (d2 - base)) / 2
    one_round_err = ((dec[0] - base) - (curs[0] - base)).abs().mean()
    assert float((two_round_mean - (curs[0] - base) - r_before / 2).abs().mean()) <= float(one_round_err)


def test_codec_kernels_match_the_cpu_oracle_bit_for_bit(K):
    """K-CODEC encode / decode / fused decode in K-AGG against fedb200/parallel/codec.py: same bytes, same scales, same
    residuals (e4m3 round-to-nearest-even, IEEE division), over two rounds of error feedback and a ragged tail block."""
    from fedb200.parallel import codec
    torch.manual_seed(3)
    n = 32 * 257 + 20
    base = torch.randn(n)
    curs = [base + 0.02 * torch.randn(n) * torch.rand(n).pow(4) for _ in range(3)]     # wide dynamic range inside blocks
    res_cpu = [torch.zeros(n) for _ in curs]
    res_gpu = [torch.zeros(n, device=DEV) for _ in curs]
    base_d = base.to(DEV)
    nb = (n + 31) // 32
    for rnd in range(2):
        qs_d, ss_d, qs_c, ss_c = [], [], [], []
        for cur, rc, rg in zip(curs, res_cpu, res_gpu):
            cur = cur + 0.001 * rnd
            q_c, s_c = codec.encode(cur, base, rc)
            q_d = torch.zeros(n, dtype=torch.uint8, device=DEV)
            s_d = torch.zeros(nb, device=DEV)
            K.codec_encode(cur.to(DEV), base_d, rg, q_d, s_d)
            torch.cuda.synchronize()
            assert torch.equal(s_d.cpu(), s_c), f"scales, round {rnd}"
            assert torch.equal(q_d.cpu(), q_c), f"fp8 bytes, round {rnd}"
            assert torch.equal(rg.cpu(), rc), f"residual, round {rnd}"
            out_d = torch.empty(n, device=DEV)
            K.codec_decode(out_d, base_d, q_d, s_d)
            assert torch.equal(out_d.cpu(), codec.decode(base, q_c, s_c))
            qs_d.append(q_d), ss_d.append(s_d), qs_c.append(q_c), ss_c.append(s_c)
        n4 = n // 4 * 4                                  # K-AGG ranges are multiples of four elements
        ctrl = K.new_ctrl(DEV)
        agg = torch.zeros(n, device=DEV)
        K.fedavg_raw([], [agg.data_ptr()], 0, n4, ctrl_local=ctrl.data_ptr(), src_q=[q.data_ptr() for q in qs_d],
                     src_scale=[x.data_ptr() for x in ss_d], base=base_d.data_ptr())
        torch.cuda.synchronize()
        assert torch.equal(agg[:n4].cpu(), codec.fedavg_decoded(base, qs_c, ss_c)[:n4]), f"fused decode, round {rnd}"


def test_sgd_matches_torch(K):
    torch.manual_seed(0)
    n = 8192
    p0 = torch.randn(n)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref], lr=0.1, momentum=0.9, weight_decay=5e-4)
    p, v = p0.clone().to(DEV), torch.zeros(n, device=DEV)
    shadow = torch.zeros(n, dtype=torch.bfloat16, device=DEV)
    for step in range(4):
        g = torch.randn(n)
        ref.grad = g.clone()
        opt.step()
        K.sgd_step(p, g.to(DEV), v, 0.1, 0.9, 5e-4, shadow_bf16=shadow)
    torch.cuda.synchronize()
    torch.testing.assert_close(p.cpu(), ref.data, rtol=1e-6, atol=1e-6)
    assert torch.equal(shadow, p.to(torch.bfloat16))


def test_cross_entropy_matches_torch(K):
    torch.manual_seed(0)
    for B, C in ((128, 10), (100, 10), (7, 1000)):
        z = torch.randn(B, C)
        y = torch.randint(0, C, (B,))
        zr = z.clone().requires_grad_(True)
        lr = torch.nn.functional.cross_entropy(zr, y)
        lr.backward()
        zd = z.to(DEV).requires_grad_(True)
        stats = torch.zeros(4, device=DEV)
        loss = K.cross_entropy(zd, y.to(DEV))
        loss.backward()
        torch.testing.assert_close(loss.cpu(), lr.detach(), rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(zd.grad.cpu(), zr.grad, rtol=1e-4, atol=1e-7)
        K.cross_entropy_raw(z.to(DEV), y.to(DEV), None, None, stats)
        assert int(stats[1]) == int((z.argmax(1) == y).sum()) and int(stats[2]) == B


def test_prepare_batch_matches_normalize(K):
    ds = D.synthetic_cifar(64, 8, seed=1)
    x = ds.train_x[:64].to(DEV)
    want = D.normalize_batch(ds.train_x[:64])
    torch.testing.assert_close(K.prepare_batch(x).cpu(), want, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(K.prepare_batch(x, nhwc=True).cpu(), want, rtol=1e-6, atol=1e-6)
    aug = K.prepare_batch(x, augment=True, seed=5).cpu()
    assert aug.shape == want.shape and not torch.equal(aug, want)
    assert torch.equal(aug, K.prepare_batch(x, augment=True, seed=5).cpu())   # pure function of the seed


# ------------------------------------------------------------------------------------ LeNet
def _lenet_pair(seed=0):
    torch.manual_seed(seed)
    ref = models.LeNet()
    gpu = models.LeNet()
    gpu.load_state_dict(ref.state_dict())
    gpu = gpu.to(DEV)
    arena = Arena.adopt(gpu, DEV, with_optimizer_state=True)
    return ref, gpu, arena


def test_lenet_forward_matches_torch(K):
    ref, _, arena = _lenet_pair()
    plan = K.LeNetPlan(arena)
    ds = D.synthetic_cifar(256, 8, seed=2)
    xu8 = ds.train_x[:200]
    xf = D.normalize_batch(xu8)
    with torch.no_grad():
        want = ref(xf)
    torch.testing.assert_close(plan.forward(xu8.to(DEV)).cpu(), want, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(plan.forward(xf.to(DEV)).cpu(), want, rtol=1e-4, atol=1e-5)


def test_lenet_gradient_matches_autograd(K):
    ref, _, arena = _lenet_pair(1)
    plan = K.LeNetPlan(arena)
    for B in (128, 80, 5):
        x = torch.randn(B, 3, 32, 32)
        y = torch.randint(0, 10, (B,))
        ref.zero_grad()
        loss = torch.nn.functional.cross_entropy(ref(x), y)
        loss.backward()
        dparam, l = plan.grad(x.to(DEV), y.to(DEV))
        torch.cuda.synchronize()
        torch.testing.assert_close(l.cpu(), loss.detach(), rtol=1e-5, atol=1e-6)
        for k, p in ref.named_parameters():
            e = arena.manifest.by_key[k]
            got = dparam[e.offset:e.offset + e.numel].view(e.shape).cpu()
            torch.testing.assert_close(got, p.grad, rtol=2e-4, atol=2e-6, msg=k)


def test_lenet_train_steps_match_torch_sgd(K):
    ref, _, arena = _lenet_pair(2)
    plan = K.LeNetPlan(arena, lr=0.1, momentum=0.9, weight_decay=5e-4)
    opt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    ds = D.synthetic_cifar(128 * 3, 8, seed=3)
    tot_loss, tot_correct = 0.0, 0
    for s in range(3):
        xu8, y = ds.train_x[s * 128:(s + 1) * 128], ds.train_y[s * 128:(s + 1) * 128]
        opt.zero_grad()
        out = ref(D.normalize_batch(xu8))
        loss = torch.nn.functional.cross_entropy(out, y)
        loss.backward()
        opt.step()
        tot_loss += float(loss)
        tot_correct += int((out.argmax(1) == y).sum())
        plan.train_step(xu8.to(DEV), y.to(DEV), step_seed=s)
    torch.cuda.synchronize()
    for k, p in ref.named_parameters():
        torch.testing.assert_close(arena.view(k).cpu(), p.data, rtol=2e-4, atol=2e-6, msg=k)
    st = plan.stats.cpu()
    assert abs(float(st[0]) - tot_loss) < 1e-3 and int(st[1]) == tot_correct and int(st[2]) == 384 and int(st[3]) == 3


@pytest.mark.parametrize("n", [1000, 503, 3])
def test_lenet_eval_counts(K, n):
    """The evaluation kernel packs five samples per CTA: sizes that are / are not multiples of five, and a lone CTA."""
    ref, _, arena = _lenet_pair(3)
    plan = K.LeNetPlan(arena)
    ds = D.synthetic_cifar(8, n, seed=4)
    stats = torch.zeros(4, device=DEV)
    plan.eval_batch(ds.test_x.to(DEV), ds.test_y.to(DEV), stats)
    with torch.no_grad():
        out = ref(D.normalize_batch(ds.test_x))
    assert int(stats[1]) == int((out.argmax(1) == ds.test_y).sum()) and int(stats[2]) == n
    assert abs(float(stats[0]) - float(torch.nn.functional.cross_entropy(out, ds.test_y))) < 1e-4


def test_gpu_trainer_pass_matches_cpu_trainer(K):
    from fedb200.control.trainer import LocalTrainer, TrainerConfig
    ds = D.synthetic_cifar(128 * 6 + 40, 300, seed=5)
    cpu = LocalTrainer(TrainerConfig(model="LeNet", device="cpu", seed=9), ds)
    gpu = LocalTrainer(TrainerConfig(model="LeNet", device=DEV, seed=9), ds)
    gpu.load_state_dict(cpu.state_dict())
    for rnd in range(2):                       # round 2 replays the captured CUDA graph
        a, b = cpu.train(1, 1, 2), gpu.train(1, 1, 2)
        assert (a.steps, a.samples) == (b.steps, b.samples)
        assert abs(a.mean_loss - b.mean_loss) < 1e-3 and abs(a.correct - b.correct) <= 2
    for k, v in cpu.state_dict().items():
        torch.testing.assert_close(gpu.state_dict()[k], v, rtol=1e-3, atol=1e-5, msg=k)
    ea, eb = cpu.test(), gpu.test()
    assert ea.samples == eb.samples == 300 and abs(ea.correct - eb.correct) <= 2
    eb2 = gpu.test()                           # graph replay of the eval pass
    assert eb2.correct == eb.correct


def test_gpu_trainer_generic_model_eval_graph(K):
    """A zoo model other than LeNet through the trainer: the evaluation pass (batches of 500, one CUDA graph) gives the
    CPU trainer's loss / accuracy, and its replay gives the same numbers again."""
    from fedb200.control.trainer import LocalTrainer, TrainerConfig
    ds = D.synthetic_cifar(128, 1100, seed=6)
    cpu = LocalTrainer(TrainerConfig(model="MobileNet", device="cpu", seed=3), ds)
    gpu = LocalTrainer(TrainerConfig(model="MobileNet", device=DEV, seed=3), ds)
    gpu.load_state_dict(cpu.state_dict())
    ea, eb = cpu.test(), gpu.test()
    assert ea.samples == eb.samples == 1100
    assert abs(ea.mean_loss - eb.mean_loss) < 5e-2 * max(1.0, abs(ea.mean_loss)) and abs(ea.correct - eb.correct) <= 30
    eb2 = gpu.test()
    assert eb2.correct == eb.correct and eb2.loss_sum == eb.loss_sum


def test_mem_manager_roundtrip_single_process(K):
    import ctypes as C
    from fedb200.ops import _lib
    L = _lib.lib()
    vmm, fd_ok, mc = C.c_int(), C.c_int(), C.c_int()
    _lib.check(L.fedb_mem_supported(0, C.byref(vmm), C.byref(fd_ok), C.byref(mc)))
    assert vmm.value == 1 and fd_ok.value == 1
    a, b = _lib.Mem(), _lib.Mem()
    _lib.check(L.fedb_mem_create(3 << 20, 0, C.byref(a)))
    _lib.check(L.fedb_mem_export_fd(C.byref(a)))
    assert a.fd >= 0 and a.bytes % (2 << 20) == 0
    _lib.check(L.fedb_mem_import_fd(a.fd, a.bytes, 0, C.byref(b)))       # second mapping of the same memory
    from fedb200.parallel.mem import tensor_from_ptr
    ta = tensor_from_ptr(a.ptr, 1024, torch.float32, 0)
    tb = tensor_from_ptr(b.ptr, 1024, torch.float32, 0)
    ta.copy_(torch.arange(1024, dtype=torch.float32))
    torch.cuda.synchronize()
    assert torch.equal(tb.cpu(), torch.arange(1024, dtype=torch.float32))
    del ta, tb
    _lib.check(L.fedb_mem_release(C.byref(b)))
    _lib.check(L.fedb_mem_release(C.byref(a)))
