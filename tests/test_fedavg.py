"""Aggregation oracle: N1-N3 semantics, flat-arena path, weighting."""
from collections import OrderedDict

import pytest
import torch

from fedb200 import models
from fedb200.parallel.arena import Arena, Manifest, split_ranges
from fedb200.parallel.fedavg import aggregate_arenas, fedavg_state_dicts


def literal_reference_allreduce(sds):
    """Line-by-line restatement of reference src/server.py:163-171."""
    opt = OrderedDict((k, v) for k, v in sds[0].items())
    for idx in range(1, len(sds)):
        for key in sds[idx]:
            opt[key] = opt[key] + sds[idx][key]
    for key in opt:
        opt[key] = opt[key] / len(sds)
    return opt


def _clients(n, name="MobileNet", counters=(195, 196, 49, 48, 49, 49, 49, 49)):
    ms = []
    for i in range(n):
        torch.manual_seed(100 + i)
        m = models.build(name)
        for k, v in m.state_dict().items():
            if k.endswith("num_batches_tracked"):
                v.fill_(counters[i % len(counters)])
            elif "running" in k:
                v.copy_(torch.randn_like(v).abs())
        ms.append(m)
    return ms


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
def test_oracle_equals_literal_reference(n):
    sds = [m.state_dict() for m in _clients(n, "LeNet" if n > 4 else "MobileNet")]
    a, b = fedavg_state_dicts(sds), literal_reference_allreduce(sds)
    for k in a:
        assert a[k].dtype == b[k].dtype and torch.equal(a[k], b[k]), k


def test_int64_counters_become_fp32_then_truncate():
    sds = [m.state_dict() for m in _clients(2)]
    avg = fedavg_state_dicts(sds)
    k = "bn1.num_batches_tracked"
    assert avg[k].dtype == torch.float32 and float(avg[k]) == 195.5
    m = models.MobileNet()
    m.load_state_dict(avg)
    assert m.state_dict()[k].dtype == torch.int64 and int(m.state_dict()[k]) == 195


@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_flat_arena_path_is_bit_exact(n):
    ms = _clients(n)
    sds = [{k: v.clone() for k, v in m.state_dict().items()} for m in ms]
    arenas = [Arena.adopt(m) for m in ms]
    out = Arena(arenas[0].manifest)
    aggregate_arenas(arenas, out)
    g = models.MobileNet()
    g.load_state_dict(fedavg_state_dicts(sds))
    for k, v in g.state_dict().items():
        assert torch.equal(v, out.view(k)), k


def test_uniform_weights_degrade_to_unweighted_bitwise():
    sds = [m.state_dict() for m in _clients(3, "LeNet")]
    a = fedavg_state_dicts(sds, weights=[6272, 6272, 6272])
    b = fedavg_state_dicts(sds)
    assert all(torch.equal(a[k], b[k]) for k in a)


def test_weighted_mean():
    sds = [m.state_dict() for m in _clients(2, "LeNet")]
    w = [100, 300]
    a = fedavg_state_dicts(sds, weights=w)
    k = "fc1.weight"
    torch.testing.assert_close(a[k], (sds[0][k] * 100 + sds[1][k] * 300) / 400)


def test_manifest_is_deterministic_and_aligned():
    m1, m2 = Manifest.from_module(models.ResNet18()), Manifest.from_module(models.ResNet18())
    assert m1.fingerprint() == m2.fingerprint()
    assert m1.n_param % 1024 == 0 and m1.n_f32 % 1024 == 0
    assert all(e.offset % 32 == 0 for e in m1.entries if e.segment != "i64")
    assert [e.key for e in m1.entries] == list(models.ResNet18().state_dict().keys())
    rs = split_ranges(m1.n_f32, 8)
    assert rs[0][0] == 0 and rs[-1][1] == m1.n_f32 and all(a % 1024 == 0 for a, _ in rs)


def test_arena_keeps_module_live():
    m = models.LeNet()
    a = Arena.adopt(m, with_optimizer_state=True)
    a.f32.zero_()
    assert float(m.fc3.weight.abs().sum()) == 0.0
    x = torch.randn(2, 3, 32, 32)
    m(x).sum().backward()
    assert m.fc3.bias.grad.data_ptr() == a.grad_view("fc3.bias").data_ptr()
    assert float(a.grad.abs().sum()) > 0


def test_roles_follow_the_live_set():
    """No rank is special on the GPU plane: the global model and its replica move to the lowest live ranks."""
    from fedb200.parallel.arena import assign_roles
    assert assign_roles(range(8)) == (0, 1)
    assert assign_roles([1, 2, 3]) == (1, 2)              # rank 0 (the usual host of the global model) is gone
    assert assign_roles([0, 2, 3]) == (0, 2)              # the usual backup is gone
    assert assign_roles([2, 3, 7]) == (2, 3)              # both are gone
    assert assign_roles([5]) == (5, 5)                    # a single survivor holds both
    assert assign_roles([0, 1, 2], pref_primary=2, pref_backup=0) == (2, 0)
    assert assign_roles([0, 1], pref_primary=1, pref_backup=1) == (1, 0)
    import pytest
    with pytest.raises(ValueError):
        assign_roles([])


def test_fs_min_barrier_is_an_allreduce_min(tmp_path):
    """The rendezvous of gRPC-driven workers (no torch.distributed group): every rank ends with the minimum flag."""
    import threading
    from fedb200.parallel.mem import fs_min_barrier
    world, out = 3, {}

    def rank_body(r):
        bar = fs_min_barrier(str(tmp_path), "t", r, world, timeout_s=20)
        f = torch.tensor([1])
        bar(f)                                  # stage 1: everybody says 1
        first = int(f)
        f = torch.tensor([0 if r == 1 else 1])
        bar(f)                                  # stage 2: rank 1 says 0
        out[r] = (first, int(f))

    ts = [threading.Thread(target=rank_body, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join(30) for t in ts]
    assert out == {0: (1, 0), 1: (1, 0), 2: (1, 0)}


def test_codec_oracle_error_feedback_and_bounds():
    """CPU restatement of K-CODEC: e4m3 keeps 3 mantissa bits (error <= 2^-4 of the block maximum), what a round drops is
    carried to the next one, and a constant signal converges to the exact value on average."""
    from fedb200.parallel import codec
    torch.manual_seed(0)
    n = 32 * 40 + 7                                   # a ragged tail block
    base = torch.randn(n)
    cur = base + 0.01 * torch.randn(n)
    res = torch.zeros(n)
    q, s = codec.encode(cur, base, res)
    assert q.dtype == torch.uint8 and q.numel() == n and s.numel() == 41
    out = codec.decode(base, q, s)
    delta = cur - base
    amax = torch.nn.functional.pad(delta.abs(), (0, 25)).view(-1, 32).amax(1, keepdim=True).expand(-1, 32).reshape(-1)[:n]
    assert float(((out - cur).abs() / amax).max()) <= 2 ** -4 + 1e-6
    torch.testing.assert_close((out - base) + res, delta, rtol=0, atol=1e-6)      # nothing is lost, only postponed
    # the same delta sent for many rounds: the running mean of what arrives converges to it (no accumulating bias)
    res = torch.zeros(n)
    got = torch.zeros(n)
    for _ in range(64):
        q, s = codec.encode(cur, base, res)
        got += codec.decode(base, q, s) - base
    assert float(((got / 64 - delta).abs() / amax).max()) < 2 ** -4 / 16
    # an all-zero block encodes to zeros with scale 1 and decodes to the base
    q, s = codec.encode(base, base, torch.zeros(n))
    assert int(q.sum()) == 0 and torch.equal(s, torch.ones_like(s)) and torch.equal(codec.decode(base, q, s), base)
    # K-AGG's fused decode: mean of the decoded deltas on top of the base
    curs = [base + 0.01 * torch.randn(n) for _ in range(3)]
    enc = [codec.encode(c, base, torch.zeros(n)) for c in curs]
    want = base + ((codec.decode(base, *enc[0]) - base) + (codec.decode(base, *enc[1]) - base) + (codec.decode(base, *enc[2]) - base)) / 3
    torch.testing.assert_close(codec.fedavg_decoded(base, [e[0] for e in enc], [e[1] for e in enc]), want, rtol=0, atol=1e-6)
