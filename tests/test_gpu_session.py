"""GpuFederation on one GPU (world = 1): the public round API that bench.py times.

The reference's round (src/server.py:120-153) with a single client is: local pass, "average" of one model, push back,
evaluation.  Checks: the streamed-upload graphs compute what the resident-input graphs compute, and the pipelined
result read-back returns the same numbers as the blocking one.
"""
import pytest
import torch

from fedb200.utils import data as D

pytestmark = pytest.mark.gpu
_TAG = iter(range(1 << 30))


def _fed(seed, **kw):
    from fedb200.parallel.session import GpuFederation
    ds = D.synthetic_cifar(128 * 20 + 17, 1003, seed=3)
    return GpuFederation(model="LeNet", rank=0, world=1, device=0, dataset=ds, seed=seed, tag=f"t-sess-{seed}-{next(_TAG)}", **kw)


@pytest.mark.parametrize("eval_mode", ["overlap", "inline"])
def test_round_is_reproducible_across_upload_modes_and_async(eval_mode):
    a, b, c = _fed(11, eval_mode="inline"), _fed(11, eval_mode=eval_mode), _fed(11, eval_mode=eval_mode)
    outs = {"sync": [], "resident": [], "async": []}
    pending = None
    for r in range(4):
        outs["sync"].append(a.run_round(upload=True, sync=True))
        if r == 0:
            outs["resident"].append(b.run_round(upload=True, sync=True))     # first round brings the inputs over
        else:
            b.run_round(upload=False, sync=False)
            outs["resident"].append(b.engine.read_round_stats())
        h = c.run_round_async(upload=True)
        if pending is not None:
            outs["async"].append(pending.result())
        pending = h
    outs["async"].append(pending.result())
    last_eval = c.drain()
    lag = 1 if eval_mode == "overlap" else 0          # the overlapped evaluation is reported one round late
    for mode in ("resident", "async"):
        for r, ((t0, e0), (t1, e1)) in enumerate(zip(outs["sync"], outs[mode])):
            assert t0.steps == t1.steps == 21 and t0.samples == t1.samples == 128 * 20 + 17
            assert t0.loss_sum == t1.loss_sum and t0.correct == t1.correct, mode
            if mode == "async" and lag:
                if r == 0:
                    assert e1.samples == 0
                    continue
                e0 = outs["sync"][r - 1][1]
            assert e0.samples == e1.samples == 1003 and e0.loss_sum == e1.loss_sum and e0.correct == e1.correct, (mode, r)
    e_last = outs["sync"][-1][1]
    assert last_eval.samples == 1003 and last_eval.loss_sum == e_last.loss_sum and last_eval.correct == e_last.correct
    pa = a.trainer.arena.f32.clone()
    assert torch.equal(pa, b.trainer.arena.f32) and torch.equal(pa, c.trainer.arena.f32)
    assert a.h2d_bytes_per_round() == (128 * 20 + 17) * (3072 + 8) + 1003 * (3072 + 8)
    for f in (a, b, c):
        f.close()


def test_overlapped_eval_of_a_zoo_model_scores_the_snapshot():
    """Generic-model path: the second-stream evaluation reads a private copy of the global model through a second
    module bound to the snapshot arena.  After a synchronous round the arena still holds that global model, so the
    inline evaluation of the arena must give the very same counters (the forward kernels are deterministic; the
    training pass of a zoo model is not bit-reproducible across runs - split-K atomics - so two federations are not
    compared with each other)."""
    from fedb200.parallel.session import GpuFederation
    ds = D.synthetic_cifar(128 * 3, 300, seed=5)
    fed = GpuFederation(model="MobileNet", rank=0, world=1, device=0, dataset=ds, seed=7, tag="t-zoo-overlap", eval_mode="overlap")
    for _ in range(3):
        tr, ev = fed.run_round(upload=True, sync=True)
        inline = fed.engine.eval_pass(upload=True, sync=True)
        assert tr.steps == 3 and ev.samples == inline.samples == 300
        assert ev.loss_sum == inline.loss_sum and ev.correct == inline.correct
    assert torch.equal(fed.engine.snap.f32, fed.trainer.arena.f32) and torch.equal(fed.engine.snap.i64, fed.trainer.arena.i64)
    fed.close()


def test_profile_records_device_phases():
    """profile=True: the round's phases are bracketed by CUDA events (and NVTX ranges); take_profile() returns their
    device times, the NVLink byte count of the aggregation and resets."""
    fed = _fed(5, profile=True, eval_mode="inline")
    fed.run_round(upload=True, sync=True)
    fed.run_round(upload=True, sync=True)
    p = fed.take_profile()
    assert p["train_ms"] > 0 and p["agg_ms"] > 0 and p["eval_ms"] > 0 and p["repl_ms"] > 0
    assert p["rounds"] == 2 and p["nvlink_bytes"] == 0 and p["arena_bytes"] == fed.group.man.n_f32 * 4   # world = 1: nothing crosses
    assert 0.0 <= p["repl_hidden_fraction"] <= 1.0
    fed.run_round(upload=False, sync=True)
    p2 = fed.take_profile()
    assert 0 < p2["train_ms"] < p["train_ms"]            # one round after the reset, against two before it
    plain = _fed(5)
    plain.run_round()
    assert plain.take_profile() == {}
    fed.close()
    plain.close()


def test_standalone_trainer_on_gpu_follows_the_cpu_run(tmp_path):
    """python -m fedb200.main on cuda: two epochs with the cosine schedule (the learning rate changes between epochs, so
    the captured pass is dropped and captured again) track the CPU run of the same seed."""
    from fedb200 import main as M
    ds = D.synthetic_cifar(128 * 4, 200, seed=2)
    common = ["--model", "LeNet", "--epochs", "2", "--lr", "0.05", "--t-max", "3", "--seed", "4"]
    cpu = M.run(common + ["--name", "c", "--checkpoint-root", str(tmp_path / "c")], dataset=ds)
    gpu = M.run(common + ["--name", "g", "--checkpoint-root", str(tmp_path / "g"), "--device", "cuda:0"], dataset=ds)
    assert gpu["last"]["steps"] == cpu["last"]["steps"] == 4 and gpu["last"]["lr"] == cpu["last"]["lr"]
    assert abs(gpu["last"]["train_loss"] - cpu["last"]["train_loss"]) < 2e-3
    assert abs(gpu["last"]["test_loss"] - cpu["last"]["test_loss"]) < 2e-3
