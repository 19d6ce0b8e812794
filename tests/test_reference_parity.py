"""End-to-end parity with the UNMODIFIED reference (SURVEY §7.4 last bullet, BASELINE config 1):
primary + backup + 2 gRPC CPU workers, LeNet, synthetic data, identical seeds, R rounds - the final
global weights of fedb200 must equal the reference's FedAvg result to fp32 tolerance.  Also checks
protocol compatibility in the mixed direction: the reference coordinator drives fedb200 workers."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baseline.ref_run import reference_available, run_reference  # noqa: E402
from fedb200.control.coordinator import Coordinator, CoordinatorConfig  # noqa: E402
from fedb200.utils import checkpoint as ckpt  # noqa: E402
from fedutil import WorkerProc, tiny_dataset  # noqa: E402

pytestmark = [pytest.mark.slow,
              pytest.mark.skipif(reference_available() is not None, reason="reference not installed in baseline/_ref")]

ROUNDS = 2


@pytest.fixture(scope="module")
def reference_result(tmp_path_factory):
    work = tmp_path_factory.mktemp("refrun")
    return run_reference(2, rounds=ROUNDS, warmup=0, model="LeNet", n_train=1024, n_test=200, data_seed=0,
                         keep_dir=str(work), worker_seeds=[11, 12], deterministic_data=True, timeout_s=600)


def test_final_global_weights_match_reference(reference_result, tmp_path):
    ref_net = ckpt.load_checkpoint(reference_result["global_ckpt"])["net"]
    ds = tiny_dataset(1024, 200, seed=0)
    ws = [WorkerProc(tmp_path, seed=s, dataset=ds) for s in (11, 12)]
    c = Coordinator(CoordinatorConfig(workers=tuple(w.address for w in ws), model="LeNet", rounds=ROUNDS,
                                      state_dir=str(tmp_path / "Primary"), membership_period_s=0.2))
    try:
        assert c.run() == ROUNDS
    finally:
        c.close()
        for w in ws:
            w.kill()
    assert list(ref_net.keys()) == list(c.global_state.keys())
    for k, v in ref_net.items():
        torch.testing.assert_close(c.global_state[k].float(), v.float(), rtol=1e-4, atol=1e-6, msg=k)


def test_reference_rounds_were_timed(reference_result):
    assert reference_result["rounds"] == ROUNDS and reference_result["rounds_per_sec"] > 0
    assert len(reference_result["round_seconds"]) == ROUNDS
