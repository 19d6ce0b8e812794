"""Mixed-version interoperability with the UNMODIFIED reference (baseline/_ref is byte-identical to
/root/reference/src) and BASELINE config 1 to the letter.  All on the CPU, all over real gRPC on localhost.

(a) the reference's coordinator (src/server.py:113-153, through baseline/ref_launcher.py) drives two ``fedb200.client``
    workers on the legacy plane; the model it writes equals the CPU oracle of the uploads it received;
(b) ``fedb200``'s coordinator drives two reference ``client.py`` workers (src/client.py:15-35);
(c) config 1: LeNet, 28x28 data padded to 32x32, primary AND backup, 3 rounds - the reference uninterrupted, fedb200 with
    the primary lost after round 1 (backup takes over) and restarted (role handed back): same final weights to fp32
    tolerance, because a clean failover must not change the arithmetic.
"""
import os
import signal
import subprocess
import sys
import threading
import time
from dataclasses import replace

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baseline.ref_run import LAUNCHER, _free_ports, _wait_port, reference_available, run_reference  # noqa: E402
from fedb200.control.coordinator import Coordinator, CoordinatorConfig  # noqa: E402
from fedb200.control.ha import BackupConfig, BackupNode, run_primary  # noqa: E402
from fedb200.parallel.fedavg import fedavg_state_dicts  # noqa: E402
from fedb200.utils import checkpoint as ckpt  # noqa: E402
from fedb200.utils import data as D  # noqa: E402
from fedutil import WorkerProc  # noqa: E402

pytestmark = [pytest.mark.slow,
              pytest.mark.skipif(reference_available() is not None, reason="reference not installed in baseline/_ref")]


def _env():
    env = dict(os.environ, PYTHONPATH=ROOT, CUDA_VISIBLE_DEVICES="", PYTHONUNBUFFERED="1", OMP_NUM_THREADS="2")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    return env


def _kill(procs):
    for p in procs:
        if p.poll() is None:
            try:
                os.killpg(p.pid, signal.SIGKILL)          # exactly the process groups started here
            except ProcessLookupError:
                pass
    for p in procs:
        try:
            p.wait(timeout=10)
        except Exception:   # noqa: BLE001
            pass


def _wait(cond, timeout=120.0, period=0.05):
    end = time.time() + timeout
    while time.time() < end:
        if cond():
            return True
        time.sleep(period)
    return False


def test_reference_coordinator_drives_fedb200_workers(tmp_path):
    """(a) reference primary -> fedb200 workers (legacy plane), fedb200 backup receives pings and the replica."""
    rounds = 2
    wports = _free_ports(2)
    (bport,) = _free_ports(1)
    work = tmp_path / "ref"
    work.mkdir()
    log = open(tmp_path / "procs.log", "w")
    procs = []
    try:
        for i, port in enumerate(wports):
            procs.append(subprocess.Popen(
                [sys.executable, "-m", "fedb200.client", "-a", f"localhost:{port}", "--model", "LeNet", "--device", "cpu",
                 "--dataset-size", "512", "--test-size", "100", "--seed", str(21 + i)],
                cwd=str(tmp_path), env=_env(), stdout=log, stderr=subprocess.STDOUT, start_new_session=True))
        assert all(_wait_port(p, 120) for p in wports), "fedb200 workers did not come up"
        ccfg = CoordinatorConfig(workers=tuple(f"localhost:{p}" for p in wports), model="LeNet", rounds=rounds,
                                 state_dir=str(tmp_path / "BackupActing"))
        backup = BackupNode(BackupConfig(port=str(bport), watchdog_s=3600, state_dir=str(tmp_path / "Backup"), coordinator=ccfg))
        backup.start()
        timing = str(work / "timing.json")
        prim = subprocess.Popen(
            [sys.executable, LAUNCHER, "primary", "--workdir", str(work), "--model", "LeNet", "--backup-port", str(bport),
             "--workers", ",".join(f"localhost:{p}" for p in wports), "--rounds", str(rounds), "--timing-out", timing],
            env=_env(), stdout=log, stderr=subprocess.STDOUT, start_new_session=True)
        procs.append(prim)
        assert prim.wait(timeout=600) == 0, open(tmp_path / "procs.log").read()[-3000:]
        # what the reference received in its last round (src/server.py:66-71 writes test_<rank>.pth) and what it made of it
        ups = [ckpt.load_checkpoint(str(work / "Primary" / f"test_{r}.pth"))["net"] for r in range(2)]
        glob = ckpt.load_checkpoint(str(work / "Primary" / "optimizedModel.pth"))["net"]
        want = fedavg_state_dicts(ups)
        assert list(glob.keys()) == list(want.keys())
        for k in want:
            assert torch.equal(glob[k], want[k]), k
        # the two uploads are different models (different seeds, different shards), so the mean is not trivial
        assert not torch.equal(ups[0]["conv1.weight"], ups[1]["conv1.weight"])
        # the reference's backup push (src/server.py:141-142) landed in a fedb200 backup
        assert backup.ping_seen or backup.last_ping > 0
        rep = ckpt.load_checkpoint(str(tmp_path / "Backup" / "optimizedModel.pth"))["net"]
        assert all(torch.equal(rep[k], glob[k]) for k in glob)
        backup.stop()
    finally:
        _kill(procs)
        log.close()


def test_fedb200_coordinator_drives_reference_workers(tmp_path):
    """(b) fedb200 primary -> the reference's client.py workers."""
    rounds = 2
    wports = _free_ports(2)
    log = open(tmp_path / "procs.log", "w")
    procs = []
    try:
        for i, port in enumerate(wports):
            wdir = tmp_path / f"w{i}"
            procs.append(subprocess.Popen(
                [sys.executable, LAUNCHER, "worker", "--workdir", str(wdir), "--model", "LeNet", "--n-train", "512",
                 "--n-test", "100", "--address", f"localhost:{port}", "--torch-seed", str(31 + i), "--deterministic-data"],
                env=_env(), stdout=log, stderr=subprocess.STDOUT, start_new_session=True))
        assert all(_wait_port(p, 300) for p in wports), "reference workers did not come up"
        c = Coordinator(CoordinatorConfig(workers=tuple(f"localhost:{p}" for p in wports), model="LeNet", rounds=rounds,
                                          state_dir=str(tmp_path / "Primary"), membership_period_s=0.5,
                                          train_timeout_s=300, send_timeout_s=300))
        seen = []
        real = c.aggregate
        c.aggregate = lambda uploads: (seen.append([u["state"] for u in uploads]), real(uploads))[1]
        try:
            assert c.run() == rounds
        finally:
            c.close()
        assert len(seen) == rounds and all(len(s) == 2 for s in seen)
        want = fedavg_state_dicts(seen[-1])
        for k, v in want.items():
            assert torch.equal(c.global_state[k], v), k
        # the reference workers hold the broadcast model (src/client.py:25-31 writes ./checkpoint/<address>.pth)
        for i, port in enumerate(wports):
            got = ckpt.load_checkpoint(str(tmp_path / f"w{i}" / "checkpoint" / f"localhost:{port}.pth"))["net"]
            assert all(torch.equal(got[k], c.global_state[k]) for k in want)
    finally:
        _kill(procs)
        log.close()


def test_config1_literal_with_primary_loss_matches_reference(tmp_path):
    """(c) BASELINE config 1."""
    rounds, seeds = 3, [41, 42]
    ref = run_reference(2, rounds=rounds, warmup=0, model="LeNet", n_train=768, n_test=100, data_seed=3, pad_from=28,
                        keep_dir=str(tmp_path / "refrun"), worker_seeds=seeds, deterministic_data=True, timeout_s=900)
    ref_net = ckpt.load_checkpoint(ref["global_ckpt"])["net"]

    ds = D.synthetic_cifar(768, 100, seed=3, pad_from=28)
    assert int(ds.train_x[:, :2].sum()) == 0 and int(ds.train_x[:, 2:30, 2:30].sum()) > 0      # 28x28 inside a zero frame
    ws = [WorkerProc(tmp_path, seed=s, dataset=ds) for s in seeds]
    (bport,) = _free_ports(1)
    ccfg = CoordinatorConfig(workers=tuple(w.address for w in ws), model="LeNet", rounds=rounds,
                             state_dir=str(tmp_path / "Primary"), backup_target=f"127.0.0.1:{bport}", replicate="sync",
                             membership_period_s=0.2)
    backup = BackupNode(BackupConfig(port=str(bport), watchdog_s=0.6, state_dir=str(tmp_path / "Backup"), coordinator=ccfg))
    backup.start()
    try:
        # the primary commits round 0 and is lost (its process ends between rounds)
        first = run_primary(replace(ccfg, rounds=1), ping_period_s=0.1)
        assert first.round == 1
        # the backup notices the silence, takes over at round 1 with a higher term and finishes the job
        assert _wait(lambda: backup.role == "ACTING" and backup.acting is not None, timeout=20)
        acting = backup.acting
        assert acting.term > first.term
        assert _wait(lambda: acting.round >= rounds, timeout=120)
        final = {k: v.clone() for k, v in acting.global_state.items()}
        # the primary is restarted: the backup hands the role back; nothing is left to run, nothing changes
        holder = {}
        t = threading.Thread(target=lambda: run_primary(ccfg, ping_period_s=0.1, on_coordinator=lambda c: holder.setdefault("c", c)),
                             daemon=True)
        t.start()
        assert _wait(lambda: "c" in holder and backup.role == "BACKUP", timeout=30)
        new = holder["c"]
        assert new.round >= rounds and new.term > acting.term
        new.stop()
        t.join(timeout=30)
    finally:
        backup.stop()
        for w in ws:
            w.kill()
    assert list(final.keys()) == list(ref_net.keys())
    for k, v in ref_net.items():
        torch.testing.assert_close(final[k].float(), v.float(), rtol=1e-4, atol=1e-6, msg=k)
