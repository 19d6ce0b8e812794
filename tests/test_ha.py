"""Primary/backup failover (SURVEY work item 9): replication, watchdog takeover that reloads the
replica (B6), hand-back on a recovering primary without B1/B2, term monotonicity."""
import threading
import time

import torch

from fedb200 import proto
from fedb200.control import descriptor as desc
from fedb200.control.coordinator import Coordinator, CoordinatorConfig
from fedb200.control.ha import BackupConfig, BackupNode, PrimaryPinger, run_primary
from fedb200.utils import checkpoint as ckpt
from fedutil import WorkerProc, free_port, tiny_dataset


def _wait(pred, timeout=20.0, step=0.05):
    end = time.time() + timeout
    while time.time() < end:
        if pred():
            return True
        time.sleep(step)
    return False


def test_failover_and_handback(tmp_path):
    ds = tiny_dataset(512, 100)
    ws = [WorkerProc(tmp_path, seed=i, dataset=ds, eval_on_send=False) for i in range(2)]
    bport = free_port()
    ccfg = CoordinatorConfig(workers=tuple(w.address for w in ws), model="LeNet", rounds=10_000,
                             state_dir=str(tmp_path / "Primary"), backup_target=f"127.0.0.1:{bport}",
                             replicate="sync", membership_period_s=0.2)
    backup = BackupNode(BackupConfig(port=str(bport), watchdog_s=0.6, state_dir=str(tmp_path / "Backup"),
                                     coordinator=ccfg))
    backup.start()
    holder = {}
    t = threading.Thread(target=lambda: holder.setdefault(
        "c", run_primary(ccfg, ping_period_s=0.1, on_coordinator=lambda c: holder.setdefault("live", c))),
        daemon=True)
    t.start()
    try:
        assert _wait(lambda: "live" in holder and holder["live"].round >= 2)
        prim = holder["live"]
        assert backup.role == "BACKUP" and backup.next_round >= 1
        # replica == primary's committed model (C8)
        rep = ckpt.load_checkpoint(str(tmp_path / "Backup" / "optimizedModel.pth"))["net"]
        assert set(rep) == set(prim.global_state)

        # --- primary "crashes": loop stops, pings stop
        prim.stop()
        t.join(timeout=30)
        term_before, round_before = prim.term, prim.round
        assert _wait(lambda: backup.role == "ACTING", timeout=10)
        assert 0.0 < backup.detect_latency_s <= 2 * 0.6 + 0.5
        acting = backup.acting
        assert acting.term > term_before                       # term monotonic
        assert acting.round >= round_before - 1 and acting.round > 0   # B6: not from round 0
        assert _wait(lambda: acting.round >= round_before + 1)  # rounds continue on the backup
        assert all(w.servicer.term == acting.term for w in ws)

        # --- primary comes back: first ping says recovering=1 -> backup hands the role back
        holder2 = {}
        t2 = threading.Thread(target=lambda: run_primary(
            ccfg, ping_period_s=0.1, on_coordinator=lambda c: holder2.setdefault("live", c)), daemon=True)
        t2.start()
        assert _wait(lambda: "live" in holder2)
        new = holder2["live"]
        assert backup.role == "BACKUP" and backup.demotions == 1
        assert new.term > acting.term
        r0 = new.round
        assert r0 >= round_before + 1                           # continues where the backup stopped
        assert _wait(lambda: new.round >= r0 + 2)
        # deposed loop cannot drive workers any more (no split brain): all workers on the new term
        assert all(w.servicer.term == new.term for w in ws)
        assert acting.stop_event.is_set()
        new.stop()
        t2.join(timeout=30)
    finally:
        backup.stop()
        for w in ws:
            w.kill()


def test_backup_down_does_not_poison_membership(tmp_path):
    """B5: a dead backup must not add entries to the worker table or stall the rounds."""
    ds = tiny_dataset(256, 100)
    ws = [WorkerProc(tmp_path, seed=i, dataset=ds, eval_on_send=False) for i in range(2)]
    cfg = CoordinatorConfig(workers=tuple(w.address for w in ws), model="LeNet", rounds=2,
                            state_dir=str(tmp_path / "Primary"), backup_target=f"127.0.0.1:{free_port()}",
                            replicate="sync", send_timeout_s=2.0, membership_period_s=0.2)
    c = Coordinator(cfg)
    try:
        assert c.run() == 2
        assert c.backup_up is False and set(c.members.snapshot()) == {w.address for w in ws}
    finally:
        c.close()
        for w in ws:
            w.kill()


def test_ping_payload_is_reference_compatible():
    """PingRequest.req stays exactly "1" then "0"; fencing data rides in metadata."""
    seen = []

    class B(proto.TrainerServicer):
        def CheckIfPrimaryUp(self, request, context):
            seen.append((request.req, desc.read_fence(context.invocation_metadata())))
            return proto.PingResponse(value=1)

    server = proto.make_server(2)
    proto.add_TrainerServicer_to_server(B(), server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()
    try:
        p = PrimaryPinger(f"127.0.0.1:{port}", lambda: (7, 3))
        assert p.ping_once()[0] and p.ping_once()[0]
        assert [s[0] for s in seen] == ["1", "0"] and seen[0][1] == (7, 3)
        p.stop()
    finally:
        server.stop(0)
