"""Control-plane logic of the GPU ("peer") plane on the CPU: the coordinator and the worker servicers are the real ones,
the NVLink endpoint is a fake that records what the RPCs ask of it and can fail on cue.

What the reference does with checkpoint payloads (src/server.py:120-153) becomes, on the peer plane, a conversation of
short descriptors: StartTrain -> "upload" reply; SendModel("global") -> wait for the in-flight aggregation + evaluate;
after a failed aggregation SendModel("reset") puts the survivors on a fresh flag epoch and the SAME round number runs
again; a re-admitted worker is "remap"-ped by the others and "pull"-s the global model; every k rounds the primary's rank
"dump"-s a reference-format checkpoint.
"""
import threading

import pytest
import time

from fedb200.control.coordinator import Coordinator, CoordinatorConfig
from fedb200.control.trainer import TrainStats
from fedb200.ops._lib import FedbError
from fedutil import WorkerProc, tiny_dataset


class FakePeer:
    """Stands in for fedb200.parallel.session.PeerWorker."""

    def __init__(self, name, fail_finish_on_call=None):
        self.name = name
        self.calls = []
        self.lock = threading.Lock()
        self.fail_finish_on_call = fail_finish_on_call
        self.finishes = 0

    def _log(self, *what):
        with self.lock:
            self.calls.append(what)

    def start_round(self, rank, world, live=None, round_id=None):
        self._log("start", rank, world, tuple(live) if live is not None else None)
        self.round_ids = getattr(self, "round_ids", []) + [round_id]
        st = TrainStats()
        st.loss_sum, st.correct, st.samples, st.steps = 2.3, 13, 128, 1
        return st

    def finish_round(self, evaluate=True):
        self.finishes += 1
        self._log("finish", self.finishes)
        if self.fail_finish_on_call == self.finishes:
            raise FedbError("K-AGG flag wait failed: status 1 (a peer never arrived)")
        ev = TrainStats()
        ev.loss_sum, ev.correct, ev.samples, ev.steps = 2.0, 20, 100, 1
        return ev if evaluate else None

    def reset(self, epoch):
        self._log("reset", epoch)

    def remap(self, r):
        self._log("remap", r)

    def pull_global(self, src=None):
        self._log("pull", src)

    def dump(self, path, what="global"):
        self._log("dump", path, what)
        open(path, "wb").write(b"stub")

    def status(self):
        return {"rank": 0, "primary_rank": 0, "backup_rank": 1, "replica_version": 0, "multicast": False}


def _peer_workers(tmp_path, n, **fake_kw):
    ds = tiny_dataset()
    ws = []
    for i in range(n):
        w = WorkerProc(tmp_path, seed=i, dataset=ds, plane="peer")
        w.servicer.peer = FakePeer(f"rank{i}", **(fake_kw if i == 0 else {}))
        ws.append(w)
    return ws


def _coord(tmp_path, ws, **kw):
    cfg = CoordinatorConfig(workers=tuple(w.address for w in ws), model="LeNet", rounds=kw.pop("rounds", 3),
                            state_dir=str(tmp_path / "Primary"), membership_period_s=0.2, heartbeat_timeout_s=1.0,
                            plane="peer", **kw)
    return Coordinator(cfg)


def _kinds(peer):
    return [c[0] for c in peer.calls]


def test_dump_descriptor_is_confined_to_the_dump_root_and_status_is_recorded(tmp_path):
    import grpc
    from fedb200 import proto
    from fedb200.control import descriptor as desc
    (w,) = _peer_workers(tmp_path, 1)
    try:
        stub = proto.TrainerStub(proto.make_channel(w.address, False))
        ok = stub.SendModel(proto.SendModelRequest(model=desc.encode("dump", path=str(tmp_path / "sub" / ".." / "g.pth"), what="client")))
        assert ok.reply == "success" and ("dump", str(tmp_path / "g.pth"), "client") in w.servicer.peer.calls
        with pytest.raises(grpc.RpcError) as e:
            stub.SendModel(proto.SendModelRequest(model=desc.encode("dump", path="/tmp/../etc/fedb200-escape.pth")))
        assert e.value.code() == grpc.StatusCode.PERMISSION_DENIED
        assert stub.SendModel(proto.SendModelRequest(model=desc.encode("status"))).reply == "success"
        rec = [r for r in w.sink.records if r["event"] == "peer_status"]
        assert rec and rec[-1]["primary_rank"] == 0 and rec[-1]["multicast"] is False
    finally:
        w.kill()


def test_descriptor_rounds_and_periodic_dump(tmp_path):
    ws = _peer_workers(tmp_path, 2)
    c = _coord(tmp_path, ws, rounds=4, checkpoint_every=2)
    try:
        assert c.run() == 4
        for r, w in enumerate(ws):
            p = w.servicer.peer
            starts = [x for x in p.calls if x[0] == "start"]
            assert [s[1:3] for s in starts] == [(r, 2)] * 4            # dense rank among live workers, world = total (N9)
            assert all(s[3] == (0, 1) for s in starts)                  # live set travels in metadata
            assert _kinds(p).count("finish") == 4 and "reset" not in _kinds(p)
        dumps = [x for x in ws[0].servicer.peer.calls if x[0] == "dump"]
        assert len(dumps) == 2 and dumps[0][1].endswith("optimizedModel.pth")   # rounds 2 and 4, on the first target only
        assert not any(x[0] == "dump" for x in ws[1].servicer.peer.calls)
    finally:
        c.close()
        for w in ws:
            w.kill()


def test_failed_aggregation_is_not_committed_and_survivors_are_reset(tmp_path):
    ws = _peer_workers(tmp_path, 2, fail_finish_on_call=2)     # rank 0's second aggregation times out in its flag wait
    c = _coord(tmp_path, ws, rounds=3)
    try:
        assert c.run() == 3                                     # three COMMITTED rounds: the failed one ran again
        for w in ws:
            k = _kinds(w.servicer.peer)
            assert k.count("start") == 4 and k.count("finish") == 4 and k.count("reset") == 1
            i = k.index("reset")
            assert k[i - 1] == "finish" and k[i + 1] == "start"  # reset sits between the failed round and its retry
        e0 = [x[1] for x in ws[0].servicer.peer.calls if x[0] == "reset"]
        e1 = [x[1] for x in ws[1].servicer.peer.calls if x[0] == "reset"]
        assert e0 == e1 and e0[0] > 1024                        # one coordinator-chosen epoch, above anything used so far
    finally:
        c.close()
        for w in ws:
            w.kill()


def test_readmitted_worker_is_remapped_and_pulls_the_global_model(tmp_path):
    ws = _peer_workers(tmp_path, 2)
    c = _coord(tmp_path, ws, rounds=2)
    try:
        assert c.run() == 2
        ws[1].kill()
        assert c.run(5) >= 2                                    # rank 1 is found dead, rounds go on with rank 0 alone
        assert any(x[0] == "start" and x[3] == (0,) for x in ws[0].servicer.peer.calls)
        ws[1].restart()
        ws[1].servicer.peer = FakePeer("rank1-restarted")
        deadline = time.time() + 10.0                           # the membership thread (0.2 s period) re-admits the worker
        while not all(c.members.snapshot().values()) and time.time() < deadline:
            time.sleep(0.05)
        assert all(c.members.snapshot().values())
        c.run(9)
        assert ("remap", 1) in ws[0].servicer.peer.calls        # the survivor maps the newcomer's fresh allocation
        assert "pull" in _kinds(ws[1].servicer.peer)            # the newcomer copies the global model out of peer memory
        last = [x for x in ws[0].servicer.peer.calls if x[0] == "start"][-1]
        assert last[3] == (0, 1)
    finally:
        c.close()
        for w in ws:
            w.kill()
