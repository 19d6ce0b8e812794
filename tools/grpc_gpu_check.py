"""gRPC control plane + GPU data plane, end to end on one box with >= 2 GPUs (FEDB200_CHECK_WORLD=8 for eight):

  coordinator (this process, no GPU context) --StartTrain/SendModel descriptors--> 2 worker processes
  (``python -m fedb200.client --plane peer``), one per GPU, whose arenas are peer-mapped; FedAvg, broadcast and
  the backup replica run as K-AGG/K-REPL kernels over NVLink.  Then: kill a worker mid-run (bounded flag wait ->
  ABORTED -> reset -> live-only rounds), restart it with --rejoin (survivor re-maps its memory, newcomer pulls the
  global model), and check the rounds continue with both.  No rank is special: the worker that HOSTS the global model
  (rank 0) is killed too - the lowest live rank takes the global model and the next one the replica, the replica's version
  follows the committed rounds, a restarted rank 0 pulls from the new holder - and, with >= 4 GPUs, ranks 0 and 1 die in
  consecutive rounds.  Finally the primary PROCESS is killed mid-round and the backup takes over.
"""
import json
import os
import signal
import socket
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fedb200.control.coordinator import Coordinator, CoordinatorConfig  # noqa: E402
from fedb200.control.ha import BackupConfig, BackupNode  # noqa: E402
from fedb200.utils import checkpoint as ckpt  # noqa: E402


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def wait_port(port, timeout=240):
    end = time.time() + timeout
    while time.time() < end:
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.5).close()
            return True
        except OSError:
            time.sleep(0.3)
    return False


WORLD = int(os.environ.get("FEDB200_CHECK_WORLD", "2"))     # workers = GPUs (config 4 of BASELINE.json uses 8)


def spawn_worker(rank, port, tag, work, rejoin=False):
    cmd = [sys.executable, "-m", "fedb200.client", "-a", f"127.0.0.1:{port}", "--plane", "peer", "--model", "LeNet",
           "--device", f"cuda:{rank}", "--peer-rank", str(rank), "--peer-world", str(WORLD), "--peer-tag", tag,
           "--dataset-size", "2048", "--test-size", "500", "--seed", str(10 + rank), "--agg-timeout", "4",
           "--dump-root", work, "--metrics", os.path.join(work, f"worker{rank}.jsonl"), "--profile"]
    if rejoin:
        cmd.append("--rejoin")
    log = open(os.path.join(work, f"worker{rank}{'-re' if rejoin else ''}.log"), "w")
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.Popen(cmd, cwd=work, stdout=log, stderr=subprocess.STDOUT, env=env, start_new_session=True)


def statuses(c, work, ranks):
    """'status' descriptor to the given workers -> their newest peer_status records (roles, replica version, checksum)."""
    from fedb200.control import descriptor as desc
    out = {}
    for r in ranks:
        assert c._send_model_to(c.cfg.workers[r], desc.encode("status")), f"status RPC to worker {r} failed"
        recs = [json.loads(line) for line in open(os.path.join(work, f"worker{r}.jsonl")) if '"peer_status"' in line]
        out[r] = recs[-1]
    return out


def check_roles(c, work, live, label):
    """After committed rounds with the live set ``live``: every survivor agrees on the holders (lowest live ranks), the
    replica on the backup rank carries the coordinator's committed-round count, all client arenas hold the same bits,
    and they are the bits of the global model the holder dumps."""
    import torch
    from fedb200.control import descriptor as desc
    st = statuses(c, work, live)
    prim = live[0]
    bak = live[1] if len(live) > 1 else live[0]
    assert all(s["primary_rank"] == prim and s["backup_rank"] == bak for s in st.values()), (label, st)
    assert len({s["arena_checksum"] for s in st.values()}) == 1, (label, st)
    if bak != prim:
        assert st[bak]["replica_version"] == c.round, (label, st[bak]["replica_version"], c.round)
    gpath = os.path.join(work, f"global-{label}.pth")
    assert c._send_model_to(c.cfg.workers[prim], desc.encode("dump", path=gpath, what="global"))
    glob = ckpt.load_checkpoint(gpath)["net"]
    for r in live:
        cpath = os.path.join(work, f"client{r}-{label}.pth")
        assert c._send_model_to(c.cfg.workers[r], desc.encode("dump", path=cpath, what="client"))
        mine = ckpt.load_checkpoint(cpath)["net"]
        assert all(torch.equal(mine[k], glob[k]) for k in glob), (label, r)
    return st


def rejoin(c, procs, ports, tag, work, r):
    procs[r] = spawn_worker(r, ports[r], tag, work, rejoin=True)
    assert wait_port(ports[r])
    end = time.time() + 60
    while time.time() < end and not c.members.snapshot()[c.cfg.workers[r]]:
        time.sleep(0.2)
    assert c.members.snapshot()[c.cfg.workers[r]], f"worker {r} was not re-admitted"


def main():
    work = tempfile.mkdtemp(prefix="fedb200-grpcgpu-")
    tag = f"g{os.getpid()}"
    ports = [free_port() for _ in range(WORLD)]
    procs = [spawn_worker(r, ports[r], tag, work) for r in range(WORLD)]
    try:
        assert all(wait_port(p) for p in ports), "workers did not come up"
        bport = free_port()
        ccfg = CoordinatorConfig(workers=tuple(f"127.0.0.1:{p}" for p in ports), model="LeNet", rounds=3,
                                 state_dir=os.path.join(work, "Primary"), backup_target=f"127.0.0.1:{bport}",
                                 plane="peer", checkpoint_every=1, membership_period_s=0.5, train_timeout_s=120,
                                 send_timeout_s=120)
        backup = BackupNode(BackupConfig(port=str(bport), watchdog_s=3600, state_dir=os.path.join(work, "Backup"),
                                         coordinator=ccfg))
        backup.start()
        c = Coordinator(ccfg, term=1)
        assert c.run() == 3
        rounds = [r for r in c.sink.records if r["event"] == "round"]
        assert all(r["uploads"] == WORLD and r["plane"] == "peer" for r in rounds), rounds
        assert all(r["bytes_up"] == 0 for r in rounds)                       # no weights in the RPC payloads
        glob = ckpt.load_checkpoint(os.path.join(work, "Primary", "optimizedModel.pth"))
        assert list(glob["net"].keys())[0] == "conv1.weight" and glob["acc"] == 1
        assert backup.next_round == 3                                          # replica metadata followed
        st = check_roles(c, work, list(range(WORLD)), "p1")
        prof = [json.loads(line) for line in open(os.path.join(work, "worker0.jsonl")) if '"round_profile"' in line]
        assert prof and all(k in prof[-1] for k in ("train_ms", "agg_ms", "eval_ms", "nvlink_bytes", "agg_gbps")), prof[-1:]
        print("device-timed phases of the last round on worker 0:",
              {k: prof[-1][k] for k in ("train_ms", "agg_ms", "repl_ms", "eval_ms", "nvlink_bytes", "agg_gbps", "repl_hidden_fraction")
               if k in prof[-1]})
        print("phase 1 ok: 3 descriptor rounds,", [round(r["total_s"], 3) for r in rounds],
              f"multicast: {str(st[0]['multicast']).lower()}, replica version {st[1 % WORLD]['replica_version']}")

        # ---- kill worker 1: the survivor's flag wait is bounded, the round is dropped, the next ones are live-only
        os.killpg(procs[1].pid, signal.SIGKILL)
        procs[1].wait()
        c.cfg.rounds = 6
        done = c.run()
        rounds = [r for r in c.sink.records if r["event"] == "round"]
        assert done == 3 and rounds[-1]["uploads"] == WORLD - 1, (done, rounds[-3:])
        assert c.members.snapshot()[ccfg.workers[1]] is False
        print("phase 2 ok: worker 1 killed, rounds continue with the survivor:", [r["uploads"] for r in rounds])

        # ---- restart it with --rejoin: membership re-admits it (remap + pull), rounds use both again
        procs[1] = spawn_worker(1, ports[1], tag, work, rejoin=True)
        assert wait_port(ports[1])
        end = time.time() + 60
        while time.time() < end and not c.members.snapshot()[ccfg.workers[1]]:
            time.sleep(0.2)
        assert c.members.snapshot()[ccfg.workers[1]], "worker 1 was not re-admitted"
        c.cfg.rounds = c.round + 2
        done = c.run()
        rounds = [r for r in c.sink.records if r["event"] == "round"]
        assert done == 2 and rounds[-1]["uploads"] == WORLD, (done, rounds[-2:])
        print("phase 3 ok: worker 1 re-admitted, uploads per round:", [r["uploads"] for r in rounds])
        check_roles(c, work, list(range(WORLD)), "p3")

        # ---- phase 3b: the worker that HOSTS the global model dies.  The round in flight is dropped (nobody's arena is
        # touched by the failed K-AGG), the survivors re-run it live-only, rank 1 now keeps the global model, the next
        # live rank the replica; a restarted rank 0 pulls from rank 1.
        os.killpg(procs[0].pid, signal.SIGKILL)
        procs[0].wait()
        c.cfg.rounds = c.round + 3
        done = c.run()
        rounds = [r for r in c.sink.records if r["event"] == "round"]
        assert done == 3 and rounds[-1]["uploads"] == WORLD - 1, (done, rounds[-3:])
        survivors = list(range(1, WORLD))
        st = check_roles(c, work, survivors, "p3b")
        print(f"phase 3b ok: worker 0 (global model host) killed; holder is now rank {st[1]['primary_rank']}, replica on rank "
              f"{st[1]['backup_rank']} at version {st[st[1]['backup_rank']]['replica_version']} == committed rounds {c.round}")
        rejoin(c, procs, ports, tag, work, 0)
        c.cfg.rounds = c.round + 2
        assert c.run() == 2
        rounds = [r for r in c.sink.records if r["event"] == "round"]
        assert rounds[-1]["uploads"] == WORLD, rounds[-2:]
        check_roles(c, work, list(range(WORLD)), "p3c")
        print("phase 3c ok: worker 0 re-admitted from the new holder; all arenas equal, roles back on ranks 0/1")

        if WORLD >= 4:
            # ---- phase 3d: ranks 0 and 1 die in consecutive rounds
            os.killpg(procs[0].pid, signal.SIGKILL)
            procs[0].wait()
            c.cfg.rounds = c.round + 1
            assert c.run() == 1
            os.killpg(procs[1].pid, signal.SIGKILL)
            procs[1].wait()
            c.cfg.rounds = c.round + 2
            assert c.run() == 2
            st = check_roles(c, work, list(range(2, WORLD)), "p3d")
            print(f"phase 3d ok: workers 0 and 1 killed in consecutive rounds; holder rank {st[2]['primary_rank']}, replica rank "
                  f"{st[2]['backup_rank']} at version {st[st[2]['backup_rank']]['replica_version']} == {c.round}")
            rejoin(c, procs, ports, tag, work, 0)
            rejoin(c, procs, ports, tag, work, 1)
            c.cfg.rounds = c.round + 2
            assert c.run() == 2
            check_roles(c, work, list(range(WORLD)), "p3e")
            print("phase 3e ok: both re-admitted")
        last_round = c.round
        c.close()
        backup.stop()

        # ---- phase 4: primary failover on the GPU plane.  A primary PROCESS is killed by fault injection right
        # after the workers trained (K-AGG launches in flight); the backup's watchdog promotes it, it resets the
        # flag protocol, bumps the term and the rounds go on from the replicated round number.
        bport2 = free_port()
        pdir = os.path.join(work, "Primary2")
        ccfg2 = CoordinatorConfig(workers=ccfg.workers, model="LeNet", rounds=10_000, state_dir=os.path.join(work, "Backup2"),
                                  plane="peer", membership_period_s=0.5, train_timeout_s=120, send_timeout_s=120)
        env = dict(os.environ, PYTHONPATH=ROOT, FEDB200_FAULTS="kill:after_train:800")
        prim = subprocess.Popen([sys.executable, "-m", "fedb200.server", "--p", "y", "--plane", "peer", "--model", "LeNet",
                                 "--workers", ",".join(ccfg.workers), "--rounds", "100000", "--backupPort", str(bport2),
                                 "--backupAddress", "127.0.0.1", "--state-dir", pdir, "--ping-period", "0.2"],
                                cwd=work, env=env, stdout=open(os.path.join(work, "primary2.log"), "w"),
                                stderr=subprocess.STDOUT, start_new_session=True)
        procs.append(prim)
        # the backup (watchdog 1.5 s) starts once the primary is up and running rounds, like an operator would
        end = time.time() + 120
        while time.time() < end:
            try:
                if json.load(open(os.path.join(pdir, "coordinator.json")))["next_round"] >= 1:
                    break
            except (OSError, ValueError, KeyError):
                pass
            time.sleep(0.05)
        backup2 = BackupNode(BackupConfig(port=str(bport2), watchdog_s=1.5, state_dir=os.path.join(work, "Backup2"),
                                          coordinator=ccfg2))
        backup2.start()
        prim.wait(timeout=120)
        if prim.returncode != 137:
            print(open(os.path.join(work, "primary2.log")).read()[-3000:])
        assert prim.returncode == 137, f"primary should have been killed by the injected fault, rc={prim.returncode}"
        t_kill = time.time()
        end = time.time() + 30
        while time.time() < end and backup2.role != "ACTING":
            time.sleep(0.05)
        assert backup2.role == "ACTING", "backup did not take over"
        t_take = time.time() - t_kill
        acting = backup2.acting
        start = acting.round
        end = time.time() + 60
        while time.time() < end and acting.round < start + 3:
            time.sleep(0.1)
        assert acting.round >= start + 3, f"acting coordinator stuck at round {acting.round}"
        assert acting.term >= 2 and start >= 100, (acting.term, start)
        print(f"phase 4 ok: primary killed mid-round, backup took over {t_take:.2f}s later at round {start}, "
              f"term {acting.term}, now at round {acting.round}")
        backup2.stop()
        print("GRPC_GPU_CHECK_OK")
    finally:
        for p in procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass
        for name in sorted(os.listdir(work)):
            if name.endswith(".log"):
                tail = open(os.path.join(work, name)).read()[-1500:]
                if "Traceback" in tail or "Error" in tail:
                    print(f"---- {name}\n{tail}")


if __name__ == "__main__":
    main()
