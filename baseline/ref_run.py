"""Process orchestration for the reference arm: primary + backup + N workers of the unmodified
reference on localhost (its own "multi-node" story, src/server.py:281-282), timed per round."""
from __future__ import annotations

import json
import os
import shutil
import signal
import socket
import subprocess
import sys
import tempfile
import time
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
LAUNCHER = os.path.join(HERE, "ref_launcher.py")


def reference_available() -> Optional[str]:
    src = os.path.join(HERE, "_ref", "fedref_src", "server.py")
    return None if os.path.exists(src) else f"{src} missing (reference not installed into baseline/_ref)"


def _free_ports(n):
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def _wait_port(port, timeout):
    end = time.time() + timeout
    while time.time() < end:
        try:
            with socket.create_connection(("127.0.0.1", port), timeout=0.5):
                return True
        except OSError:
            time.sleep(0.25)
    return False


def run_reference(n_workers: int, rounds: int, warmup: int, model: str = "LeNet", n_train: int = 50000,
                  n_test: int = 10000, data_seed: int = 0, pad_from: Optional[int] = None, compress: bool = False,
                  gpu_workers: bool = False, timeout_s: float = 3600.0, keep_dir: Optional[str] = None,
                  worker_seeds=None, log=None, deterministic_data: bool = False) -> dict:
    """Runs ``warmup + rounds`` reference rounds; returns timing of the last ``rounds`` ones plus the
    path of the final global checkpoint (when ``keep_dir`` is given)."""
    total = warmup + rounds
    work = keep_dir or tempfile.mkdtemp(prefix="fedref-")
    os.makedirs(work, exist_ok=True)
    ports = _free_ports(n_workers + 1)
    wports, bport = ports[:n_workers], ports[-1]
    workers = ",".join(f"localhost:{p}" for p in wports)
    cores = os.cpu_count() or 8
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = str(max(1, cores // max(n_workers, 1)))
    env["PYTHONUNBUFFERED"] = "1"
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    common = ["--model", model, "--n-train", str(n_train), "--n-test", str(n_test), "--data-seed", str(data_seed)]
    if pad_from:
        common += ["--pad-from", str(pad_from)]
    if compress:
        common += ["--compress", "Y"]
    if deterministic_data:
        common += ["--deterministic-data"]
    procs = []
    logf = open(os.path.join(work, "reference.log"), "w") if log is None else log

    def spawn(role, extra, wenv):
        p = subprocess.Popen([sys.executable, LAUNCHER, role, "--workdir", work] + common + extra,
                             stdout=logf, stderr=subprocess.STDOUT, env=wenv, start_new_session=True)
        procs.append(p)
        return p

    try:
        for i, port in enumerate(wports):
            wenv = dict(env)
            if gpu_workers:
                wenv["CUDA_VISIBLE_DEVICES"] = str(i)
            else:
                wenv["CUDA_VISIBLE_DEVICES"] = ""      # strict reference: CPU workers
            extra = ["--address", f"localhost:{port}"] + (["--dp-shim"] if gpu_workers else [])
            if worker_seeds is not None:
                extra += ["--torch-seed", str(worker_seeds[i])]
            spawn("worker", extra, wenv)
        cenv = dict(env)
        # R-gpu: the workers' checkpoints hold CUDA tensors, which the reference's coordinator un-pickles with a plain
        # torch.load (src/server.py:158-160) - it needs a visible GPU for that, or it dies in its receive thread and hangs
        cenv["CUDA_VISIBLE_DEVICES"] = "0" if gpu_workers else ""
        for port in wports:
            if not _wait_port(port, 600):
                raise RuntimeError(f"reference worker on port {port} did not come up (see {work}/reference.log)")
        # Ordering matters twice: the backup's watchdog promotes it after 10-20 s without a ping
        # (src/server.py:254-264), so the primary is started first and is already pinging when the backup
        # comes up; and the primary must find the backup listening before round 0, because a failed backup
        # push poisons its worker table (clients[None] = False, src/server.py:75,142) - the
        # launcher holds run() until the backup port answers.
        timing = os.path.join(work, "timing.json")
        prim = spawn("primary", ["--backup-port", str(bport), "--workers", workers, "--rounds", str(total),
                                 "--timing-out", timing], cenv)
        time.sleep(1.0)     # the primary starts pinging first; it waits for the backup before round 0
        spawn("backup", ["--backup-port", str(bport), "--workers", workers, "--rounds", str(total)], cenv)
        prim.wait(timeout=timeout_s)
        if prim.returncode != 0 or not os.path.exists(timing):
            raise RuntimeError(f"reference primary failed (rc={prim.returncode}); see {work}/reference.log")
        with open(timing) as f:
            t = json.load(f)
        starts, end = t["round_starts"], t["end"]
        if len(starts) != total:
            raise RuntimeError(f"reference ran {len(starts)} rounds, expected {total}")
        seconds = end - starts[warmup]
        return {"rounds": rounds, "seconds": seconds, "rounds_per_sec": rounds / seconds,
                "round_seconds": [b - a for a, b in zip(starts, starts[1:] + [end])], "workdir": work,
                "global_ckpt": os.path.join(work, "Primary", "optimizedModel.pth")}
    finally:
        for p in procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGKILL)     # exact process groups we started
                except ProcessLookupError:
                    pass
        for p in procs:
            try:
                p.wait(timeout=10)
            except Exception:
                pass
        if log is None:
            logf.close()
        if keep_dir is None:
            shutil.rmtree(work, ignore_errors=True)
