"""Runs the UNMODIFIED reference (baseline/_ref/fedref_src, byte-identical to /root/reference/src)
headless and offline.  Harness code, not product code.

Nothing in the reference is edited.  The launcher only prepares the process it runs in:
  * ``multipledispatch`` is not installed here -> a tiny in-memory dispatch-by-arity/type shim is put in
    ``sys.modules`` before ``main.py`` imports it (src/main.py:16,104,128,167,193);
  * ``utils.py:45`` shells out to ``stty size`` -> ``os.popen`` answers "24 80" for that command;
  * there is no network -> ``torchvision.datasets.CIFAR10`` is replaced by a seeded synthetic dataset
    of the same shape that applies the reference's own transforms exactly like torchvision does;
  * ``./checkpoint`` is created (src/main.py:231-239 needs it, B7);
  * ``--model NAME`` rebinds ``models.MobileNet`` to another constructor of the reference's own zoo
    BEFORE main/server resolve the name (the reference selects models by editing source);
  * coordinator only: the worker table (src/server.py:281-282) and the round count
    (src/server.py:120) are hard-coded, so for N != 2 workers / R != 20 rounds the launcher replays the
    reference's own ``__main__`` block (src/server.py:286-293) call for call with N addresses and
    shadows ``range`` inside the server module.

roles:  worker  -> runs client.py's __main__;   primary / backup -> runs server.py.
"""
from __future__ import annotations

import argparse
import builtins
import io
import json
import os
import runpy
import sys
import threading
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.path.join(HERE, "_ref", "fedref_src")


def install_multipledispatch_shim():
    mod = types.ModuleType("multipledispatch")
    registry = {}

    def dispatch(*types_):
        def deco(fn):
            table = registry.setdefault(fn.__name__, {})
            table[tuple(types_)] = fn

            def caller(*args):
                for sig, impl in table.items():
                    if len(sig) == len(args) and all(isinstance(a, t) for a, t in zip(args, sig)):
                        return impl(*args)
                raise NotImplementedError(f"no {fn.__name__} for {[type(a) for a in args]}")
            caller.__name__ = fn.__name__
            return caller
        return deco

    mod.dispatch = dispatch
    sys.modules["multipledispatch"] = mod


def patch_stty():
    real = os.popen

    def popen(cmd, *a, **k):
        if isinstance(cmd, str) and cmd.strip().startswith("stty size"):
            return io.StringIO("24 80")
        return real(cmd, *a, **k)
    os.popen = popen


def install_synthetic_cifar(n_train, n_test, seed, pad_from, deterministic=False):
    import numpy as np
    import torch
    import torchvision
    from PIL import Image
    sys.path.insert(0, os.path.dirname(HERE))
    from fedb200.utils.data import synthetic_cifar
    ds = synthetic_cifar(n_train, n_test, seed=seed, pad_from=pad_from)

    class SyntheticCIFAR10(torch.utils.data.Dataset):
        def __init__(self, root=None, train=True, download=False, transform=None, target_transform=None):
            x, y = (ds.train_x, ds.train_y) if train else (ds.test_x, ds.test_y)
            self.data, self.targets = x.numpy(), [int(v) for v in y]
            if deterministic and transform is not None:
                # parity runs only: drop the Random* stages so both systems see the same pixels
                import torchvision.transforms as T
                transform = T.Compose([t for t in transform.transforms if not type(t).__name__.startswith("Random")])
            self.transform, self.target_transform = transform, target_transform

        def __getitem__(self, index):           # same steps as torchvision.datasets.CIFAR10.__getitem__
            img, target = Image.fromarray(self.data[index]), self.targets[index]
            if self.transform is not None:
                img = self.transform(img)
            if self.target_transform is not None:
                target = self.target_transform(target)
            return img, target

        def __len__(self):
            return len(self.data)

    torchvision.datasets.CIFAR10 = SyntheticCIFAR10


def rebind_model(name):
    if not name or name == "MobileNet":
        return
    import models                         # the reference's package (REF_SRC is first on sys.path)
    models.MobileNet = getattr(models, name)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("role", choices=("worker", "primary", "backup"))
    ap.add_argument("--address", default="localhost:50051")
    ap.add_argument("--workers", default="localhost:50051,localhost:50052")
    ap.add_argument("--rounds", type=int, default=20)
    ap.add_argument("--model", default="MobileNet")
    ap.add_argument("--compress", default=None)
    ap.add_argument("--backup-port", default="8080")
    ap.add_argument("--n-train", type=int, default=50000)
    ap.add_argument("--n-test", type=int, default=10000)
    ap.add_argument("--data-seed", type=int, default=0)
    ap.add_argument("--pad-from", type=int, default=None)
    ap.add_argument("--dp-shim", action="store_true", help="R-gpu variant: DataParallel -> identity (B4)")
    ap.add_argument("--workdir", default=None)
    ap.add_argument("--timing-out", default=None, help="primary: write per-round start times as JSON")
    ap.add_argument("--torch-seed", type=int, default=None)
    ap.add_argument("--deterministic-data", action="store_true",
                    help="parity tests: skip the reference's Random* transforms")
    args = ap.parse_args()

    if args.workdir:
        os.makedirs(args.workdir, exist_ok=True)
        os.chdir(args.workdir)
    os.makedirs("checkpoint", exist_ok=True)
    sys.path.insert(0, REF_SRC)
    install_multipledispatch_shim()
    patch_stty()
    import torch
    if args.torch_seed is not None:
        torch.manual_seed(args.torch_seed)
    if args.dp_shim:
        torch.nn.DataParallel = lambda m, *a, **k: m

    if args.role == "worker":
        install_synthetic_cifar(args.n_train, args.n_test, args.data_seed, args.pad_from, args.deterministic_data)
        rebind_model(args.model)
        sys.argv = ["client.py", "-a", args.address] + (["-c", args.compress] if args.compress else [])
        runpy.run_path(os.path.join(REF_SRC, "client.py"), run_name="__main__")
        return

    rebind_model(args.model)
    import importlib
    sys.argv = ["server.py"]
    server = importlib.import_module("server")       # imported as a module: its __main__ block does not run
    server.args = types.SimpleNamespace(compressFlag=args.compress, p="y" if args.role == "primary" else "n",
                                        backupAddress="localhost", backupPort=args.backup_port)
    server.compressFlag = args.compress == "Y"
    for w in args.workers.split(","):
        server.clients[w] = True                      # src/server.py:281-282 with N entries
    rounds = args.rounds
    server.range = lambda *a: builtins.range(rounds) if a == (20,) else builtins.range(*a)   # src/server.py:120

    if args.role == "primary":
        starts = []
        real_print = builtins.print

        def stamped_print(*a, **k):                   # "Starting epoch i" is printed at src/server.py:121
            if a and a[0] == "Starting epoch":
                starts.append(time.time())
            elif a and isinstance(a[0], str) and a[0].startswith(("Client status", "Recovering")):
                return
            real_print(*a, **k)
        server.print = stamped_print
        # harness ordering only: the reference's start-up is racy in two ways - a backup promotes itself
        # after 10-20 s without pings (src/server.py:254-264) and a primary whose backup push fails
        # poisons its own worker table - so this process is spawned first (its imports overlap
        # the backup's) and waits here until the backup listens before it replays the main block.
        import socket
        deadline = time.time() + 300
        while time.time() < deadline:
            try:
                socket.create_connection(("127.0.0.1", int(args.backup_port)), timeout=0.5).close()
                break
            except OSError:
                time.sleep(0.1)
        # replay of src/server.py:286-293
        server.ConnectToBackupServer("localhost", args.backup_port)
        server.mountPoint = "Primary"
        os.makedirs("Primary", exist_ok=True)
        t = threading.Thread(target=server.pingBackupServer, daemon=True)
        t.start()
        server.run()
        end = time.time()
        if args.timing_out:
            with open(args.timing_out, "w") as f:
                json.dump({"round_starts": starts, "end": end}, f)
        sys.stdout.flush()
        os._exit(0)                                   # the reference never exits on its own (B12)
    else:
        import signal
        server.print = lambda *a, **k: None if (a and a[0] == "Ping received") else builtins.print(*a, **k)
        # replay of src/server.py:294-301
        server.mountPoint = "Backup"
        os.makedirs("Backup", exist_ok=True)
        signal.signal(signal.SIGUSR1, server.handler)
        t = threading.Thread(target=server.CheckingIfPrimaryServerUp, daemon=True)
        t.start()
        server.serve(args.backup_port)


if __name__ == "__main__":
    main()
