"""Command lines.  The reference's flags keep their spelling and meaning:

  server:  --p y|n   -c/--compressFlag Y   --backupAddress HOST   --backupPort PORT
  client:  -a/--address HOST:PORT   -c/--compressFlag Y   -r/--resume   --lr LR

and everything the reference hard-codes (B10) became a flag: --workers, --rounds, --model,
--batch-size, --watchdog, --plane, --device, --dataset-size, --data-root, --seed, --metrics.
"""
from __future__ import annotations

import argparse
import logging
import sys

from ..utils.metrics import JsonlSink, setup_logging


def _common(p: argparse.ArgumentParser):
    p.add_argument("-c", "--compressFlag", help="Y enables gzip on coordinator<->worker channels")
    p.add_argument("--model", default="MobileNet", help="model zoo name (reference default MobileNet)")
    p.add_argument("--metrics", default=None, help="JSONL metrics file")
    p.add_argument("--seed", type=int, default=None)


def server_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="fedb200 coordinator (primary or backup)")
    _common(p)
    p.add_argument("--p", default="n", help="y = run as primary, anything else = backup")
    p.add_argument("--backupAddress", default="localhost")
    p.add_argument("--backupPort", default="8080")
    p.add_argument("--workers", default="localhost:50051,localhost:50052",
                   help="comma separated worker addresses")
    p.add_argument("--rounds", type=int, default=20)
    p.add_argument("--watchdog", type=float, default=10.0, help="backup failure-detection window (s)")
    p.add_argument("--ping-period", type=float, default=1.0)
    p.add_argument("--replicate", choices=("sync", "async"), default="async")
    p.add_argument("--weighted", action="store_true", help="sample-count weighted FedAvg")
    p.add_argument("--no-backup", action="store_true", help="primary without a backup")
    p.add_argument("--state-dir", default=None, help="default Primary/ or Backup/")
    p.add_argument("--train-timeout", type=float, default=3600.0)
    p.add_argument("--plane", choices=("legacy", "peer"), default="legacy",
                   help="peer: workers exchange weights over NVLink peer memory, RPCs carry descriptors")
    p.add_argument("--checkpoint-every", type=int, default=0, help="peer plane: dump optimizedModel.pth every k rounds")
    return p


def server_main(argv=None) -> int:
    from .coordinator import CoordinatorConfig
    from .ha import BackupConfig, BackupNode, run_primary
    args = server_parser().parse_args(argv)
    setup_logging()
    compress = args.compressFlag == "Y"
    print(f"Compression {args.compressFlag} enabled")
    primary = args.p == "y"
    ccfg = CoordinatorConfig(
        workers=tuple(w for w in args.workers.split(",") if w), rounds=args.rounds, model=args.model,
        compress=compress, state_dir=args.state_dir or ("Primary" if primary else "Backup"),
        backup_target=None if (args.no_backup or not primary) else f"{args.backupAddress}:{args.backupPort}",
        replicate=args.replicate, weighted=args.weighted, metrics_path=args.metrics, plane=args.plane,
        checkpoint_every=args.checkpoint_every,
        train_timeout_s=args.train_timeout, send_timeout_s=args.train_timeout)
    sink = JsonlSink(args.metrics)
    if primary:
        print("Primary triggered")
        coord = run_primary(ccfg, ping_period_s=args.ping_period, sink=sink)
        return 0 if not coord.deposed else 3
    print("Backup triggered")
    node = BackupNode(BackupConfig(port=args.backupPort, watchdog_s=args.watchdog,
                                   state_dir=ccfg.state_dir, coordinator=ccfg), sink=sink)
    node.start(install_signal=True)
    try:
        node.wait()
    except KeyboardInterrupt:
        node.stop()
    return 0


def client_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="fedb200 training worker")
    _common(p)
    p.add_argument("-a", "--address", default="temp", help="listen address and checkpoint stem")
    p.add_argument("-r", "--resume", action="store_true", help="resume from ./checkpoint/<address>.pth")
    p.add_argument("--lr", type=float, default=0.1)
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--device", default=None, help="cpu | cuda | cuda:N (default: cuda if available)")
    p.add_argument("--plane", choices=("legacy", "peer"), default="legacy")
    p.add_argument("--dataset-size", type=int, default=50000)
    p.add_argument("--test-size", type=int, default=10000)
    p.add_argument("--data-root", default=None,
                   help="directory holding cifar-10-batches-py: train on the real CIFAR-10 instead of the synthetic set")
    p.add_argument("--data-seed", type=int, default=0)
    p.add_argument("--pad-from", type=int, default=None, help="draw SxS images, zero-pad to 32x32")
    p.add_argument("--augment", action="store_true", help="RandomCrop(32,4)+flip like the reference")
    p.add_argument("--no-eval", action="store_true", help="skip the eval inside SendModel")
    p.add_argument("--progress", action="store_true", help="reference-style per-batch progress bar when stdout is a TTY (CPU trainer)")
    p.add_argument("--peer-rank", type=int, default=0, help="peer plane: this worker's index in the coordinator's --workers list")
    p.add_argument("--peer-world", type=int, default=1, help="peer plane: number of configured workers")
    p.add_argument("--peer-tag", default="fed", help="peer plane: job tag shared by all workers (names the unix sockets)")
    p.add_argument("--codec", choices=("none", "fp8"), default="none",
                   help="peer plane meaning of -c: none = lossless fp32 words, fp8 = block-scaled e4m3 deltas")
    p.add_argument("--agg-timeout", type=float, default=20.0, help="peer plane: bound of the device-flag waits (s)")
    p.add_argument("--profile", action="store_true",
                   help="peer plane: NVTX ranges + device-timed phases of every round (train/agg/repl/eval ms, NVLink bytes, "
                        "GB/s) as 'round_profile' records in --metrics; synchronises the GPU once per round")
    p.add_argument("--no-multicast", action="store_true",
                   help="peer plane: broadcast with unicast peer stores even where NVLS multicast is available")
    p.add_argument("--dump-root", default=".", help="peer plane: 'dump' descriptors may only write below this directory")
    p.add_argument("--rejoin", action="store_true", help="peer plane: this worker restarts into a running job")
    return p


def client_main(argv=None) -> int:
    import torch
    from ..utils import checkpoint as ckpt
    from ..utils import data as D
    from .trainer import LocalTrainer, TrainerConfig
    from .worker import WorkerConfig, serve_worker
    args = client_parser().parse_args(argv)
    setup_logging()
    device = args.device or ("cuda" if torch.cuda.is_available() else "cpu")
    print(f"Client is running on {args.address} ")
    print(f"Compression {args.compressFlag} enabled")
    ds = D.load_cifar10(args.data_root) if args.data_root else \
        D.synthetic_cifar(args.dataset_size, args.test_size, seed=args.data_seed, pad_from=args.pad_from)
    tcfg = TrainerConfig(model=args.model, lr=args.lr, batch_size=args.batch_size, device=device,
                         seed=args.seed, augment=args.augment, resume=args.resume, progress=args.progress,
                         checkpoint_path=ckpt.worker_checkpoint_path(args.address))
    wcfg = WorkerConfig(address=args.address, compress=args.compressFlag == "Y", plane=args.plane,
                        eval_on_send=not args.no_eval, dump_root=args.dump_root)
    if args.plane == "peer":
        from ..parallel.session import GpuFederation, PeerWorker
        dev = torch.device(device)
        fed = GpuFederation(model=args.model, rank=args.peer_rank, world=args.peer_world,
                            device=dev.index if dev.index is not None else args.peer_rank, dataset=ds, seed=args.seed,
                            tag=args.peer_tag, lr=args.lr, batch_size=args.batch_size, augment=args.augment,
                            codec=args.codec, use_multicast=not (args.no_multicast or args.rejoin), timeout_ns=int(args.agg_timeout * 1e9),
                            persistent_server=True, rejoin=args.rejoin, eval_mode="inline", profile=args.profile)
        fed.trainer.cfg.checkpoint_path = tcfg.checkpoint_path
        serve_worker(fed.trainer, wcfg, sink=JsonlSink(args.metrics), peer_endpoint=PeerWorker(fed), block=True)
        return 0
    trainer = LocalTrainer(tcfg, ds)
    serve_worker(trainer, wcfg, sink=JsonlSink(args.metrics), block=True)
    return 0
