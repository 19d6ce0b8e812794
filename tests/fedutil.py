"""Helpers for the in-process federation tests: tiny dataset, LeNet workers on free ports."""
import socket

from fedb200.control.trainer import LocalTrainer, TrainerConfig
from fedb200.control.worker import WorkerConfig, serve_worker
from fedb200.utils import data as D
from fedb200.utils.metrics import JsonlSink


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def tiny_dataset(n_train=1024, n_test=200, seed=0):
    return D.synthetic_cifar(n_train, n_test, seed=seed)


class WorkerProc:
    """An in-process worker (its own trainer + gRPC server) that can be killed and restarted."""

    def __init__(self, tmp_path, seed, port=None, model="LeNet", dataset=None, plane="legacy",
                 eval_on_send=True, compress=False):
        self.port = port or free_port()
        self.address = f"127.0.0.1:{self.port}"
        self.tmp_path, self.seed, self.model, self.plane = tmp_path, seed, model, plane
        self.dataset = dataset or tiny_dataset()
        self.eval_on_send, self.compress = eval_on_send, compress
        self.sink = JsonlSink(None)
        self.trainer = LocalTrainer(TrainerConfig(model=model, device="cpu", seed=seed), self.dataset)
        self.server = self.servicer = None
        self.start()

    def start(self):
        cfg = WorkerConfig(address=self.address, plane=self.plane, eval_on_send=self.eval_on_send,
                           compress=self.compress, checkpoint_root=str(self.tmp_path / "checkpoint"),
                           dump_root=str(self.tmp_path))
        self.server, self.servicer = serve_worker(self.trainer, cfg, sink=self.sink, block=False)

    def kill(self):
        self.server.stop(0).wait()

    def restart(self):
        self.start()
