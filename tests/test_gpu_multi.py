"""Multi-GPU checks (skipped with fewer than 2 GPUs): the peer plane under torchrun and the gRPC-driven GPU plane
with worker loss / re-admission and primary failover.  The heavy lifting lives in tools/ so the same scripts can
be launched by hand with torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_gpus() < 2, reason="needs >= 2 GPUs")
def test_peer_plane_under_torchrun():
    n = min(_gpus(), 4)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "multigpu_check.py")],
                         capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert "MULTIGPU_CHECK_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.skipif(_gpus() < 2, reason="needs >= 2 GPUs")
def test_grpc_control_plane_drives_gpu_data_plane_with_failures():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "grpc_gpu_check.py")], capture_output=True, text=True,
                         timeout=900, cwd=ROOT)
    assert "GRPC_GPU_CHECK_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
