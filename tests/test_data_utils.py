"""Dataset and initialisation utilities (reference: src/main.py:43-56 CIFAR-10 loading, src/utils.py:15-42 helpers)."""
import os
import pickle

import numpy as np
import torch

from fedb200 import models
from fedb200.parallel.arena import Arena
from fedb200.utils import data as D
from fedb200.utils.init import init_params


def _fake_cifar(root, per_batch=20):
    base = os.path.join(root, "cifar-10-batches-py")
    os.makedirs(base)
    rng = np.random.default_rng(0)
    blobs = {}
    for name in [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]:
        data = rng.integers(0, 256, size=(per_batch, 3072), dtype=np.uint8)
        labels = rng.integers(0, 10, size=per_batch).tolist()
        with open(os.path.join(base, name), "wb") as f:
            pickle.dump({"data": data, "labels": labels}, f)
        blobs[name] = (data, labels)
    return blobs


def test_cifar10_archive_loader(tmp_path):
    blobs = _fake_cifar(str(tmp_path))
    ds = D.load_cifar10(str(tmp_path))
    assert ds.train_x.shape == (100, 32, 32, 3) and ds.train_x.dtype == torch.uint8 and ds.test_x.shape == (20, 32, 32, 3)
    data, labels = blobs["data_batch_2"]
    # archive rows are planar RGB (3 x 1024); the dataset is HWC like torchvision's .data
    want = torch.from_numpy(data[3].reshape(3, 32, 32).transpose(1, 2, 0).copy())
    assert torch.equal(ds.train_x[20 + 3], want) and int(ds.train_y[20 + 3]) == labels[3]
    assert D.get_dataset(str(tmp_path)).name == "cifar10" and D.get_dataset(None, 64, 16).train_x.shape[0] == 64


def test_mean_and_std_statistic():
    x = torch.randint(0, 256, (50, 32, 32, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(1))
    mean, std = D.get_mean_and_std(x)
    f = x.float().div(255).permute(0, 3, 1, 2)
    want_m = torch.stack([torch.stack([f[n, c].mean() for n in range(50)]).mean() for c in range(3)])
    want_s = torch.stack([torch.stack([f[n, c].std() for n in range(50)]).mean() for c in range(3)])
    torch.testing.assert_close(mean, want_m)
    torch.testing.assert_close(std, want_s)


def test_init_params_on_an_adopted_model():
    torch.manual_seed(0)
    m = models.build("ResNet18")
    arena = Arena.adopt(m, "cpu")
    init_params(m)
    sd = arena.state_dict(to_cpu=True)                     # the arena sees the new values: parameters are views into it
    assert float(sd["bn1.weight"].min()) == 1.0 and float(sd["bn1.bias"].abs().max()) == 0.0
    assert float(sd["linear.bias"].abs().max()) == 0.0 and float(sd["linear.weight"].std()) < 2e-3
    w = sd["layer1.0.conv1.weight"]                        # Kaiming normal over fan_out = 64 * 9
    assert abs(float(w.std()) - (2.0 / (64 * 9)) ** 0.5) < 5e-3
