"""Datasets for the local trainer.

The reference trains on torchvision CIFAR-10 with RandomCrop(32, 4) + RandomHorizontalFlip +
Normalize on the CPU, per sample, even for batches a worker then skips (src/main.py:36-56,
142-145).  There is no network here, so the default dataset is synthetic and CIFAR-shaped:
uint8 HWC images and int64 labels from a seeded generator, held once in (pinned) host memory.
Normalisation and augmentation happen per batch - on the device when the trainer is on a GPU
(csrc/augment.cu), vectorised on the CPU otherwise.

Labels are a deterministic function of the image (a random linear probe) so that training
actually reduces the loss and accuracy is a meaningful signal in the tests.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)


@dataclass
class ImageDataset:
    train_x: torch.Tensor   # uint8 [N, 32, 32, 3]
    train_y: torch.Tensor   # int64 [N]
    test_x: torch.Tensor
    test_y: torch.Tensor
    name: str = "synthetic"

    def pin(self) -> "ImageDataset":
        if torch.cuda.is_available():
            for f in ("train_x", "train_y", "test_x", "test_y"):
                t = getattr(self, f)
                if not t.is_pinned():
                    setattr(self, f, t.pin_memory())
        return self

    def num_train_batches(self, batch_size: int) -> int:
        return (self.train_x.shape[0] + batch_size - 1) // batch_size


def synthetic_cifar(n_train: int = 50000, n_test: int = 10000, seed: int = 0,
                    pad_from: Optional[int] = None, classes: int = 10, learnable: bool = False,
                    signal: float = 0.35) -> ImageDataset:
    """Seeded CIFAR-shaped data.  ``pad_from=28`` draws 28x28 images and zero-pads them to 32x32
    (BASELINE config 1: LeNet needs 3x32x32).

    ``learnable``: class-conditional images - every class has a smooth random pattern (an 8x8 field, bilinearly
    enlarged) that is mixed into per-pixel noise with weight ``signal`` - so that a few federated rounds move the
    accuracy well above chance and accuracy COMPARISONS between two runs mean something (the accuracy gate of the
    lossy codec, tools/codec_gate.py).  The default keeps the byte-identical data of the benchmarks (uniform noise
    with labels from a random linear probe, which a small CNN barely learns)."""
    g = torch.Generator().manual_seed(seed)
    side = pad_from or 32
    probe = torch.randn(3 * side * side, classes, generator=g)
    if learnable:
        coarse = torch.randn(classes, 3, 8, 8, generator=g)
        pattern = torch.nn.functional.interpolate(coarse, size=(side, side), mode="bilinear", align_corners=False)
        pattern = pattern.permute(0, 2, 3, 1).contiguous()                 # [classes, side, side, 3]

    def draw(n):
        if learnable:
            y = torch.randint(0, classes, (n,), generator=g)
            noise = torch.randn(n, side, side, 3, generator=g)
            x = (127.5 + 64.0 * (signal * pattern[y] + noise)).clamp_(0, 255).to(torch.uint8)
        else:
            x = torch.randint(0, 256, (n, side, side, 3), dtype=torch.uint8, generator=g)
            y = ((x.reshape(n, -1).float() / 255.0 - 0.5) @ probe).argmax(1)
        if side != 32:
            lo = (32 - side) // 2
            full = torch.zeros(n, 32, 32, 3, dtype=torch.uint8)
            full[:, lo:lo + side, lo:lo + side] = x
            x = full
        return x, y

    tx, ty = draw(n_train)
    vx, vy = draw(n_test)
    return ImageDataset(tx, ty, vx, vy, name=f"synthetic{side}")


def normalize_batch(x_u8: torch.Tensor, device=None) -> torch.Tensor:
    """uint8 NHWC -> fp32 NCHW, (x/255 - mean)/std: ToTensor + Normalize of the reference."""
    x = x_u8.to(device) if device is not None else x_u8
    x = x.permute(0, 3, 1, 2).float().div_(255.0)
    mean = torch.tensor(CIFAR_MEAN, device=x.device).view(1, 3, 1, 1)
    std = torch.tensor(CIFAR_STD, device=x.device).view(1, 3, 1, 1)
    return (x - mean) / std


def augment_batch_cpu(x_u8: torch.Tensor, gen: torch.Generator) -> torch.Tensor:
    """RandomCrop(32, padding=4) + RandomHorizontalFlip on a uint8 NHWC batch."""
    n = x_u8.shape[0]
    padded = torch.zeros(n, 40, 40, 3, dtype=torch.uint8)
    padded[:, 4:36, 4:36] = x_u8
    dx = torch.randint(0, 9, (n,), generator=gen)
    dy = torch.randint(0, 9, (n,), generator=gen)
    flip = torch.rand(n, generator=gen) < 0.5
    out = torch.empty_like(x_u8)
    for i in range(n):
        img = padded[i, dy[i]:dy[i] + 32, dx[i]:dx[i] + 32]
        out[i] = img.flip(1) if flip[i] else img
    return out


def shard_batches(num_batches: int, rank: int, world: int):
    """The reference's shard rule (src/main.py:140-145): over the UNSHUFFLED batch sequence a
    counter is advanced before the test, so 0-based batch ``b`` belongs to rank ``(b+1) % world``.
    world=2: rank0 gets b=1,3,.. (195 of 391), rank1 gets b=0,2,.. (196)."""
    return [b for b in range(num_batches) if (b + 1) % world == rank]


def load_cifar10(root: str) -> ImageDataset:
    """The real CIFAR-10 from the standard python archive (``<root>/cifar-10-batches-py/data_batch_1..5, test_batch``, the
    files torchvision's ``CIFAR10(root, download=True)`` leaves behind - src/main.py:43-56).  Nothing is downloaded; the
    pickles are read directly, so torchvision is not needed.  Images come back as uint8 HWC like the synthetic set."""
    import os
    import pickle

    import numpy as np

    base = os.path.join(root, "cifar-10-batches-py")
    if not os.path.isdir(base):
        base = root

    def read(names):
        xs, ys = [], []
        for n in names:
            with open(os.path.join(base, n), "rb") as f:
                d = pickle.load(f, encoding="latin1")
            xs.append(np.asarray(d["data"], dtype=np.uint8).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1))
            ys.append(np.asarray(d.get("labels", d.get("fine_labels")), dtype=np.int64))
        return torch.from_numpy(np.ascontiguousarray(np.concatenate(xs))), torch.from_numpy(np.concatenate(ys))

    tx, ty = read([f"data_batch_{i}" for i in range(1, 6)])
    vx, vy = read(["test_batch"])
    return ImageDataset(tx, ty, vx, vy, name="cifar10")


def get_dataset(root: Optional[str] = None, n_train: int = 50000, n_test: int = 10000, seed: int = 0) -> ImageDataset:
    """CIFAR-10 from ``root`` when given, else the synthetic stand-in (this image has no network)."""
    return load_cifar10(root) if root else synthetic_cifar(n_train, n_test, seed=seed)


def get_mean_and_std(images_u8: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-channel mean of the per-image means and of the per-image standard deviations of a uint8 [N, H, W, 3] set
    scaled to [0, 1] - the statistic the reference's (unused, broken: it never imports torch) ``get_mean_and_std``
    describes (src/utils.py:15-27), computed in one vectorised pass instead of a batch-size-1 DataLoader loop."""
    x = images_u8.to(torch.float32).div_(255.0).flatten(1, 2)         # [N, H*W, 3]
    return x.mean(1).mean(0), x.std(1, unbiased=True).mean(0)
