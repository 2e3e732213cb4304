"""CIFAR-10 storage: the whole split as one uint8 tensor ``[N,32,32,3]`` plus int64 labels.

The reference downloads CIFAR-10 through torchvision on every rank (ref ``Part 2a/main.py:36-37,
48-49``).  Here the split is a single tensor that can sit in host (pinned) or device memory, so
a batch is a gather by index instead of a Python loop over PIL images.  Two sources:

* ``load_cifar10(root, train)`` reads the standard ``cifar-10-batches-py`` pickles if they are
  already on disk (no network access is attempted);
* ``synthetic_cifar10(n, seed)`` makes CIFAR-shaped random images with learnable structure
  (class-dependent colour bias), used by benchmarks and tests.
"""
from __future__ import annotations

import dataclasses
import os
import pickle
from typing import Optional

import numpy as np
import torch


@dataclasses.dataclass
class ImageSplit:
    images: torch.Tensor    # uint8 [N,32,32,3] (HWC, like the PIL images torchvision hands out)
    labels: torch.Tensor    # int64 [N]
    name: str = "cifar10"

    def __len__(self) -> int:
        return int(self.images.shape[0])

    def to(self, device, pin: bool = False) -> "ImageSplit":
        images, labels = self.images, self.labels
        if str(device) != str(images.device):
            images, labels = images.to(device), labels.to(device)
        elif pin and images.device.type == "cpu" and torch.cuda.is_available():
            images, labels = images.pin_memory(), labels.pin_memory()
        return ImageSplit(images, labels, self.name)


def synthetic_cifar10(n: int, seed: int = 0, train: bool = True) -> ImageSplit:
    g = torch.Generator()
    g.manual_seed(seed * 2 + (0 if train else 1) + 12345)
    labels = torch.randint(0, 10, (n,), generator=g, dtype=torch.int64)
    base = torch.randint(0, 256, (n, 32, 32, 3), generator=g, dtype=torch.int32)
    # class-dependent tint so that a model can actually learn something from the data
    tint = torch.randint(-40, 41, (10, 1, 1, 3), generator=g, dtype=torch.int32)
    images = (base + tint[labels]).clamp_(0, 255).to(torch.uint8)
    return ImageSplit(images, labels, "synthetic")


def load_cifar10(root: str = "./data", train: bool = True) -> ImageSplit:
    folder = os.path.join(root, "cifar-10-batches-py")
    if not os.path.isdir(folder):
        raise FileNotFoundError(
            f"{folder} not found. This framework never downloads data (the GPU boxes have no "
            f"network); place the extracted CIFAR-10 python archive there or pass --synthetic.")
    names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
    xs, ys = [], []
    for nm in names:
        with open(os.path.join(folder, nm), "rb") as fh:
            entry = pickle.load(fh, encoding="latin1")
        xs.append(np.asarray(entry["data"], dtype=np.uint8))
        ys.extend(entry.get("labels", entry.get("fine_labels")))
    data = np.concatenate(xs).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)  # -> NHWC
    return ImageSplit(torch.from_numpy(np.ascontiguousarray(data)),
                      torch.tensor(ys, dtype=torch.int64), "cifar10")


def get_split(train: bool, synthetic: bool, root: str = "./data", size: Optional[int] = None,
              seed: int = 0) -> ImageSplit:
    if synthetic:
        return synthetic_cifar10(size or (50_000 if train else 10_000), seed, train)
    split = load_cifar10(root, train)
    if size is not None and size < len(split):
        split = ImageSplit(split.images[:size].clone(), split.labels[:size].clone(), split.name)
    return split
