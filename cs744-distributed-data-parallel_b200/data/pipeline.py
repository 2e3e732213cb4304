"""Batch pipeline: index batches -> gathered uint8 images -> augmented, normalised tensors.

Replaces the reference's ``DataLoader(num_workers=2, pin_memory=True)`` over PIL transforms
(ref ``Part 1/main.py:82-109``) with batched tensor ops:

* train transform = RandomCrop(32, padding=4) -> RandomHorizontalFlip -> ToTensor -> Normalize
* test transform  = ToTensor -> Normalize

Per-sample crop offsets and flip bits are drawn once per batch on the host; the image math is
one fused op over the batch, either torch ops (any device) or the native ``augment`` kernel
(``ops.augment``) that writes NHWC bf16/fp32 directly for the sm_100a convolutions.
"""
from __future__ import annotations

import dataclasses
from typing import Iterator, List, Optional, Sequence

import torch

from ..config import CIFAR_MEAN, CIFAR_STD
from .cifar import ImageSplit
from .sampler import batches_per_epoch

PAD = 4


@dataclasses.dataclass
class RawBatch:
    """What the loader hands to a backend: untransformed pixels + the augmentation draw."""
    images: torch.Tensor            # uint8 [B,32,32,3]
    labels: torch.Tensor            # int64 [B]
    crop_xy: Optional[torch.Tensor]  # int32 [B,2] offsets in [0, 2*PAD], None => no crop
    flip: Optional[torch.Tensor]     # uint8 [B] 1 => mirror horizontally, None => no flip

    @property
    def size(self) -> int:
        return int(self.labels.shape[0])

    def to(self, device, non_blocking: bool = True) -> "RawBatch":
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)
        return RawBatch(mv(self.images), mv(self.labels), mv(self.crop_xy), mv(self.flip))

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size()
                   for t in (self.images, self.labels, self.crop_xy, self.flip) if t is not None)


def normalize_consts(device, dtype=torch.float32):
    mean = torch.tensor(CIFAR_MEAN, device=device, dtype=dtype)
    std = torch.tensor(CIFAR_STD, device=device, dtype=dtype)
    return mean, std


def transform_batch(batch: RawBatch, layout: str = "nchw",
                    dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """Torch implementation of the train/test transform (the oracle for the native kernel)."""
    img = batch.images
    b = img.shape[0]
    x = img.to(torch.float32) / 255.0                         # ToTensor
    if batch.crop_xy is not None:
        padded = torch.zeros((b, 32 + 2 * PAD, 32 + 2 * PAD, 3), dtype=x.dtype, device=x.device)
        padded[:, PAD:PAD + 32, PAD:PAD + 32] = x
        ar = torch.arange(32, device=x.device)
        ox = batch.crop_xy[:, 0].long().to(x.device)
        oy = batch.crop_xy[:, 1].long().to(x.device)
        rows = (oy[:, None] + ar[None, :])                    # [B,32]
        cols = (ox[:, None] + ar[None, :])
        bi = torch.arange(b, device=x.device)[:, None, None]
        x = padded[bi, rows[:, :, None], cols[:, None, :]]
    if batch.flip is not None:
        f = batch.flip.to(x.device).bool()
        x = torch.where(f[:, None, None, None], x.flip(2), x)
    mean, std = normalize_consts(x.device)
    x = (x - mean) / std                                      # Normalize (NHWC broadcast)
    if layout == "nchw":
        x = x.permute(0, 3, 1, 2).contiguous()
    return x.to(dtype)


class BatchLoader:
    """Iterates one epoch of ``RawBatch``es for one rank.

    ``len(loader)`` is the number of batches (what the reference prints as "Size of training
    set"), the last batch is short when the shard is not a multiple of the batch size.
    """

    def __init__(self, split: ImageSplit, sampler, batch_size: int, train: bool,
                 aug_seed: int = 0, drop_last: bool = False) -> None:
        if batch_size <= 0:
            raise ValueError("batch_size must be positive (world size larger than global batch?)")
        self.split = split
        self.sampler = sampler
        self.batch_size = int(batch_size)
        self.train = train
        self.drop_last = drop_last
        self._aug = torch.Generator()
        self._aug.manual_seed(1_000_003 * (aug_seed + 1))

    @property
    def dataset(self) -> ImageSplit:
        return self.split

    def __len__(self) -> int:
        return batches_per_epoch(len(self.sampler), self.batch_size, self.drop_last)

    def index_batches(self) -> List[torch.Tensor]:
        idx = torch.as_tensor(self.sampler.indices(), dtype=torch.int64)
        chunks = list(idx.split(self.batch_size))
        if self.drop_last and chunks and chunks[-1].numel() < self.batch_size:
            chunks.pop()
        return chunks

    def draw_augmentation(self, b: int):
        crop = torch.randint(0, 2 * PAD + 1, (b, 2), generator=self._aug, dtype=torch.int32)
        flip = (torch.rand(b, generator=self._aug) < 0.5).to(torch.uint8)
        return crop, flip

    def __iter__(self) -> Iterator[RawBatch]:
        dev = self.split.images.device
        for idx in self.index_batches():
            sel = idx.to(dev)
            images = self.split.images.index_select(0, sel)
            labels = self.split.labels.index_select(0, sel)
            if self.train:
                crop, flip = self.draw_augmentation(idx.numel())
            else:
                crop, flip = None, None
            yield RawBatch(images, labels, crop, flip)
