"""Sampler / loader parity with torch's samplers."""
import math

import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

from cs744_ddp_b200.data import (BatchLoader, GlobalShuffleSampler, SequentialSampler,
                                 ShardedSampler, synthetic_cifar10, transform_batch)
from cs744_ddp_b200.data.pipeline import RawBatch


class _Len:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("n", [50_000, 1003])
def test_sharded_sampler_matches_distributed_sampler(world, n):
    for rank in range(world):
        ref = list(DistributedSampler(_Len(n), num_replicas=world, rank=rank))
        ours = ShardedSampler(n, world, rank).indices()
        assert ours == ref


def test_global_shuffle_matches_random_sampler():
    torch.manual_seed(0)
    ref = list(torch.utils.data.RandomSampler(_Len(1000)))
    torch.manual_seed(0)
    ours = GlobalShuffleSampler(1000).indices()
    assert ours == ref


@pytest.mark.parametrize("world,per_rank,last", [(1, 256, 80), (2, 128, 40), (4, 64, 20), (8, 32, 10)])
def test_iterations_per_epoch_and_last_batch(world, per_rank, last):
    s = ShardedSampler(50_000, world, 0)
    split = synthetic_cifar10(64)            # content irrelevant: only index math is checked
    loader = BatchLoader(split, s, per_rank, True)
    assert len(loader) == 196
    chunks = loader.index_batches()
    assert len(chunks) == 196 and chunks[-1].numel() == last


def test_test_loader_last_batch_is_16():
    for per_rank, nb in [(256, 40), (128, 79), (64, 157), (32, 313)]:
        loader = BatchLoader(synthetic_cifar10(8), SequentialSampler(10_000), per_rank, False)
        assert len(loader) == nb
        assert loader.index_batches()[-1].numel() == 16


def test_transform_matches_torchvision_semantics():
    split = synthetic_cifar10(16, seed=3)
    mean = torch.tensor([125.3, 123.0, 113.9]) / 255
    std = torch.tensor([63.0, 62.1, 66.7]) / 255
    # test transform: ToTensor + Normalize
    b = RawBatch(split.images, split.labels, None, None)
    x = transform_batch(b)
    ref = (split.images.float() / 255 - mean) / std
    assert torch.allclose(x, ref.permute(0, 3, 1, 2), atol=1e-6)
    # train transform with explicit draws: crop from the 4-padded image, then flip
    crop = torch.tensor([[0, 0], [8, 8], [4, 4], [3, 7]] * 4, dtype=torch.int32)
    flip = torch.tensor([0, 1] * 8, dtype=torch.uint8)
    x = transform_batch(RawBatch(split.images, split.labels, crop, flip))
    for i in range(16):
        img = split.images[i].float() / 255
        pad = torch.zeros(40, 40, 3)
        pad[4:36, 4:36] = img
        ox, oy = int(crop[i, 0]), int(crop[i, 1])
        c = pad[oy:oy + 32, ox:ox + 32]
        if flip[i]:
            c = c.flip(1)
        c = ((c - mean) / std).permute(2, 0, 1)
        assert torch.allclose(x[i], c, atol=1e-6)
    # identity crop (offset 4,4) without flip equals the test transform
    assert torch.allclose(x[2], ref[2].permute(2, 0, 1), atol=1e-6)


def test_loader_batches_are_deterministic_per_seed():
    split = synthetic_cifar10(100)
    a = [b.crop_xy.clone() for b in BatchLoader(split, ShardedSampler(100, 2, 0), 16, True, aug_seed=1)]
    b = [b.crop_xy.clone() for b in BatchLoader(split, ShardedSampler(100, 2, 0), 16, True, aug_seed=1)]
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    assert len(a) == math.ceil(50 / 16)
