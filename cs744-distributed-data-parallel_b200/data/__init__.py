"""Data path: CIFAR storage, sharded samplers, the batched augmentation pipeline."""
from .cifar import ImageSplit, get_split, synthetic_cifar10, load_cifar10  # noqa: F401
from .sampler import ShardedSampler, GlobalShuffleSampler, SequentialSampler  # noqa: F401
from .pipeline import BatchLoader, RawBatch, transform_batch  # noqa: F401
