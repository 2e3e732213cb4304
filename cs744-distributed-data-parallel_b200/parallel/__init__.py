"""Data-parallel machinery: strategies, bucket planning, process-group bootstrap, P2P engine."""
from .strategies import (GradSync, NoSync, GatherScatterSync, AllReduceSync,  # noqa: F401
                         BucketedOverlapSync, make_strategy)
from .buckets import plan_buckets, arena_layout, Bucket  # noqa: F401
from .launch import init_process, shutdown  # noqa: F401
