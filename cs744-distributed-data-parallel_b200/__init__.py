"""B200-native data-parallel training framework (VGG / CIFAR-10, four DP strategies).

Capabilities mirror rawahars/CS744-Distributed-Data-Parallel (reference ``src/Part {1,2a,2b,3}``):

* Part 1  – single-process trainer                         (``sync="none"``)
* Part 2a – rank-0 gather -> mean -> scatter averaging      (``sync="gather_scatter"``)
* Part 2b – per-parameter all-reduce averaging              (``sync="allreduce"``)
* Part 3  – bucketed all-reduce overlapped with backward    (``sync="bucketed"``)

Two execution back-ends sit behind one :class:`Trainer`:

* ``torch``  – plain PyTorch ops (CPU + gloo, or CUDA + NCCL).  It is the numerical oracle and
  the "Part 1 plumbing on CPU" configuration.
* ``native`` – the sm_100a engine in ``csrc/``: tcgen05/TMEM/TMA implicit-GEMM convolutions,
  fused BN/ReLU/pool, classifier + cross-entropy, fused SGD, and a peer-to-peer gradient
  engine over NVLink.  Loaded through a C ABI (``include/ddp_b200.h``) with ctypes.

Import as ``cs744_ddp_b200`` (see the alias package at the repository root).
"""

__version__ = "0.1.0"

from .config import TrainConfig, SYNC_STRATEGIES  # noqa: F401
