"""Seeding (ref ``Part 2a/main.py:20-21``): torch and numpy seeded with the same value on every
rank, which is what keeps the replicas' initial weights identical in Parts 2a/2b."""
import numpy as np
import torch


def seed_everything(seed: int = 0) -> None:
    torch.manual_seed(seed)
    np.random.seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
