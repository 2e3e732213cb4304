"""Checkpoint / resume in the torch ``state_dict`` format.

The reference never saves anything; the contract is therefore "what
``torch.save(model.state_dict())`` of the reference model would contain": the 58 keys of
SURVEY 2.5.3, fp32, OIHW conv weights.  Files written here load into the reference ``VGG11()``
with ``strict=True``; files written from a DDP-wrapped reference model (keys prefixed
``module.``) load here.  A full training checkpoint additionally carries the SGD momentum
buffers (``optimizer.state_dict()`` layout: tensors keyed by parameter index) and progress
counters.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Optional

import torch

from ..models.vgg import strip_module_prefix

FORMAT_VERSION = 1


def save_model(path: str, state_dict: Dict[str, torch.Tensor], ddp_prefix: bool = False) -> None:
    """Write exactly a model ``state_dict`` (what the reference's format means)."""
    sd = {("module." + k if ddp_prefix else k): v.detach().to("cpu") for k, v in state_dict.items()}
    _atomic_save(sd, path)


def save_training(path: str, model_sd: Dict[str, torch.Tensor], optim_sd: Dict[str, Any],
                  epoch: int, iteration: int, extra: Optional[Dict[str, Any]] = None) -> None:
    blob = {
        "format_version": FORMAT_VERSION,
        "model": {k: v.detach().to("cpu") for k, v in model_sd.items()},
        "optimizer": _to_cpu(optim_sd),
        "epoch": int(epoch),
        "iteration": int(iteration),
        "extra": extra or {},
    }
    _atomic_save(blob, path)


def load(path: str, allow_pickle: bool = False) -> Dict[str, Any]:
    """Returns ``{"model": sd, "optimizer": sd|None, "epoch": int, "iteration": int}`` for
    either file flavour (bare model state_dict or full training checkpoint).

    Files are read with torch's restricted unpickler (``weights_only=True``): a checkpoint holds
    only tensors, numbers, strings, dicts and lists, and this function is advertised for files
    produced elsewhere (the reference model, a DDP-wrapped model), so nothing in them may execute
    code.  ``allow_pickle=True`` is the explicit opt-in for a trusted legacy file that needs the
    full unpickler."""
    blob = torch.load(path, map_location="cpu", weights_only=not allow_pickle)
    if isinstance(blob, dict) and "model" in blob and "format_version" in blob:
        blob["model"] = strip_module_prefix(blob["model"])
        return blob
    return {"model": strip_module_prefix(blob), "optimizer": None, "epoch": 0, "iteration": 0,
            "extra": {}}


def _to_cpu(obj):
    if torch.is_tensor(obj):
        return obj.detach().to("cpu")
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def _atomic_save(obj, path: str) -> None:
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    tmp = path + ".tmp"
    torch.save(obj, tmp)
    os.replace(tmp, path)
