"""VGG family for 3x32x32 inputs / 10 classes, and the static facts the engine derives from it.

Parity contract (ref ``Part 1/model.py:3-50``): the ``state_dict`` has exactly the reference's
58 keys for VGG-11 (``layers.<i>.{weight,bias,running_mean,running_var,num_batches_tracked}``
and ``fc1.{weight,bias}``), fp32, conv weights OIHW, so a checkpoint written here loads into the
reference model with ``strict=True`` and vice versa.

Unlike the reference (which only exposes ``VGG11()``), all four configurations in its table
are constructible here.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Sequence, Tuple

import torch
from torch import nn

POOL = "M"

# channel plan per variant; "M" = 2x2/stride-2 max-pool   (ref Part 1/model.py:3-8)
VGG_PLANS: Dict[str, Tuple] = {
    "VGG11": (64, POOL, 128, POOL, 256, 256, POOL, 512, 512, POOL, 512, 512, POOL),
    "VGG13": (64, 64, POOL, 128, 128, POOL, 256, 256, POOL, 512, 512, POOL, 512, 512, POOL),
    "VGG16": (64, 64, POOL, 128, 128, POOL, 256, 256, 256, POOL, 512, 512, 512, POOL,
              512, 512, 512, POOL),
    "VGG19": (64, 64, POOL, 128, 128, POOL, 256, 256, 256, 256, POOL, 512, 512, 512, 512, POOL,
              512, 512, 512, 512, POOL),
}
NUM_CLASSES = 10
IMAGE_HW = 32
IN_CHANNELS = 3
FLATTEN_FEATURES = 512


@dataclasses.dataclass(frozen=True)
class ConvBlock:
    """One conv3x3+BN+ReLU(+pool) stage as the native engine sees it."""
    index: int          # 0-based conv index
    seq_conv: int       # position of the Conv2d inside ``layers`` (state_dict key)
    seq_bn: int         # position of the BatchNorm2d
    cin: int
    cout: int
    hw: int             # input (= conv output) spatial size
    pool: bool          # a 2x2 max-pool follows the ReLU

    @property
    def out_hw(self) -> int:
        return self.hw // 2 if self.pool else self.hw


def conv_blocks(name: str = "VGG11") -> List[ConvBlock]:
    plan = VGG_PLANS[name]
    blocks: List[ConvBlock] = []
    cin, hw, seq = IN_CHANNELS, IMAGE_HW, 0
    for i, item in enumerate(plan):
        if item == POOL:
            seq += 1
            continue
        pool = i + 1 < len(plan) and plan[i + 1] == POOL
        blocks.append(ConvBlock(len(blocks), seq, seq + 1, cin, int(item), hw, pool))
        seq += 3
        cin = int(item)
        if pool:
            hw //= 2
    return blocks


class VGG(nn.Module):
    """``layers`` (Sequential of conv/BN/ReLU/pool) followed by ``fc1``; same keys as the ref."""

    def __init__(self, name: str = "VGG11") -> None:
        super().__init__()
        if name not in VGG_PLANS:
            raise KeyError(f"unknown VGG variant {name!r}; choose from {sorted(VGG_PLANS)}")
        self.variant = name
        stages: List[nn.Module] = []
        width = IN_CHANNELS
        for item in VGG_PLANS[name]:
            if item == POOL:
                stages.append(nn.MaxPool2d(2, 2))
                continue
            stages += [nn.Conv2d(width, item, 3, 1, 1, bias=True), nn.BatchNorm2d(item),
                       nn.ReLU(inplace=True)]
            width = item
        self.layers = nn.Sequential(*stages)
        self.fc1 = nn.Linear(FLATTEN_FEATURES, NUM_CLASSES)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        feats = self.layers(x)
        return self.fc1(feats.flatten(1))


def VGG11() -> VGG:
    return VGG("VGG11")


def VGG13() -> VGG:
    return VGG("VGG13")


def VGG16() -> VGG:
    return VGG("VGG16")


def VGG19() -> VGG:
    return VGG("VGG19")


def build_model(name: str = "VGG11") -> VGG:
    return VGG(name)


# ----------------------------------------------------------------------------------------------
# Static parameter table (declaration order == ``model.parameters()`` order)
# ----------------------------------------------------------------------------------------------

@dataclasses.dataclass(frozen=True)
class ParamSpec:
    name: str
    shape: Tuple[int, ...]
    kind: str           # conv_w | conv_b | bn_w | bn_b | fc_w | fc_b
    block: int          # conv block index, or -1 for the classifier

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


def param_specs(name: str = "VGG11") -> List[ParamSpec]:
    specs: List[ParamSpec] = []
    for b in conv_blocks(name):
        specs.append(ParamSpec(f"layers.{b.seq_conv}.weight", (b.cout, b.cin, 3, 3), "conv_w", b.index))
        specs.append(ParamSpec(f"layers.{b.seq_conv}.bias", (b.cout,), "conv_b", b.index))
        specs.append(ParamSpec(f"layers.{b.seq_bn}.weight", (b.cout,), "bn_w", b.index))
        specs.append(ParamSpec(f"layers.{b.seq_bn}.bias", (b.cout,), "bn_b", b.index))
    specs.append(ParamSpec("fc1.weight", (NUM_CLASSES, FLATTEN_FEATURES), "fc_w", -1))
    specs.append(ParamSpec("fc1.bias", (NUM_CLASSES,), "fc_b", -1))
    return specs


def grad_ready_order(name: str = "VGG11") -> List[str]:
    """Order in which gradients become final during backward: the classifier
    first (bias, then weight), then each conv block from last to first with the BN pair before
    the conv pair and weight before bias inside each pair."""
    order = ["fc1.bias", "fc1.weight"]
    for b in reversed(conv_blocks(name)):
        order += [f"layers.{b.seq_bn}.weight", f"layers.{b.seq_bn}.bias",
                  f"layers.{b.seq_conv}.weight", f"layers.{b.seq_conv}.bias"]
    return order


def buffer_names(name: str = "VGG11") -> List[str]:
    out: List[str] = []
    for b in conv_blocks(name):
        out += [f"layers.{b.seq_bn}.running_mean", f"layers.{b.seq_bn}.running_var",
                f"layers.{b.seq_bn}.num_batches_tracked"]
    return out


def state_dict_keys(name: str = "VGG11") -> List[str]:
    """Key order of ``model.state_dict()`` (module order)."""
    keys: List[str] = []
    for b in conv_blocks(name):
        keys += [f"layers.{b.seq_conv}.weight", f"layers.{b.seq_conv}.bias",
                 f"layers.{b.seq_bn}.weight", f"layers.{b.seq_bn}.bias",
                 f"layers.{b.seq_bn}.running_mean", f"layers.{b.seq_bn}.running_var",
                 f"layers.{b.seq_bn}.num_batches_tracked"]
    keys += ["fc1.weight", "fc1.bias"]
    return keys


# ----------------------------------------------------------------------------------------------
# Layout helpers: checkpoint layout (OIHW) <-> engine layout (K,R,S,C == O,H,W,I)
# ----------------------------------------------------------------------------------------------

def oihw_to_krsc(w: torch.Tensor) -> torch.Tensor:
    return w.permute(0, 2, 3, 1).contiguous()


def krsc_to_oihw(w: torch.Tensor) -> torch.Tensor:
    return w.permute(0, 3, 1, 2).contiguous()


def strip_module_prefix(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Accept checkpoints written from a DDP-wrapped model (every key prefixed ``module.``)."""
    if sd and all(k.startswith("module.") for k in sd):
        return {k[len("module."):]: v for k, v in sd.items()}
    return dict(sd)


def forward_flops_per_image(name: str = "VGG11") -> float:
    f = 0.0
    for b in conv_blocks(name):
        f += 2.0 * b.hw * b.hw * b.cout * 9 * b.cin
    return f + 2.0 * FLATTEN_FEATURES * NUM_CLASSES


def train_flops_per_image(name: str = "VGG11") -> float:
    """fprop + dgrad + wgrad; the first conv needs no dgrad."""
    blocks = conv_blocks(name)
    fwd = forward_flops_per_image(name)
    first = 2.0 * blocks[0].hw ** 2 * blocks[0].cout * 9 * blocks[0].cin
    return 3.0 * fwd - first
