"""Bucket planner vs torch's own assignment routine."""
import torch
import torch.distributed as dist

from cs744_ddp_b200.models import vgg
from cs744_ddp_b200.parallel.buckets import MIB, arena_layout, plan_buckets


def _ready(name="VGG11"):
    specs = {s.name: s for s in vgg.param_specs(name)}
    order = vgg.grad_ready_order(name)
    return order, [specs[n].numel for n in order]


def test_ddp_plan_matches_torch_assignment():
    m = vgg.VGG11()
    rev = list(reversed(list(m.named_parameters())))
    ref, _ = dist._compute_bucket_assignment_by_size([p for _, p in rev], [1 * MIB, 25 * MIB])
    ours = plan_buckets([n for n, _ in rev], [p.numel() for _, p in rev], "ddp")
    assert [list(b.names) for b in ours] == [[rev[i][0] for i in g] for g in ref]
    assert [b.nbytes for b in ours] == [9_463_848, 27_156_480, 304_128]


def test_per_layer_and_single_plans():
    order, numels = _ready()
    single = plan_buckets(order, numels, "single")
    assert len(single) == 1 and single[0].nbytes == 36_924_456
    layer_of = {}
    for b in vgg.conv_blocks("VGG11"):
        for suffix in ("weight", "bias"):
            layer_of[f"layers.{b.seq_conv}.{suffix}"] = b.index
            layer_of[f"layers.{b.seq_bn}.{suffix}"] = b.index
    layer_of["fc1.weight"] = layer_of["fc1.bias"] = 99
    per = plan_buckets(order, numels, "per_layer", layer_of=layer_of)
    assert len(per) == 9
    assert per[0].names == ("fc1.bias", "fc1.weight")
    assert sum(b.nbytes for b in per) == 36_924_456


def test_arena_layout_alignment_and_ranges():
    order, numels = _ready()
    lay = arena_layout(order, numels)
    assert all(s.offset % 64 == 0 for s in lay.slots)
    assert lay.total >= 9_231_114 and lay.total % 64 == 0
    buckets = plan_buckets(order, numels, "ddp")
    ranges = lay.bucket_ranges(buckets)
    assert ranges[0][0] == 0
    for (o0, n0), (o1, _) in zip(ranges, ranges[1:]):
        assert o0 + n0 == o1
    assert ranges[-1][0] + ranges[-1][1] == lay.total
