#!/usr/bin/env python
"""Run-to-run noise of the first-step gradients: the same engine on the same batch three times (lr 0) and a second
engine with the same weights.  GRAPH=0/1, DT=bf16|fp32, DDPB_TF32, DDPB_POISON as usual.  Findings: tests/test_engine_gpu.py
(test_sub_bucket_hand_off_computes_the_same_step)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_engine_gpu import _cfg, _batches, _Fused, _T, _rel, DEV
from cs744_ddp_b200.backends import TorchBackend
from cs744_ddp_b200.engine import NativeBackend
from cs744_ddp_b200.ops.native import check
torch.manual_seed(0)
tb = TorchBackend(_cfg(backend="torch", dtype="fp32", global_batch=32), DEV)
graph = os.environ.get("GRAPH", "1") == "1"
dtype = os.environ.get("DT", "bf16")
mk = lambda: NativeBackend(_cfg(backend="native", dtype=dtype, use_cuda_graph=graph, global_batch=32, bucket_plan="per_layer"), DEV, init_state=tb.state_dict())
raws = _batches(2, 32, seed=3)
def grads_of(be, n=1):
    check(be._lib.ddpb_engine_set_lr(be._handle, ctypes.c_float(0.0)))
    out = []
    for _ in range(n):
        l = float(be.train_step(*be.prepare(raws[0]), _Fused() if graph else _T()))
        torch.cuda.synchronize()
        out.append(({k: v.clone() for k, v in be.grads().items()}, l))
    return out
a = mk(); ga = grads_of(a, 3)
b = mk(); gb = grads_of(b, 1)
keys = ["fc1.weight", "layers.26.weight", "layers.25.weight", "layers.15.weight", "layers.0.weight"]
print("dtype", dtype, "graph", graph, "losses", [g[1] for g in ga], gb[0][1])
for k in keys:
    print(f"{k:18s} same engine run0-run1 {_rel(ga[1][0][k], ga[0][0][k]):.2e}  run1-run2 {_rel(ga[2][0][k], ga[1][0][k]):.2e}   other engine {_rel(gb[0][0][k], ga[0][0][k]):.2e}")
