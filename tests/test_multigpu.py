"""Multi-GPU: the peer-to-peer engine against NCCL / the torch oracle (needs >= 2 GPUs)."""
import os
import socket
import tempfile

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

WORLD = 2
STEPS = 3
B = 16


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _T:
    is_async = True

    def mark(self, i):
        pass


def _worker(rank, port, sync, comm, dtype, outdir, plan):
    import torch.distributed as dist
    from cs744_ddp_b200.config import TrainConfig
    from cs744_ddp_b200.data import synthetic_cifar10
    from cs744_ddp_b200.data.pipeline import RawBatch
    from cs744_ddp_b200.engine import NativeBackend
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=WORLD, device_id=torch.device("cuda", rank))
    torch.manual_seed(0)
    cfg = TrainConfig(num_nodes=WORLD, rank=rank, sync=sync, comm=comm, dtype=dtype, device="cuda",
                      backend="native", global_batch=B * WORLD, synthetic=True, bucket_plan=plan, lr=0.01)
    be = NativeBackend(cfg, torch.device("cuda", rank))
    split = synthetic_cifar10(STEPS * B, seed=10 + rank)
    losses = []
    for s in range(STEPS):
        raw = RawBatch(split.images[s * B:(s + 1) * B], split.labels[s * B:(s + 1) * B], None, None)
        losses.append(float(be.train_step(*be.prepare(raw), _T())))
    torch.cuda.synchronize()
    sd = {k: v.cpu() for k, v in be.state_dict().items()}
    mom = {i: st["momentum_buffer"].cpu() for i, st in be.optimizer_state_dict()["state"].items()}
    torch.save({"sd": sd, "losses": losses, "comm": be.comm_name(), "mom": mom},
               os.path.join(outdir, f"{sync}_{comm}_{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def _run(sync, comm, dtype="fp32", plan="per_layer"):
    import torch.multiprocessing as mp
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(_free_port(), sync, comm, dtype, d, plan), nprocs=WORLD, join=True)
        return [torch.load(os.path.join(d, f"{sync}_{comm}_{r}.pt"), weights_only=False) for r in range(WORLD)]


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


@pytest.mark.parametrize("sync", ["gather_scatter", "allreduce", "bucketed"])
def test_p2p_strategy_matches_nccl_baseline(sync):
    p2p = _run(sync, "p2p")
    nccl = _run(sync, "nccl")
    assert p2p[0]["comm"] == "p2p" and nccl[0]["comm"] == "nccl"
    for r in range(WORLD):
        for a, b in zip(p2p[r]["losses"], nccl[r]["losses"]):
            assert abs(a - b) < 2e-3 * max(1.0, abs(b)), (sync, r, p2p[r]["losses"], nccl[r]["losses"])
    # replicas hold bit-identical weights after the fused reduce+SGD
    for k, v in p2p[0]["sd"].items():
        if "running" in k or "num_batches" in k:
            continue
        assert torch.equal(v, p2p[1]["sd"][k]), (sync, k)
        if not (k.endswith(".bias") and "layers" in k):
            assert _rel(v, nccl[0]["sd"][k]) < 5e-3, (sync, k)
    # momentum shards were gathered: every rank reports the full optimizer state
    for i, m in p2p[0]["mom"].items():
        assert torch.equal(m, p2p[1]["mom"][i]), (sync, i)


def test_nvls_multicast_reduce_matches_nccl_baseline():
    """K-S3: the in-switch reduction (multimem.ld_reduce / multimem.st) where the box exposes NVSwitch
    multicast; skipped (not failed) when it does not."""
    from cs744_ddp_b200.parallel.vmm import VmmArena
    if not VmmArena.supported(torch.device("cuda", 0)):
        pytest.skip("this device/driver does not report multicast support")
    os.environ["DDPB_NVLS"] = "1"
    try:
        got = _run("bucketed", "auto")
    finally:
        os.environ.pop("DDPB_NVLS", None)
    if got[0]["comm"] != "nvls":
        pytest.skip(f"multicast arena could not be set up on this box (fell back to {got[0]['comm']})")
    nccl = _run("bucketed", "nccl")
    for r in range(WORLD):
        for a, b in zip(got[r]["losses"], nccl[r]["losses"]):
            assert abs(a - b) < 2e-3 * max(1.0, abs(b)), (got[r]["losses"], nccl[r]["losses"])
    for k, v in got[0]["sd"].items():
        if "running" in k or "num_batches" in k:
            continue
        assert torch.equal(v, got[1]["sd"][k]), k
        if not (k.endswith(".bias") and "layers" in k):
            assert _rel(v, nccl[0]["sd"][k]) < 5e-3, k


def test_bucketed_bf16_ddp_plan_runs_and_replicas_agree():
    out = _run("bucketed", "p2p", dtype="bf16", plan="ddp")
    for k, v in out[0]["sd"].items():
        if "running" in k or "num_batches" in k:
            continue
        assert torch.equal(v, out[1]["sd"][k]), k
    assert all(abs(a - b) < 1.0 for a, b in zip(out[0]["losses"], out[1]["losses"]))


def _stress_worker(rank, port, outdir):
    """Many fused reduce+SGD calls in every mode, with random host sleeps and random device-side spins in
    front of each one so that the ranks reach every cross-GPU barrier in a different order every time."""
    import random
    import time
    import torch.distributed as dist
    from cs744_ddp_b200.ops import native as nv
    from cs744_ddp_b200.parallel.p2p import SymmetricArena
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=WORLD, device_id=dev)
    n = 64 * 1024 + 192
    arena = SymmetricArena(n, dev, with_shadow=True)
    rng = random.Random(1234 + rank)            # a different delay schedule on every rank
    g = torch.Generator(device="cpu").manual_seed(7)
    lr, mu, wd = 0.05, 0.9, 1e-4
    bad = 0
    # one phase per mode (the two-shot and root modes shard / centralise the momentum, so modes do not mix)
    for phase, mode in enumerate([nv.P2P_ONESHOT, nv.P2P_TWOSHOT, nv.P2P_ROOT]):
        expect = torch.zeros(arena.numel)
        mom_ref = torch.zeros(arena.numel)
        torch.cuda.synchronize(); dist.barrier()
        arena.master.zero_(); arena.mom.zero_()
        torch.cuda.synchronize(); dist.barrier()
        for it in range(20):
            grads = [torch.randn(arena.numel, generator=g) for _ in range(WORLD)]     # same numbers on all ranks
            arena.grad.copy_(grads[rank].to(dev))
            time.sleep(rng.random() * 0.004)                         # host jitter: 0 - 4 ms
            torch.cuda._sleep(int(rng.random() * 4_000_000))         # device jitter: 0 - ~2 ms of spinning
            arena.sync(mode, 0, arena.numel, lr=lr, momentum=mu, wd=wd, slot=phase * 4 + it % 4)
            mean = sum(grads) / WORLD + wd * expect
            mom_ref = mu * mom_ref + mean
            expect = expect - lr * mom_ref
            torch.cuda.synchronize()
            err = float((arena.master.cpu() - expect).abs().max())
            if err > 1e-4 * (1.0 + float(expect.abs().max())):
                bad += 1
    torch.save({"bad": bad}, os.path.join(outdir, f"stress_{rank}.pt"))
    dist.barrier()
    arena.close()
    dist.destroy_process_group()


def test_sync_kernels_survive_random_arrival_order():
    """racecheck cannot see peer accesses, so the flag/epoch barrier protocol is stressed with
    randomised host and device delays; every call must still produce the exact averaged SGD update."""
    import torch.multiprocessing as mp
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_stress_worker, args=(_free_port(), d), nprocs=WORLD, join=True)
        res = [torch.load(os.path.join(d, f"stress_{r}.pt"), weights_only=False) for r in range(WORLD)]
    assert all(r["bad"] == 0 for r in res), res
