"""The native step engine against the torch back-end (the reference's computation) on one GPU."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = torch.device("cuda", 0)


class _T:            # region timer stub (forces the split forward/backward path)
    is_async = True

    def mark(self, i):
        pass


class _Fused:        # no ``is_async``: the engine runs the whole step as one graph
    def mark(self, i):
        pass


def _cfg(**kw):
    from cs744_ddp_b200.config import TrainConfig
    base = dict(synthetic=True, global_batch=16, device="cuda", log_every=4, lr=0.01)
    base.update(kw)
    return TrainConfig(**base)


def _batches(n, b, seed=0):
    from cs744_ddp_b200.data import synthetic_cifar10
    from cs744_ddp_b200.data.pipeline import RawBatch
    split = synthetic_cifar10(n * b, seed=seed)
    g = torch.Generator().manual_seed(seed)
    out = []
    for i in range(n):
        crop = torch.randint(0, 9, (b, 2), generator=g, dtype=torch.int32)
        flip = (torch.rand(b, generator=g) < 0.5).to(torch.uint8)
        out.append(RawBatch(split.images[i * b:(i + 1) * b], split.labels[i * b:(i + 1) * b], crop, flip))
    return out


@pytest.fixture(scope="module")
def no_tf32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _pair(dtype, graph, b=16, tf32=False):
    """torch fp32 oracle + native engine with the same weights.  ``tf32`` selects, for the fp32 engine, the
    tensor-core kernels (tcgen05 kind::tf32, the default) or the exact-fp32 SIMT kernels (DDPB_TF32=0)."""
    import os
    from cs744_ddp_b200.backends import TorchBackend
    from cs744_ddp_b200.engine import NativeBackend
    torch.manual_seed(0)
    tb = TorchBackend(_cfg(backend="torch", dtype="fp32", global_batch=b), DEV)
    old = os.environ.get("DDPB_TF32")
    os.environ["DDPB_TF32"] = "1" if tf32 else "0"
    try:
        nb = NativeBackend(_cfg(backend="native", dtype=dtype, use_cuda_graph=graph, global_batch=b), DEV,
                           init_state=tb.state_dict())
    finally:
        if old is None:
            os.environ.pop("DDPB_TF32", None)
        else:
            os.environ["DDPB_TF32"] = old
    return tb, nb


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


@pytest.mark.parametrize("tf32", [False, True])
@pytest.mark.parametrize("graph", [False, True])
def test_fp32_engine_matches_torch_training(no_tf32, graph, tf32):
    """fp32 engine against torch in strict fp32.  With the exact SIMT kernels the two differ by summation order
    only; with the tensor-core kernels every convolution reads its operands as TF32 (10-bit mantissa, ~8e-4
    relative per product sum, measured in test_conv_tf32) - the loss bounds below are those of the exact path
    widened by that amount per layer on the way forward and back."""
    tb, nb = _pair("fp32", graph, tf32=tf32)
    tb.set_train(True)
    for step, raw in enumerate(_batches(4, 16)):
        lt = tb.train_step(*tb.prepare(raw), _T())
        ln = nb.train_step(*nb.prepare(raw), _T())
        print(f"[eng] fp32 graph={graph} step {step}: torch {float(lt):.6f} native {float(ln):.6f}")
        # training is chaotic: rounding-level differences grow step by step (torch itself is not
        # run-to-run reproducible at the 1e-3 level by step 3)
        tol = (2e-3 if step < 2 else 3e-2) if not tf32 else (1e-2 if step < 2 else 5e-2)
        if step == 0:       # diagnostic: which tensors differ most after the first update
            sd_t, sd_n = tb.state_dict(), nb.state_dict()
            worst = sorted(((_rel(sd_n[k], sd_t[k]), k) for k in sd_t if "num_batches" not in k), reverse=True)[:3]
            print("[eng] fp32 after step 0, largest relative differences: " + ", ".join(f"{k} {r:.2e}" for r, k in worst))
        assert abs(float(lt) - float(ln)) < tol * max(1.0, abs(float(lt))), step
    sd_t, sd_n = tb.state_dict(), nb.state_dict()
    assert list(sd_t.keys()) == list(sd_n.keys())
    for k in sd_t:
        if "num_batches" in k:
            assert int(sd_t[k]) == int(sd_n[k]) == 4
            continue
        err = _rel(sd_n[k], sd_t[k])
        # conv biases sit in front of a BN: their gradient is numerical noise.  BatchNorm
        # biases start at zero and one borderline ReLU decision (see
        # test_first_update_is_reproducible_across_fresh_engines) moves them by ~10 % after four steps.
        small = k.endswith(".bias") and sd_t[k].numel() > 10 and "layers" in k
        # TF32 perturbs every pre-activation by ~1e-3, which flips the ReLU decision of the elements that close to
        # zero: BatchNorm biases (which start at 0, so their value IS the accumulated gradient) move by tens of per
        # cent, weights by a few per cent after four steps; the per-kernel bound is test_conv_tf32's 2e-3
        tol = (0.2 if small else 2e-2) if not tf32 else (1.0 if small else 0.1)
        assert err < tol, (k, err)


@pytest.mark.parametrize("tf32", [False, True])
def test_fp32_gradients_match_autograd(no_tf32, tf32):
    """Exact-fp32 kernels: gradients equal strict-fp32 autograd up to summation order.  TF32 kernels: judged
    against what torch's OWN TF32 convolutions (cudnn.allow_tf32) achieve on the same batch - the early layers
    are ill-conditioned at initialisation, so any TF32 path is a few per cent off there."""
    import copy
    tb, nb = _pair("fp32", False, tf32=tf32)
    raw = _batches(1, 16, seed=3)[0]
    data, target = tb.prepare(raw)
    tb.optimizer.zero_grad()
    loss = tb.criterion(tb.model(data), target)
    loss.backward()
    ref = {n: p.grad.clone() for n, p in tb.model.named_parameters()}
    theirs = {}
    if tf32:
        m2 = copy.deepcopy(tb.model)
        m2.zero_grad()
        torch.backends.cudnn.allow_tf32 = True
        try:
            tb.criterion(m2(data), target).backward()
        finally:
            torch.backends.cudnn.allow_tf32 = False
        theirs = {n: _rel(p.grad, ref[n]) for n, p in m2.named_parameters()}
    # run only forward + backward on the engine (lr 0 keeps the weights, gradients stay in the arena)
    nb.cfg.lr = 0.0
    from cs744_ddp_b200.ops.native import check, stream_handle
    check(nb._lib.ddpb_engine_set_lr(nb._handle, __import__("ctypes").c_float(0.0)))
    nb.train_step(*nb.prepare(raw), _T())
    got = nb.grads()
    for name, g in ref.items():
        if name.endswith(".bias") and "layers" in name and g.abs().max() < 1e-5:
            continue                                   # conv bias before BN: pure noise
        err = _rel(got[name], g)
        if not tf32:
            print(f"[eng] grad {name}: rel {err:.2e}")
            assert err < 2e-3, (name, err)
        else:
            print(f"[eng] grad tf32 {name}: native {err:.2e}  torch allow_tf32 {theirs[name]:.2e}")
            assert err < max(2e-2, 3.0 * theirs[name]), (name, err, theirs[name])


@pytest.mark.parametrize("graph", [False, True])
def test_bf16_tensor_core_engine_tracks_fp32(no_tf32, graph):
    tb, nb = _pair("bf16", graph, b=32)
    tb.set_train(True)
    lt_hist, ln_hist = [], []
    for step, raw in enumerate(_batches(6, 32, seed=1)):
        lt_hist.append(float(tb.train_step(*tb.prepare(raw), _T())))
        ln_hist.append(float(nb.train_step(*nb.prepare(raw), _Fused() if graph else _T())))
    print("[eng] bf16 losses torch ", ["%.4f" % v for v in lt_hist])
    print("[eng] bf16 losses native", ["%.4f" % v for v in ln_hist])
    assert abs(lt_hist[0] - ln_hist[0]) < 3e-2 * max(1.0, lt_hist[0])
    for a, b in zip(lt_hist, ln_hist):
        assert abs(a - b) < 0.25 * max(1.0, a)
    w_t = tb.state_dict()["layers.8.weight"]
    w_n = nb.state_dict()["layers.8.weight"]
    assert _rel(w_n, w_t) < 5e-2


@pytest.mark.parametrize("parts", [2, 4])
def test_sub_bucket_hand_off_computes_the_same_step(no_tf32, monkeypatch, parts):
    """K-B1 below bucket granularity: with DDPB_SUB_BUCKETS the large layers' weight gradients are launched per
    output-channel range and each range is its own bucket.  Same gradients (first step, lr 0: the ranges fill
    disjoint rows of dW; only the split-K summation order may differ) and the same training trajectory."""
    import ctypes
    from cs744_ddp_b200.engine import NativeBackend
    from cs744_ddp_b200.ops.native import check
    torch.manual_seed(0)
    tb, _ = _pair("bf16", True, b=32)
    mk = lambda: NativeBackend(_cfg(backend="native", dtype="bf16", use_cuda_graph=True, global_batch=32,
                                    bucket_plan="per_layer"), DEV, init_state=tb.state_dict())
    ref, ref2 = mk(), mk()            # two identical engines: their difference is the run-to-run noise floor
    monkeypatch.setenv("DDPB_SUB_BUCKETS", str(parts))
    monkeypatch.setenv("DDPB_SUB_BUCKETS_FORCE", "1")            # one GPU: the plan is normally only cut on 2+ GPUs
    monkeypatch.setenv("DDPB_SUB_BUCKET_MIN", "200000")          # blocks 2..7
    cut = NativeBackend(_cfg(backend="native", dtype="bf16", use_cuda_graph=True, global_batch=32, bucket_plan="per_layer"),
                        DEV, init_state=tb.state_dict())
    n_ref, n_cut = len(ref.buckets()), len(cut.buckets())
    assert n_cut > n_ref, (n_ref, n_cut)
    raws = _batches(4, 32, seed=3)
    for be in (ref, ref2, cut):
        check(be._lib.ddpb_engine_set_lr(be._handle, ctypes.c_float(0.0)))
        be.train_step(*be.prepare(raws[0]), _Fused())
    g_ref, g_ref2, g_cut = ref.grads(), ref2.grads(), cut.grads()
    # The bf16 step is not bit-reproducible: the order of the fp32 atomics behind the BatchNorm statistics moves a mean
    # by an ulp, which flips a few bf16 roundings per layer (forward noise ~1e-4, visible in the 4th digit of the loss),
    # and at random initialisation with batch 32 the gradients are small differences of large per-sample terms
    # (measured with tools/grad_noise.py: the SAME engine on the SAME batch twice differs by 0.6 % in fc1.weight and
    # 10-16 % in the conv weights; the exact-fp32 engine by 3e-6 .. 3e-4).  The cut engine must sit at that noise floor.
    worst, worst_noise = 0.0, 0.0
    for k in g_ref:
        noise, got = _rel(g_ref2[k], g_ref[k]), _rel(g_cut[k], g_ref[k])
        worst, worst_noise = max(worst, got), max(worst_noise, noise)
        assert got < max(1e-4, 3.0 * noise + 1e-3), (k, got, noise)
    print(f"[eng] sub-buckets x{parts}: worst gradient difference to the uncut engine {worst:.2e} (noise floor {worst_noise:.2e})")
    for be in (ref, cut):
        check(be._lib.ddpb_engine_set_lr(be._handle, ctypes.c_float(0.01)))
    for raw in raws[1:]:
        a = float(ref.train_step(*ref.prepare(raw), _Fused()))
        b = float(cut.train_step(*cut.prepare(raw), _Fused()))
        # two bf16 engines drift apart by a few per cent within three updates at this batch size (same noise as
        # above, amplified by training): a wrong half of a weight gradient would show up as tens of per cent
        assert abs(a - b) < 0.1 * max(1.0, abs(a)), (a, b)


def test_bf16_first_step_gradients_close_to_fp32(no_tf32):
    """bf16 gradients vs the fp32 autograd gradients, judged against what torch's own bf16 autocast
    achieves on the same batch (early layers are ill-conditioned at init: any bf16 path is noisy)."""
    import copy
    import ctypes
    from cs744_ddp_b200.ops.native import check
    tb, nb = _pair("bf16", False, b=32)
    raw = _batches(1, 32, seed=5)[0]
    data, target = tb.prepare(raw)
    tb.criterion(tb.model(data), target).backward()
    ac = copy.deepcopy(tb.model)
    ac.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = tb.criterion(ac(data), target)
    loss.backward()
    check(nb._lib.ddpb_engine_set_lr(nb._handle, ctypes.c_float(0.0)))
    nb.train_step(*nb.prepare(raw), _T())
    got = nb.grads()
    cosf = lambda a, b: float(torch.nn.functional.cosine_similarity(a.flatten().float(), b.flatten().float(), dim=0))
    for (name, p), (_, q) in zip(tb.model.named_parameters(), ac.named_parameters()):
        if "weight" not in name:
            continue
        ours, theirs = cosf(got[name], p.grad), cosf(q.grad, p.grad)
        print(f"[eng] bf16 grad {name}: cos(native, fp32) {ours:.5f}  cos(torch autocast, fp32) {theirs:.5f}")
        assert ours > min(0.98, theirs - 0.03), (name, ours, theirs)


def test_eval_matches_torch(no_tf32):
    tb, nb = _pair("fp32", False)
    tb.set_train(False)
    nb.set_train(False)
    raw = _batches(1, 16, seed=7)[0]
    raw.crop_xy = None
    raw.flip = None
    lt, ct = tb.eval_step(*tb.prepare(raw))
    ln, cn = nb.eval_step(*nb.prepare(raw))
    assert abs(float(lt) - float(ln)) < 1e-3 * max(1.0, float(lt))
    assert int(ct) == int(cn)


def test_state_dict_is_the_reference_format(ref_model_module, tmp_path):
    from cs744_ddp_b200.engine import NativeBackend
    from cs744_ddp_b200.utils import checkpoint as ckpt
    torch.manual_seed(0)
    nb = NativeBackend(_cfg(backend="native", dtype="bf16"), DEV)
    nb.train_step(*nb.prepare(_batches(1, 16)[0]), _T())
    path = str(tmp_path / "m.pt")
    ckpt.save_model(path, nb.state_dict())
    ref = ref_model_module.VGG11()
    ref.load_state_dict(torch.load(path), strict=True)            # loads into the reference model
    for k, v in ref.state_dict().items():
        assert v.dtype == nb.state_dict()[k].dtype
    # and back: a fresh engine initialised from that file continues identically
    osd = nb.optimizer_state_dict()
    nb2 = NativeBackend(_cfg(backend="native", dtype="bf16"), DEV, init_state=ckpt.load(path)["model"])
    nb2.load_optimizer_state_dict(osd)
    raw = _batches(1, 16, seed=9)[0]
    l1 = float(nb.train_step(*nb.prepare(raw), _T()))
    l2 = float(nb2.train_step(*nb2.prepare(raw), _T()))
    assert abs(l1 - l2) < 2e-3 * max(1.0, abs(l1))      # atomics + bf16 rounding: not bit-reproducible
    for k, v in nb.state_dict().items():
        if v.dtype.is_floating_point and v.numel() > 1 and not (k.endswith(".bias") and "layers" in k):
            # float atomics + bf16 rounding: runs are not bit-reproducible, and the earliest layers
            # (ill-conditioned gradients) amplify that the most
            assert _rel(nb2.state_dict()[k], v) < 2e-2, k


def test_trainer_end_to_end_native():
    import re
    from cs744_ddp_b200.trainer import Trainer
    from cs744_ddp_b200.utils.logger import RunLogger
    lines = []
    cfg = _cfg(backend="native", dtype="bf16", train_size=16 * 9, test_size=40, max_iters=8)
    t = Trainer(cfg, RunLogger("1", 4, sink=lines.append))
    res = t.fit()
    assert res["total"] == 40
    assert re.match(r"Training loss after 4 epochs is [-\d.e]+$", lines[0])
    assert re.match(r"Forward Pass time in iter 8 is [-\d.e]+$", lines[2])
    assert t.backend.launch_count() > 0
    assert "libddpb200.so" in open("/proc/self/maps").read()


@pytest.mark.parametrize("dtype", ["bf16", "fp32"])
def test_window_loss_is_the_mean_of_the_step_losses_native(dtype):
    """ref Part 1/main.py:45-50.  The engine returns a view of its loss scalar that the next step
    overwrites; the printed window mean must still be the mean of the individual step losses."""
    from cs744_ddp_b200.trainer import Trainer
    from cs744_ddp_b200.utils.logger import RunLogger
    cfg = _cfg(backend="native", dtype=dtype, train_size=16 * 9, test_size=16, max_iters=8)
    t = Trainer(cfg, RunLogger("1", 4, sink=lambda s: None))
    be, seen, step = t.backend, [], t.backend.train_step

    def spy(data, target, timer):
        loss = step(data, target, timer)
        seen.append(be.loss_to_host(loss))            # this step's own value, copied before the next step runs
        return loss

    be.train_step = spy
    t.train_epoch(0)
    vals = [h.item() for h in seen[4:]] if len(seen) > 4 else []
    recs = t.logger.records
    assert len(recs) == 2 and len(seen) == 8
    want = sum(vals) / 4                                # (only 4 pinned slots: the second window is checkable)
    assert abs(recs[1]["loss"] - want) <= 1e-6 * max(1.0, abs(want)), (recs[1]["loss"], want)
    assert len(set(vals)) > 1                           # the steps really had different losses


def test_eval_after_training_matches_torch(no_tf32):
    """Running statistics produced by the engine's own BatchNorm forward (momentum 0.1, unbiased
    variance, ref Part 1/model.py:24) must evaluate like torch's after the same training steps
    (ref test_model, Part 1/main.py:61-76)."""
    tb, nb = _pair("fp32", False, b=32)
    tb.set_train(True)
    for step, raw in enumerate(_batches(20, 32, seed=5)):
        lt = tb.train_step(*tb.prepare(raw), _T())
        ln = nb.train_step(*nb.prepare(raw), _T())
        if step == 0:
            # identical weights went in: the running statistics of the first step must agree tightly
            sd_t, sd_n = tb.state_dict(), nb.state_dict()
            for k in sd_t:
                if k.endswith("running_mean") or k.endswith("running_var"):
                    assert _rel(sd_n[k], sd_t[k]) < 1e-3, (k, _rel(sd_n[k], sd_t[k]))
    print(f"[eng] after 20 steps: train loss torch {float(lt):.4f} native {float(ln):.4f}")
    sd_t, sd_n = tb.state_dict(), nb.state_dict()
    for k in sd_t:
        if k.endswith("running_mean") or k.endswith("running_var"):
            # 20 chaotic steps (lr 0.01, bf16-free but atomics-ordered sums): the two trainings drift apart
            # at the 1e-1 level and the statistics follow the weights; only gross errors are caught here,
            # the tight comparisons are the step-0 check above and the shared-state evaluations below
            assert _rel(sd_n[k], sd_t[k]) < 0.5, (k, _rel(sd_n[k], sd_t[k]))
    # same weights + same statistics in both models: the evaluation itself must agree tightly
    nb.load_state_dict(sd_t)
    tb.set_train(False)
    nb.set_train(False)
    for raw in _batches(2, 32, seed=8):
        raw.crop_xy = None
        raw.flip = None
        le_t, c_t = tb.eval_step(*tb.prepare(raw))
        le_n, c_n = nb.eval_step(*nb.prepare(raw))
        print(f"[eng] eval torch {float(le_t):.5f}/{int(c_t)} native {float(le_n):.5f}/{int(c_n)}")
        assert abs(float(le_t) - float(le_n)) < 2e-3 * max(1.0, float(le_t))
        assert abs(int(c_t) - int(c_n)) <= 1
    # and the engine's OWN trained state evaluates like torch loaded with that state
    tb2, nb2 = _pair("fp32", False, b=32)
    for raw in _batches(20, 32, seed=5):
        nb2.train_step(*nb2.prepare(raw), _T())
    tb2.load_state_dict(nb2.state_dict())
    tb2.set_train(False)
    nb2.set_train(False)
    raw = _batches(1, 32, seed=8)[0]
    raw.crop_xy = None
    raw.flip = None
    le_t, c_t = tb2.eval_step(*tb2.prepare(raw))
    le_n, c_n = nb2.eval_step(*nb2.prepare(raw))
    print(f"[eng] eval of the native-trained state: torch {float(le_t):.5f}/{int(c_t)} native {float(le_n):.5f}/{int(c_n)}")
    assert abs(float(le_t) - float(le_n)) < 2e-3 * max(1.0, float(le_t))


def test_bucket_plans_cover_arena():
    from cs744_ddp_b200.engine import NativeBackend
    for plan, n in (("ddp", 3), ("per_layer", 9), ("single", 1)):
        nb = NativeBackend(_cfg(backend="native", dtype="bf16", bucket_plan=plan, sync="bucketed"), DEV)
        b = nb.buckets()
        assert len(b) == n, (plan, b)
        assert b[0][0] == 0 and all(x[0] + x[1] == y[0] for x, y in zip(b, b[1:]))
        assert b[-1][0] + b[-1][1] == int(nb.info.arena_elems)
        del nb


@pytest.mark.parametrize("variant", ["VGG13", "VGG16"])
def test_other_vgg_variants_run_on_the_native_engine(no_tf32, variant):
    """The reference's table lists VGG13/16/19 too; their 64->64 layers use the tensor-core forward /
    dgrad with a SIMT wgrad (the wgrad tile is 128 output channels wide)."""
    from cs744_ddp_b200.backends import TorchBackend
    from cs744_ddp_b200.engine import NativeBackend
    torch.manual_seed(0)
    tb = TorchBackend(_cfg(backend="torch", dtype="fp32", model=variant, global_batch=8), DEV)
    nb = NativeBackend(_cfg(backend="native", dtype="bf16", model=variant, global_batch=8), DEV,
                       init_state=tb.state_dict())
    raw = _batches(1, 8, seed=11)[0]
    lt = float(tb.train_step(*tb.prepare(raw), _T()))
    ln = float(nb.train_step(*nb.prepare(raw), _T()))
    print(f"[eng] {variant}: torch {lt:.4f} native bf16 {ln:.4f}")
    assert abs(lt - ln) < 5e-2 * max(1.0, lt)
    assert list(nb.state_dict().keys()) == list(tb.state_dict().keys())


def test_profile_step_reports_a_consistent_timeline():
    from cs744_ddp_b200.engine import NativeBackend
    nb = NativeBackend(_cfg(backend="native", dtype="bf16", global_batch=32, sync="bucketed"), DEV)
    raw = _batches(1, 32)[0]
    nb.train_step(*nb.prepare(raw), _T())
    tl = nb.profile_step(*nb.prepare(raw))
    assert 0 < tl["forward_ms"] < tl["backward_compute_end_ms"] <= tl["step_ms"]
    assert 0 <= tl["exposed_comm_frac"] < 0.5
    ready = [b["ready_ms"] for b in tl["buckets"]]
    assert ready == sorted(ready) and len(ready) == len(nb.buckets())
    assert all(b["done_ms"] >= b["ready_ms"] for b in tl["buckets"])


def test_staged_batches_and_async_loss_match_the_direct_path():
    """stage() (copy stream, double-buffered) + loss_to_host() give the same training as prepare() on host batches."""
    from cs744_ddp_b200.data.pipeline import RawBatch
    tb, nb1 = _pair("bf16", True)
    _, nb2 = _pair("bf16", True)
    raws = _batches(4, 16, seed=7)
    pinned = [RawBatch(r.images.pin_memory(), r.labels.pin_memory(), r.crop_xy.pin_memory(), r.flip.pin_memory())
              for r in raws]
    direct = [float(nb1.train_step(*nb1.prepare(r), _Fused())) for r in raws]
    handles, staged = [], nb2.stage(pinned[0])
    for i in range(len(pinned)):
        data, target = nb2.prepare(staged)
        handles.append(nb2.loss_to_host(nb2.train_step(data, target, _Fused())))
        if i + 1 < len(pinned):
            staged = nb2.stage(pinned[i + 1])          # queued while step i is still running
    piped = [h.item() for h in handles]                 # four pinned slots: all four are still valid
    print("[eng] direct", [f"{v:.4f}" for v in direct], "pipelined", [f"{v:.4f}" for v in piped])
    for step, (a, b) in enumerate(zip(direct, piped)):
        # bf16 + atomics: two runs of the same training drift apart step by step (see the bf16 test above)
        tol = 5e-3 if step == 0 else (3e-2 if step == 1 else 0.15)
        assert abs(a - b) < tol * max(1.0, abs(a)), (step, a, b)


def test_first_update_is_reproducible_across_fresh_engines(no_tf32, monkeypatch):
    """Same weights + same batch on fresh engines: the first update must be the same every time up to the
    order of the floating-point atomics.  That order can move a BatchNorm mean by one ulp, and this
    batch has an element whose pre-ReLU value sits within an ulp of zero (channel 302 of block 6): in
    ~10 % of runs its ReLU decision flips in the recomputed backward mask and the gradients of that
    channel - and through dX of everything earlier - change by ~1 %.  Both outcomes are legitimate
    (torch lands on one of them); anything larger than that one flip is a bug."""
    from cs744_ddp_b200.backends import TorchBackend
    from cs744_ddp_b200.engine import NativeBackend
    torch.manual_seed(0)
    tb = TorchBackend(_cfg(backend="torch", dtype="fp32", global_batch=16), DEV)
    init = {k: v.clone() for k, v in tb.state_dict().items()}
    raw = _batches(1, 16)[0]
    ref, flips = None, 0
    monkeypatch.setenv("DDPB_TF32", "0")        # the exact-fp32 kernels: this test is about the order of fp32 atomics
    for it in range(8):
        nb = NativeBackend(_cfg(backend="native", dtype="fp32", use_cuda_graph=False, global_batch=16), DEV,
                           init_state=init)
        nb.train_step(*nb.prepare(raw), _T())
        torch.cuda.synchronize()
        # BatchNorm biases start at zero, so their value after one step IS the (scaled) gradient
        cur = {k: v.float().clone() for k, v in nb.state_dict().items() if k.endswith(".bias") and "layers" in k
               and int(k.split(".")[1]) in (1, 5, 9, 12, 16, 19, 23, 26)}
        if ref is None:
            ref = cur
        else:
            worst = max((_rel(cur[k], ref[k]), k) for k in cur)
            assert worst[0] < 5e-2, (it, worst)
            assert _rel(cur["layers.26.bias"], ref["layers.26.bias"]) < 1e-4     # above the borderline element: exact
            flips += worst[0] > 1e-4
        del nb
    print(f"[eng] {flips} of 7 fresh engines took the other ReLU decision")
