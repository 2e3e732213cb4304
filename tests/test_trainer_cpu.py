"""End-to-end on CPU: Part 1 plumbing, log lines, checkpoint/resume (BASELINE config 1)."""
import re

import torch

from cs744_ddp_b200.config import TrainConfig
from cs744_ddp_b200.trainer import Trainer
from cs744_ddp_b200.utils.logger import RunLogger


def _cfg(**kw):
    base = dict(synthetic=True, train_size=96, test_size=32, global_batch=8, device="cpu",
                backend="torch", log_every=4, max_iters=8, cpu_threads=2)
    base.update(kw)
    return TrainConfig(**base)


def test_part1_runs_and_prints_reference_lines():
    lines = []
    cfg = _cfg()
    t = Trainer(cfg, RunLogger("1", cfg.log_every, sink=lines.append))
    res = t.fit()
    assert res["total"] == 32
    pat = [r"Training loss after 4 epochs is [-\d.e]+$",
           r"Training loss after 8 epochs is [-\d.e]+$",
           r"Forward Pass time in iter 8 is [-\d.e]+$",
           r"Backward Pass time in iter 8 is [-\d.e]+$",
           r"Average Pass time in iter 8 is [-\d.e]+$",
           r"Training time after 1 epoch is [-\d.e]+$",
           r"Test set: Average loss: \d+\.\d{4}, Accuracy: \d+/32 \(\d+%\)\n$"]
    assert len(lines) == len(pat)
    for line, p in zip(lines, pat):
        assert re.match(p, line), (line, p)


def test_training_reduces_loss_on_learnable_data():
    cfg = _cfg(train_size=64, global_batch=16, max_iters=None, epochs=1, log_every=2)
    t = Trainer(cfg, RunLogger("1", 2, sink=lambda s: None))
    first = None
    for epoch in range(6):
        recs = t.train_epoch(epoch)
    losses = [r["loss"] for r in t.logger.records]
    assert losses[-1] < losses[0]


def test_checkpoint_resume_continues_identically(tmp_path):
    path = str(tmp_path / "ck.pt")
    quiet = lambda s: None
    a = Trainer(_cfg(max_iters=4, checkpoint=path), RunLogger("1", 4, sink=quiet))
    a.fit()
    # continue 2 more iterations from memory vs. from the file
    b = Trainer(_cfg(max_iters=4, resume=path), RunLogger("1", 4, sink=quiet))
    for k, v in a.backend.state_dict().items():
        assert torch.equal(v, b.backend.state_dict()[k]), k
    sa, sb = a.backend.optimizer_state_dict(), b.backend.optimizer_state_dict()
    for i in sa["state"]:
        assert torch.equal(sa["state"][i]["momentum_buffer"], sb["state"][i]["momentum_buffer"])
    assert b.global_iteration == 4


def test_resume_continues_with_the_next_epoch(tmp_path):
    """The checkpoint stores the number of COMPLETED epochs; --epochs stays the total."""
    path = str(tmp_path / "ck.pt")
    lines = []
    a = Trainer(_cfg(max_iters=2, epochs=1, checkpoint=path), RunLogger("1", 4, sink=lambda s: None))
    a.fit()
    b = Trainer(_cfg(max_iters=2, epochs=3, resume=path), RunLogger("1", 4, sink=lines.append))
    assert b.start_epoch == 1 and b.global_iteration == 2
    b.fit()
    done = [l for l in lines if l.startswith("Training time after")]
    assert [l.split()[3] for l in done] == ["2", "3"]            # epochs 2 and 3 ran, epoch 1 did not
    assert b.global_iteration == 6
    c = Trainer(_cfg(max_iters=2, epochs=1, resume=path), RunLogger("1", 4, sink=lambda s: None))
    assert c.fit() == {}                                           # nothing left to train


def test_checkpoint_loader_refuses_pickled_code(tmp_path):
    """Foreign checkpoints are read with the restricted unpickler: a file that smuggles a callable fails."""
    import pickle

    import pytest

    from cs744_ddp_b200.utils import checkpoint as ckpt

    class Boom:
        def __reduce__(self):
            return (print, ("arbitrary code ran",))

    bad = tmp_path / "bad.pt"
    torch.save({"layers.0.weight": Boom()}, bad)
    with pytest.raises(pickle.UnpicklingError):
        ckpt.load(str(bad))


def test_window_loss_is_the_mean_of_the_step_losses():
    """ref Part 1/main.py:45-50: the printed value is sum(loss of each of the 20 iterations) / 20."""
    cfg = _cfg(max_iters=8)
    t = Trainer(cfg, RunLogger("1", cfg.log_every, sink=lambda s: None))
    seen = []
    step = t.backend.train_step

    def spy(data, target, timer):
        loss = step(data, target, timer)
        seen.append(float(loss))
        return loss

    t.backend.train_step = spy
    t.train_epoch(0)
    recs = t.logger.records
    for w in range(2):
        want = sum(seen[4 * w: 4 * w + 4]) / 4
        assert abs(recs[w]["loss"] - want) <= 1e-6 * max(1.0, abs(want)), (w, recs[w]["loss"], want)


def test_eval_mode_quirk_flag():
    quiet = lambda s: None
    t = Trainer(_cfg(epochs=2, max_iters=2, reproduce_eval_mode_quirk=True), RunLogger("1", 4, sink=quiet))
    t.fit()
    assert not t.backend.model.training
    t2 = Trainer(_cfg(epochs=1, max_iters=2), RunLogger("1", 4, sink=quiet))
    t2.train_epoch(0)
    assert t2.backend.model.training
