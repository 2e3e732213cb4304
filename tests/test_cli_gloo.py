"""The main.py-style entry points, started the way the reference is (one process per node, by hand),
over gloo / loopback on CPU ("Gloo over loopback is the de facto fake cluster")."""
import os
import re
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _launch(part, rank, world, port, extra=()):
    cmd = [sys.executable, "-m", f"cs744_ddp_b200.parts.part{part}", "--master", "127.0.0.1", "--num-nodes", str(world),
           "--rank", str(rank), "--synthetic", "--train-size", "96", "--test-size", "16", "--batch-size", "8",
           "--device", "cpu", "--backend", "torch", "--port", str(port), *extra]
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    return subprocess.Popen(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)


@pytest.mark.parametrize("part,word", [("2a", "epochs"), ("2b", "epochs"), ("3", "iterations")])
def test_two_rank_cli_run_prints_reference_lines(part, word):
    port = _free_port()
    procs = [_launch(part, r, 2, port) for r in range(2)]
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (out, err) in zip(procs, outs):
        assert p.returncode == 0, err[-2000:]
        lines = out.splitlines()
        assert lines[0] == "Size of training set is 12"         # 96 / 2 ranks / batch 4
        assert lines[1] == "Size of test set is 4"
        assert any(re.match(r"Training time after 1 epoch is [\d.e-]+$", l) for l in lines)
        assert any(re.match(r"Test set: Average loss: \d+\.\d{4}, Accuracy: \d+/16 \(\d+%\)$", l) for l in lines)
    # both ranks evaluate the full test set on identical weights (Part 3 also shares BN statistics)
    if part == "3":
        t0 = [l for l in outs[0][0].splitlines() if l.startswith("Test set")]
        t1 = [l for l in outs[1][0].splitlines() if l.startswith("Test set")]
        assert t0 == t1


def test_two_rank_run_writes_one_checkpoint(tmp_path):
    """Saving is a collective every rank takes part in (sharded optimizer state on the GPU paths);
    rank 0 alone writes the file, and no rank hangs."""
    ck = tmp_path / "p3.pt"
    port = _free_port()
    procs = [_launch("3", r, 2, port, ("--checkpoint", str(ck))) for r in range(2)]
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (out, err) in zip(procs, outs):
        assert p.returncode == 0, err[-2000:]
    import torch
    blob = torch.load(ck, weights_only=True)
    assert len(blob["model"]) == 58 and blob["epoch"] == 1 and len(blob["optimizer"]["state"]) == 34


def test_part1_cli_needs_no_flags_and_checkpoints(tmp_path):
    ck = tmp_path / "p1.pt"
    cmd = [sys.executable, "-m", "cs744_ddp_b200.parts.part1", "--synthetic", "--train-size", "64", "--test-size", "16",
           "--batch-size", "16", "--device", "cpu", "--backend", "torch", "--checkpoint", str(ck)]
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "Size of training set" not in res.stdout              # Part 1 prints no loader sizes
    assert "Test set: Average loss" in res.stdout
    import torch
    blob = torch.load(ck, weights_only=False)
    assert len(blob["model"]) == 58 and "optimizer" in blob
