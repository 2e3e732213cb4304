"""Golden-string tests for the log lines and the CLI contract (C24)."""
import pytest

from cs744_ddp_b200.cli import build_parser, config_from_args
from cs744_ddp_b200.config import TrainConfig
from cs744_ddp_b200.utils.logger import RunLogger


def _capture(part):
    lines = []
    return RunLogger(part, 20, sink=lines.append), lines


def test_window_lines_parts_1_2a_2b():
    log, lines = _capture("2a")
    for it in range(1, 41):
        log.step(it, 2.0, 0.5, 1.0, 1.5)
    assert lines == [
        "Training loss after 20 epochs is 2.0",
        "Training loss after 40 epochs is 2.0",
        "Forward Pass time in iter 40 is 0.5",
        "Backward Pass time in iter 40 is 1.0",
        "Average Pass time in iter 40 is 1.5",
    ]


def test_window_lines_part3_says_iterations():
    log, lines = _capture("3")
    for it in range(1, 21):
        log.step(it, 13.13917064666748, 0, 0, 0)
    assert lines == ["Training loss after 20 iterations is 13.13917064666748"]


def test_static_lines():
    log, lines = _capture("2b")
    log.loader_sizes(196, 79)
    log.epoch_time(1, 12.5)
    log.test_result(2.30449, 62, 512)
    assert lines == [
        "Size of training set is 196",
        "Size of test set is 79",
        "Training time after 1 epoch is 12.5",
        "Test set: Average loss: 2.3045, Accuracy: 62/512 (12%)\n",
    ]
    log1, lines1 = _capture("1")
    log1.loader_sizes(196, 40)
    assert lines1 == []          # Part 1 prints no loader sizes


def test_cli_reference_flags_and_alias():
    cfg = config_from_args("2a", ["--master", "10.0.0.1", "--num-nodes", "4", "--rank", "2"])
    assert (cfg.master, cfg.num_nodes, cfg.rank, cfg.epochs) == ("10.0.0.1", 4, 2, 1)
    assert cfg.port == "6585" and cfg.global_batch == 256 and cfg.sync == "gather_scatter"
    assert cfg.per_rank_batch == 64
    cfg = config_from_args("3", ["--master-ip", "h", "--num-nodes", "8", "--rank", "7", "--epochs", "3"])
    assert cfg.master == "h" and cfg.epochs == 3 and cfg.sync == "bucketed" and cfg.per_rank_batch == 32
    assert config_from_args("2b", ["--master", "h", "--num-nodes", "3", "--rank", "0"]).per_rank_batch == 85


def test_cli_required_flags():
    with pytest.raises(SystemExit):
        build_parser("2b").parse_args(["--num-nodes", "2", "--rank", "0"])
    cfg = config_from_args("1", [])           # Part 1 needs no flags
    assert cfg.sync == "none" and cfg.num_nodes == 1


def test_config_defaults_match_reference_constants():
    c = TrainConfig()
    assert (c.lr, c.momentum, c.weight_decay, c.seed, c.cpu_threads, c.log_every) == (0.1, 0.9, 1e-4, 0, 4, 20)
    with pytest.raises(ValueError):
        TrainConfig(sync="ring")
