"""The reference's log lines, byte for byte, plus an optional JSON line.

Lines and where the reference prints them:

* ``Size of training set is {}`` / ``Size of test set is {}``      ref ``Part 2a/main.py:46,55``
* ``Training loss after {} epochs is {}``   (Parts 1/2a/2b)        ref ``Part 1/main.py:49``
* ``Training loss after {} iterations is {}``   (Part 3)           ref ``Part 3/main.py:105``
* ``Forward|Backward|Average Pass time in iter {} is {}``          ref ``Part 1/main.py:52-54``
* ``Training time after {} epoch is {}``                           ref ``Part 1/main.py:120``
* ``Test set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\\n`` ref ``Part 1/main.py:74-76``

Every ``log_every`` (20) iterations the mean loss of the window is printed; the three timings
are printed from the second window on (the first is treated as warm-up and discarded).
"""
from __future__ import annotations

import json
import sys
from typing import Callable, Dict, List, Optional


def _fmt(v) -> str:
    # the reference formats a 0-dim tensor, which prints as a plain python float
    return "{}".format(float(v))


class RunLogger:
    def __init__(self, part: str = "1", log_every: int = 20, rank: int = 0,
                 sink: Optional[Callable[[str], None]] = None, json_lines: bool = False,
                 prefix_rank: bool = False) -> None:
        self.part = part
        self.log_every = int(log_every)
        self.rank = rank
        self.json_lines = json_lines
        self.prefix = f"[rank {rank}] " if prefix_rank else ""
        self._sink = sink or self._stdout
        self.records: List[Dict] = []
        self.reset_window()

    @staticmethod
    def _stdout(line: str) -> None:
        sys.stdout.write(line + "\n")
        sys.stdout.flush()

    def emit(self, line: str) -> None:
        self._sink(self.prefix + line)

    # ---- static lines ---------------------------------------------------------------
    def loader_sizes(self, train_batches: int, test_batches: int) -> None:
        """Printed by the distributed parts only (Part 1 prints neither)."""
        if self.part != "1":
            self.emit("Size of training set is {}".format(train_batches))
            self.emit("Size of test set is {}".format(test_batches))

    def train_loader_size(self, n: int) -> None:
        if self.part != "1":
            self.emit("Size of training set is {}".format(n))

    def test_loader_size(self, n: int) -> None:
        if self.part != "1":
            self.emit("Size of test set is {}".format(n))

    def epoch_time(self, epoch_1based: int, seconds: float) -> None:
        self.emit("Training time after {} epoch is {}".format(epoch_1based, seconds))

    def test_result(self, avg_loss: float, correct: int, total: int) -> None:
        self.emit("Test set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n".format(
            float(avg_loss), int(correct), int(total), 100.0 * correct / total))

    # ---- the 20-iteration window -----------------------------------------------------
    def reset_window(self) -> None:
        self._loss = 0.0
        self._fwd = 0.0
        self._bwd = 0.0
        self._tot = 0.0

    def window(self, iter_number: int, mean_loss: float, fwd_mean: float, bwd_mean: float,
               tot_mean: float, extra: Optional[Dict] = None) -> None:
        """Print one closed window (``iter_number`` is a multiple of ``log_every``)."""
        word = "iterations" if self.part == "3" else "epochs"
        self.emit("Training loss after {} {} is {}".format(iter_number, word, _fmt(mean_loss)))
        rec = {"iter": iter_number, "loss": float(mean_loss)}
        if iter_number != self.log_every:
            self.emit("Forward Pass time in iter {} is {}".format(iter_number, fwd_mean))
            self.emit("Backward Pass time in iter {} is {}".format(iter_number, bwd_mean))
            self.emit("Average Pass time in iter {} is {}".format(iter_number, tot_mean))
            rec.update(forward_s=fwd_mean, backward_s=bwd_mean, total_s=tot_mean)
        if extra:
            rec.update(extra)
        self.records.append(rec)
        if self.json_lines:
            self.emit(json.dumps({"window": rec, "rank": self.rank}))

    def step(self, iter_number: int, loss: float, forward_s: float, backward_s: float,
             total_s: float, extra: Optional[Dict] = None) -> bool:
        """Accumulate one iteration (1-based ``iter_number``); returns True when a window closed."""
        self._loss += float(loss)
        self._fwd += forward_s
        self._bwd += backward_s
        self._tot += total_s
        if iter_number % self.log_every != 0:
            return False
        n = float(self.log_every)
        self.window(iter_number, self._loss / n, self._fwd / n, self._bwd / n, self._tot / n, extra)
        self.reset_window()
        return True
