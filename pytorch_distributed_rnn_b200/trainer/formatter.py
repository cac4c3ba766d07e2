"""Log-line contract of the trainers.

These strings are an interface: the reference's experiment driver stores stdout/stderr and
its analysis notebooks regex the performance line
(``0: Memory Usage: (\\d+\\.\\d+), Training Duration: (\\d+\\.\\d+)`` —
evaluation/Experiments.ipynb:49), so every message here renders byte-for-byte like
src/motion/trainer/formatter.py:6-27.
"""
from __future__ import annotations

import re
from typing import Dict, Iterator, Optional


def percentage(current, overall) -> float:
    return 100.0 * (current / overall)


class TrainingMessageFormatter:
    RANK = "Rank: {:02d}   "

    def __init__(self, num_epochs: int, rank: int = 0):
        self.num_epochs = num_epochs
        self.rank = rank

    def epoch_start_message(self, epoch: int) -> str:
        return (self.RANK + "Start Epoch {}").format(self.rank, epoch)

    def train_progress_message(self, batch_idx: int, batches: int, training_examples: int, correct, loss: float) -> str:
        done = batch_idx + 1
        head = (self.RANK + "Train Batch: {}/{} ({:.0f}%)").format(self.rank, done, batches, percentage(done, batches))
        body = "\tLoss: {:.6f}\tAcc: {}/{} ({:.0f}%)".format(
            loss, correct, training_examples, percentage(float(correct), training_examples))
        return head + body

    def evaluation_message(self, accuracy: float, examples: int, epoch: Optional[int], eval_loss: float, total_correct) -> str:
        if epoch is None:
            prefix = "Test Evaluation:\t"
        else:
            prefix = "Evaluation Epoch: {}/{} ({:.0f}%)\t".format(
                epoch + 1, self.num_epochs, percentage(epoch + 1, self.num_epochs))
        return prefix + "Loss: {:.4f}\t Accuracy: {}/{} ({:.0f}%)\n".format(
            eval_loss, total_correct, examples, 100.0 * accuracy)

    def performance_message(self, memory, duration) -> str:
        return f"{self.rank}: Memory Usage: {memory}, Training Duration: {duration}"


# ---------------------------------------------------------------------------
# The inverse direction: parsers for the lines above.  The reference only ever parses the performance line
# (one regex inside a notebook); here every message has a parser, so logs of either implementation can be
# turned back into numbers (bench/analysis.py, tests/test_formatter.py round-trips) without copying regexes.
# ---------------------------------------------------------------------------
_RANK = r"Rank: (?P<rank>\d+)   "
PATTERNS: Dict[str, "re.Pattern[str]"] = {
    "epoch_start": re.compile(_RANK + r"Start Epoch (?P<epoch>\d+)"),
    "train_progress": re.compile(
        _RANK + r"Train Batch: (?P<batch>\d+)/(?P<batches>\d+) \((?P<pct>\d+)%\)\tLoss: (?P<loss>-?[\d.]+(?:e-?\d+)?|nan|inf)"
        r"\tAcc: (?P<correct>\d+)/(?P<examples>\d+) \((?P<acc_pct>\d+)%\)"),
    "evaluation": re.compile(
        r"(?:Test Evaluation:|Evaluation Epoch: (?P<epoch>\d+)/(?P<epochs>\d+) \((?P<pct>\d+)%\))\t"
        r"Loss: (?P<loss>-?[\d.]+|nan|inf)\t Accuracy: (?P<correct>\d+)/(?P<examples>\d+) \((?P<acc_pct>\d+)%\)"),
    "performance": re.compile(r"(?P<rank>\d+): Memory Usage: (?P<memory>\d+\.\d+), Training Duration: (?P<duration>\d+\.\d+)"),
}
_INT = {"rank", "epoch", "epochs", "batch", "batches", "pct", "correct", "examples", "acc_pct"}


def parse_line(line: str) -> Optional[dict]:
    """Return ``{"kind": ..., <fields as int / float>}`` for a trainer log line, or None if it is something else."""
    for kind, pat in PATTERNS.items():
        m = pat.search(line)
        if m is None:
            continue
        out = {"kind": kind}
        for k, v in m.groupdict().items():
            if v is not None:
                out[k] = int(v) if k in _INT else float(v)
        return out
    return None


def parse_log(text: str) -> Iterator[dict]:
    """All recognised trainer lines of a captured stdout / stderr, in order."""
    for line in text.splitlines():
        rec = parse_line(line)
        if rec is not None:
            yield rec
