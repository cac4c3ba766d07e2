"""Log-line contract of the trainers.

These strings are an interface: the reference's experiment driver stores stdout/stderr and
its analysis notebooks regex the performance line
(``0: Memory Usage: (\\d+\\.\\d+), Training Duration: (\\d+\\.\\d+)`` —
evaluation/Experiments.ipynb:49), so every message here renders byte-for-byte like
src/motion/trainer/formatter.py:6-27.
"""
from __future__ import annotations

from typing import Optional


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
