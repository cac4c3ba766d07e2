"""Log-line goldens (format contract of src/motion/trainer/formatter.py and the notebooks' regex)."""
import re

from pytorch_distributed_rnn_b200.trainer.formatter import TrainingMessageFormatter, percentage


def test_epoch_and_progress_lines():
    f = TrainingMessageFormatter(5, rank=3)
    assert f.epoch_start_message(0) == "Rank: 03   Start Epoch 0"
    assert (f.train_progress_message(0, 5, 1440, 246, 1.7947154)
            == "Rank: 03   Train Batch: 1/5 (20%)\tLoss: 1.794715\tAcc: 246/1440 (17%)")


def test_evaluation_lines():
    f = TrainingMessageFormatter(4)
    assert f.evaluation_message(0.25, 100, None, 1.23456, 25) == "Test Evaluation:\tLoss: 1.2346\t Accuracy: 25/100 (25%)\n"
    assert (f.evaluation_message(0.5, 10, 1, 0.5, 5)
            == "Evaluation Epoch: 2/4 (50%)\tLoss: 0.5000\t Accuracy: 5/10 (50%)\n")


def test_performance_line_matches_notebook_regex():
    # evaluation/Experiments.ipynb:49
    line = TrainingMessageFormatter(1, rank=0).performance_message(385.12, 165.37)
    m = re.search(r"0: Memory Usage: (\d+\.\d+), Training Duration: (\d+\.\d+)", line)
    assert m and float(m.group(1)) == 385.12 and float(m.group(2)) == 165.37
    assert percentage(1, 4) == 25.0
