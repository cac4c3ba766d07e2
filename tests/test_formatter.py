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


def test_every_message_round_trips_through_its_parser():
    from pytorch_distributed_rnn_b200.trainer.formatter import parse_line, parse_log
    f = TrainingMessageFormatter(5, rank=3)
    assert parse_line(f.epoch_start_message(2)) == {"kind": "epoch_start", "rank": 3, "epoch": 2}
    rec = parse_line(f.train_progress_message(batch_idx=1, batches=5, training_examples=1440, correct=246, loss=1.794715))
    assert rec == {"kind": "train_progress", "rank": 3, "batch": 2, "batches": 5, "pct": 40, "loss": 1.794715,
                   "correct": 246, "examples": 1440, "acc_pct": 17}
    ev = parse_line(f.evaluation_message(0.25, 400, 1, 1.5, 100))
    assert ev["kind"] == "evaluation" and ev["epoch"] == 2 and ev["epochs"] == 5 and ev["correct"] == 100 and ev["loss"] == 1.5
    tv = parse_line(f.evaluation_message(0.5, 10, None, 0.75, 5))
    assert tv["kind"] == "evaluation" and "epoch" not in tv and tv["examples"] == 10
    pf = parse_line(f.performance_message(123.456, 3.595))
    assert pf == {"kind": "performance", "rank": 3, "memory": 123.456, "duration": 3.595}
    text = "\n".join(["noise", f.epoch_start_message(0), "INFO:root:" + f.performance_message(1.5, 2.5)])
    assert [r["kind"] for r in parse_log(text)] == ["epoch_start", "performance"]
    assert parse_line("unrelated line") is None
