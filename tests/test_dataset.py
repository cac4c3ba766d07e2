"""Preprocessor + cache on a tiny fake UCI-HAR tree (parity: src/motion/{processor,dataset}.py)."""
import numpy as np
import torch

from pytorch_distributed_rnn_b200.data import MotionDataProcessor, MotionDataset, synthetic


def _fake_uci(root, n_train=250, n_test=17, T=128):
    rng = np.random.RandomState(0)
    for split, n in (("train", n_train), ("test", n_test)):
        d = root / split / "Inertial Signals"
        d.mkdir(parents=True)
        for s, sig in enumerate(MotionDataProcessor.INPUT_SIGNAL_TYPES):
            arr = rng.randn(n, T).astype(np.float32) + s
            with open(d / f"{sig}{split}.txt", "w") as f:
                for row in arr:
                    f.write("  " + "  ".join(f"{v: .7e}" for v in row) + "\n")  # UCI-HAR's double-space layout
        with open(root / split / f"y_{split}.txt", "w") as f:
            for v in rng.randint(1, 7, n):
                f.write(f"{v}\n")


def test_preprocess_shapes_split_and_cache(tmp_path):
    _fake_uci(tmp_path)
    train, valid, test = MotionDataset.load(tmp_path, validation_fraction=0.1, seed=3)
    assert train.features.shape == (192, 128, 9) and train.features.dtype == torch.float32  # 225 -> multiple of 96
    assert train.labels.shape == (192, 1) and train.labels.dtype == torch.int64
    assert valid.features.shape[0] == 25 and test.features.shape[0] == 17
    assert int(train.labels.min()) >= 0 and int(train.labels.max()) <= 5
    assert train.seq_length == 128 and train.num_features == 9
    # channel s was generated with mean s: transpose (signal, N, T) -> (N, T, signal) is right
    assert abs(float(test.features[:, :, 4].mean()) - 4.0) < 0.2
    for name in ("X_train", "y_train", "X_validation", "y_validation", "X_test", "y_test"):
        assert (tmp_path / f"{name}.pt").exists()
    again = MotionDataset.load(tmp_path)  # second load hits the cache
    assert torch.equal(again[0].features, train.features)


def test_synthetic_layout(tmp_path):
    synthetic.write_dataset(tmp_path, train=96, valid=8, test=8)
    tr, va, te = MotionDataset.load(tmp_path)
    assert tr.features.shape == (96, 128, 9) and tr.labels.shape == (96, 1)
    assert len(MotionDataset.LABELS) == 6 and synthetic.TRAIN_SIZE == 6912
