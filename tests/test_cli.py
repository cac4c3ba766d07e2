"""Flag system parity with src/motion/main.py:17-36 and param_server/__init__.py:13-35."""
from pathlib import Path

import pytest

from pytorch_distributed_rnn_b200 import cli


def test_defaults_match_reference():
    a = cli.build_parser().parse_args(["local"])
    assert (a.stacked_layer, a.hidden_units, a.epochs, a.batch_size) == (2, 32, 100, 1440)
    assert (a.validation_fraction, a.learning_rate, a.dropout, a.log, a.num_threads) == (0.1, 0.0025, 0.1, "INFO", 4)
    assert a.seed is None and a.no_validation is False and a.output_path is None
    assert isinstance(a.checkpoint_directory, Path) and isinstance(a.dataset_path, Path)


def test_subcommands_and_ps_flags():
    p = cli.build_parser()
    for name in ("local", "distributed", "horovod"):
        assert callable(p.parse_args(["--epochs", "1", "--seed", "123456789", "--no-validation", name]).func)
    a = p.parse_args(["parameter-server", "--world-size", "3", "--rank", "1"])
    assert (a.world_size, a.rank, a.master_address, a.master_port) == (3, 1, "localhost", "29500")
    with pytest.raises(SystemExit):
        p.parse_args([])  # sub-command is required
    with pytest.raises(SystemExit):
        p.parse_args(["parameter-server"])  # --world-size / --rank are required


def test_local_run_resume_and_periodic_checkpoint(tiny_dataset_dir, tmp_path):
    common = ["--dataset-path", str(tiny_dataset_dir), "--batch-size", "96", "--seed", "7", "--device", "cpu",
              "--checkpoint-directory", str(tmp_path / "ck"), "--history-path", str(tmp_path / "h.json")]
    t = cli.main([*common, "--epochs", "1", "--save-every", "1", "local"])
    assert (tmp_path / "ck" / "best-model.pt").exists() and (tmp_path / "ck" / "checkpoint-epoch-1.pt").exists()
    t2 = cli.main([*common, "--epochs", "2", "--resume", str(tmp_path / "ck" / "checkpoint-epoch-1.pt"), "local"])
    assert t2.start_epoch == 1 and t2.optimizer.step_count == 2 * t.optimizer.step_count


def test_gru_cell_trains_and_checkpoints_load_into_stock_gru(tiny_dataset_dir, tmp_path):
    """`--cell gru` (an extension: the reference only builds nn.LSTM) keeps the checkpoint key names and loads
    into a stock torch GRU + Linear."""
    import torch
    from torch import nn
    t = cli.main(["--dataset-path", str(tiny_dataset_dir), "--batch-size", "96", "--seed", "7", "--device", "cpu",
                  "--cell", "gru", "--epochs", "1", "--checkpoint-directory", str(tmp_path / "ck"),
                  "--history-path", str(tmp_path / "h.json"), "local"])
    assert t.model.cell == "gru"
    ck = torch.load(tmp_path / "ck" / "best-model.pt", map_location="cpu", weights_only=False)
    state = ck["model_state"]
    assert set(state) == {f"lstm.{w}_l{l}" for w in ("weight_ih", "weight_hh", "bias_ih", "bias_hh") for l in range(2)} | {"fc.weight", "fc.bias"}
    gru = nn.GRU(9, 32, 2, batch_first=True)
    gru.load_state_dict({k[len("lstm."):]: v for k, v in state.items() if k.startswith("lstm.")})
    x = torch.randn(4, 128, 9)
    with torch.no_grad():
        want = gru(x)[0][:, -1] @ state["fc.weight"].t() + state["fc.bias"]
        got = t.model(x)
    assert torch.allclose(got, want, atol=1e-5)


def test_network_sweep_sequences_rules_and_resets_the_communicator(caplog):
    """`--delay-sweep-us` / `--loss-sweep-pct` (fabfile.py:146-191 in one job): a marked warm-up run, then one run per
    rule with exactly that rule active, and a clean communicator afterwards."""
    import logging
    import types
    from pytorch_distributed_rnn_b200 import trainer as tr

    class Comm:
        def __init__(self):
            self.delay, self.loss, self.log = 0, 0.0, []

        def set_delay_ns(self, ns):
            self.delay = ns

        def set_loss(self, pct, rto_ns=0):
            self.loss = pct

    class T:
        start_epoch = 7

        def __init__(self):
            self.comm, self.runs = Comm(), []

        def _get_comm(self):
            return self.comm

        def train(self, epochs):
            self.runs.append((self.comm.delay, self.comm.loss, self.start_epoch))

    t = T()
    with caplog.at_level(logging.INFO):
        tr._network_sweep(types.SimpleNamespace(epochs=1), t, delays_us=[0, 50], losses_pct=[0.5, 15])
    assert t.runs == [(0, 0.0, 7), (0, 0.0, 0), (50000, 0.0, 0), (0, 0.5, 0), (0, 15, 0)]
    assert (t.comm.delay, t.comm.loss) == (0, 0.0)
    marks = [r.getMessage() for r in caplog.records if r.getMessage().startswith("Injected")]
    assert marks == ["Injected warmup: 0 us", "Injected delay: 0 us", "Injected delay: 50 us", "Injected loss: 0.5 %", "Injected loss: 15 %"]

    class NoComm(T):
        def _get_comm(self):
            return None
    with pytest.raises(SystemExit):
        tr._network_sweep(types.SimpleNamespace(epochs=1), NoComm(), delays_us=[1])
