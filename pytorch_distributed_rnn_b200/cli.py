"""Command line — same flags, defaults and sub-commands as src/motion/main.py:15-43.

    python -m pytorch_distributed_rnn_b200 [global flags] {local,distributed,horovod,parameter-server} [...]

Global flags must precede the sub-command, exactly like the reference (see any recorded
command in results_baseline_1.json).  ``--dropout`` and ``--num-threads`` are parsed and
ignored the way the reference ignores them (quirks Q1, Q3); ``--validation-fraction`` is
honoured (Q2).  Extensions, all optional: ``--cell``, ``--device``, ``--deferred-logging``,
``--save-every``, ``--resume``, ``--history-path``, ``--synthetic``.
"""
from __future__ import annotations

import argparse
import logging
import sys
from pathlib import Path

import torch

PKG_DIR = Path(__file__).absolute().parent
DEFAULT_CHECKPOINT_DIR = PKG_DIR / "models_ckpt"
DEFAULT_DATASET_PATH = PKG_DIR / "data_cache"


# The reference's global flags (name, default, type) — an interface: recorded experiment commands must replay.
REFERENCE_FLAGS = (
    ("checkpoint-directory", DEFAULT_CHECKPOINT_DIR, Path),
    ("dataset-path", DEFAULT_DATASET_PATH, Path),
    ("output-path", None, Path),
    ("stacked-layer", 2, int),
    ("hidden-units", 32, int),
    ("epochs", 100, int),
    ("validation-fraction", 0.1, float),
    ("batch-size", 1440, int),
    ("learning-rate", 0.0025, float),
    ("dropout", 0.1, float),          # parsed, unused (quirk Q1)
    ("log", "INFO", str),
    ("num-threads", 4, int),          # parsed, unused (quirk Q3)
    ("seed", None, int),
)
# Extensions (name, argparse keyword arguments)
EXTENSION_FLAGS = (
    ("cell", dict(default="lstm", choices=["lstm", "gru"])),
    ("device", dict(default=None, help="cpu | cuda | cuda:N (default: this rank's GPU if any)")),
    ("deferred-logging", dict(action="store_true",
                              help="render per-batch lines once per epoch from the device ring buffer (no per-step host sync)")),
    ("save-every", dict(default=0, type=int, help="also write checkpoint-epoch-N.pt every N epochs")),
    ("resume", dict(default=None, type=Path, help="checkpoint to resume from")),
    ("history-path", dict(default=None, type=Path)),
    ("synthetic", dict(action="store_true",
                       help="write a synthetic dataset of the reference's shape to --dataset-path if none is there")),
    ("delay-sweep-us", dict(default=None, type=str,
                            help="network experiment in one process: comma list of delays (us) injected before every flag "
                                 "publication of the in-kernel collectives; the run is repeated once per value after an "
                                 "untimed warm-up, each with its own performance line (fabfile.py:130-191 ran one job per rule)")),
    ("loss-sweep-pct", dict(default=None, type=str,
                            help="same experiment for message loss: comma list of percentages; a publication is dropped "
                                 "with that probability and pays a retransmission timeout (B200RNN_LOSS_RTO_NS, default "
                                 "200 us) before it goes out (the reference's `tc netem loss N%%` rule)")),
)


def build_parser() -> argparse.ArgumentParser:
    from . import trainer
    from .parallel import param_server

    parser = argparse.ArgumentParser(prog="pytorch_distributed_rnn_b200", description="B200-native distributed RNN training")
    for name, default, kind in REFERENCE_FLAGS:
        parser.add_argument(f"--{name}", default=default, type=kind)
    parser.add_argument("--no-validation", action="store_true")
    for name, kwargs in EXTENSION_FLAGS:
        parser.add_argument(f"--{name}", **kwargs)
    commands = parser.add_subparsers(title="Available commands", metavar="command [options ...]")
    commands.required = True
    param_server.add_sub_command(commands)
    trainer.add_sub_commands(commands)
    return parser


def _ensure_synthetic_dataset(args) -> None:
    """``--synthetic``: rank 0 writes the dataset if it is missing (atomically, marker file last); the other ranks
    wait for the marker and fail loudly when it does not appear."""
    import time
    from .data import synthetic
    from .data.dataset import _TensorCache
    from .parallel.process_group import env_rank_info
    if synthetic.is_complete(args.dataset_path):
        return
    if _TensorCache(args.dataset_path).complete() and not any(Path(args.dataset_path).glob(".*.pt.tmp*")):
        return   # a dataset written by someone else (e.g. the real preprocessed UCI-HAR cache): use it as it is
    rank, world, _ = env_rank_info()
    if rank == 0:
        synthetic.write_dataset(args.dataset_path, seed=args.seed or 0)
        return
    deadline = time.monotonic() + 120.0
    while not synthetic.is_complete(args.dataset_path):
        if time.monotonic() > deadline:
            raise RuntimeError(f"--synthetic: rank 0 did not finish writing {args.dataset_path} within 120 s")
        time.sleep(0.1)


def main(argv=None):
    args = build_parser().parse_args(argv)
    if args.seed is not None:
        torch.manual_seed(args.seed)
    if args.synthetic:
        _ensure_synthetic_dataset(args)
    return args.func(args)


if __name__ == "__main__":
    sys.dont_write_bytecode = True
    main()
