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


def build_parser() -> argparse.ArgumentParser:
    from . import trainer
    from .parallel import param_server

    parser = argparse.ArgumentParser(prog="pytorch_distributed_rnn_b200", description="B200-native distributed RNN training")
    parser.add_argument("--checkpoint-directory", default=DEFAULT_CHECKPOINT_DIR, type=Path)
    parser.add_argument("--dataset-path", default=DEFAULT_DATASET_PATH, type=Path)
    parser.add_argument("--output-path", default=None, type=Path)
    parser.add_argument("--stacked-layer", default=2, type=int)
    parser.add_argument("--hidden-units", default=32, type=int)
    parser.add_argument("--epochs", default=100, type=int)
    parser.add_argument("--validation-fraction", default=0.1, type=float)
    parser.add_argument("--batch-size", default=1440, type=int)
    parser.add_argument("--learning-rate", default=0.0025, type=float)
    parser.add_argument("--dropout", default=0.1, type=float)
    parser.add_argument("--log", default="INFO")
    parser.add_argument("--num-threads", default=4, type=int)
    parser.add_argument("--seed", default=None, type=int)
    parser.add_argument("--no-validation", action="store_true")
    # extensions
    parser.add_argument("--cell", default="lstm", choices=["lstm", "gru"])
    parser.add_argument("--device", default=None, help="cpu | cuda | cuda:N (default: this rank's GPU if any)")
    parser.add_argument("--deferred-logging", action="store_true",
                        help="render per-batch lines once per epoch from the device ring buffer (no per-step host sync)")
    parser.add_argument("--save-every", default=0, type=int, help="also write checkpoint-epoch-N.pt every N epochs")
    parser.add_argument("--resume", default=None, type=Path, help="checkpoint to resume from")
    parser.add_argument("--history-path", default=None, type=Path)
    parser.add_argument("--synthetic", action="store_true",
                        help="write a synthetic dataset of the reference's shape to --dataset-path if none is there")

    sub_parser = parser.add_subparsers(title="Available commands", metavar="command [options ...]")
    sub_parser.required = True
    param_server.add_sub_command(sub_parser)
    trainer.add_sub_commands(sub_parser)
    return parser


def main(argv=None):
    parser = build_parser()
    args = parser.parse_args(argv)
    if args.seed is not None:
        torch.manual_seed(args.seed)
    if args.synthetic:
        from .data import synthetic
        from .data.dataset import MotionDataset
        fp, _ = MotionDataset.get_data_path(args.dataset_path, "train")
        if not fp.exists():
            from .parallel.process_group import env_rank_info
            rank, world, _ = env_rank_info()
            if rank == 0:
                synthetic.write_dataset(args.dataset_path, seed=args.seed or 0)
            elif world > 1:
                import time
                tp, _ = MotionDataset.get_data_path(args.dataset_path, "test")
                for _ in range(600):
                    if tp.exists():
                        break
                    time.sleep(0.1)
    return args.func(args)


if __name__ == "__main__":
    sys.dont_write_bytecode = True
    main()
