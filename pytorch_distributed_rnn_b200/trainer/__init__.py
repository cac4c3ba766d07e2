"""Trainer family, the sub-command registry and the run driver behind the CLI.

Capability parity with src/motion/trainer/__init__.py (sub-commands ``local`` / ``distributed`` /
``horovod``, :10-18; a run = set the log level, load the three datasets, optionally drop validation and
test, build the model and the trainer, train, write ``history.json``, :21-60).  Deviations:
``--validation-fraction`` is honoured (Q2), only rank 0 writes ``history.json`` (Q8), the numpy split RNG
follows ``--seed`` (Q14).  Extensions reach the trainer through optional CLI attributes (``--cell``,
``--device``, ``--deferred-logging``, ``--save-every``, ``--resume``, ``--history-path``).
"""
from __future__ import annotations

import json
import logging
from functools import partial
from typing import Dict, Type

from ..data.dataset import MotionDataset
from ..models.motion import MotionModel
from .base import Trainer
from .ddp import DDPTrainer
from .distributed import DistributedTrainer
from .formatter import TrainingMessageFormatter
from .horovod import HorovodTrainer

__all__ = ["Trainer", "DistributedTrainer", "DDPTrainer", "HorovodTrainer", "TrainingMessageFormatter",
           "TRAINERS", "add_sub_commands", "train"]

# sub-command -> trainer class (the parameter-server sub-command registers itself in parallel/param_server.py)
TRAINERS: Dict[str, Type[Trainer]] = {"local": Trainer, "distributed": DDPTrainer, "horovod": HorovodTrainer}

# optional CLI attributes forwarded to the trainer constructor: attribute name -> (keyword, default)
_OPTIONAL = {"device": ("device", None), "deferred_logging": ("deferred_logging", False),
             "save_every": ("save_every", 0), "resume": ("resume", None)}


def add_sub_commands(sub_parser) -> None:
    for name, cls in TRAINERS.items():
        sub_parser.add_parser(name).set_defaults(func=partial(_run, trainer_cls=cls))


def _run(args, trainer_cls):
    return train(args, trainer_cls)


def _datasets(args):
    splits = MotionDataset.load(args.dataset_path, output_path=args.output_path,
                                validation_fraction=args.validation_fraction, seed=args.seed)
    names = ("Training", "Validation", "Test")
    keep = 1 if args.no_validation else 3
    for name, ds in list(zip(names, splits))[:keep]:
        logging.info(f"{name} set of size {len(ds)}")
    return tuple(ds if i < keep else None for i, ds in enumerate(splits))


def _write_history(args, train_history, validation_history) -> None:
    path = getattr(args, "history_path", None) or "history.json"
    with open(path, "w") as fh:
        json.dump({"train_history": train_history, "validation_history": validation_history}, fh)


DELAY_MARK = "Injected delay: {} us"      # bench/sweep.py splits a sweep's log at these lines
LOSS_MARK = "Injected loss: {} %"
WARMUP_MARK = "Injected warmup: 0 us"     # the untimed first run of a sweep


def _network_sweep(args, trainer, delays_us=(), losses_pct=()):
    """The reference's network experiment (tc netem delay / loss on every host, one job per rule:
    fabfile.py:130-191) as ONE job: after an untimed warm-up run (first-use costs: module load, buffer
    allocation, behind an ``Injected warmup`` marker), the same ``--epochs`` training is repeated once per injected rule; every repetition logs the
    usual performance line behind a ``Injected delay: N us`` / ``Injected loss: P %`` marker.  A delay is spun
    on the device in front of every flag publication of the in-kernel collectives (`SymmComm.set_delay_ns`),
    i.e. it lengthens every message's latency; a loss makes that publication pay a retransmission timeout with
    the given probability (`SymmComm.set_loss`)."""
    comm = trainer._get_comm()
    if comm is None or not hasattr(comm, "set_delay_ns"):
        raise SystemExit("--delay-sweep-us / --loss-sweep-pct need the in-kernel NVLink communicator "
                         "(a multi-GPU distributed / horovod run)")
    # warm-up at the normal log level (the first INFO-level epoch pays one-off costs of its own: ~3 ms on B200, as much
    # as a whole epoch); bench/sweep.py drops the segment behind this marker
    logging.info(WARMUP_MARK)
    trainer.train(epochs=args.epochs)
    rules = [(DELAY_MARK, d, lambda v: comm.set_delay_ns(v * 1000)) for d in delays_us]
    rules += [(LOSS_MARK, p, lambda v: comm.set_loss(v)) for p in losses_pct]
    for mark, value, apply in rules:
        comm.set_delay_ns(0)
        comm.set_loss(0.0)
        apply(value)
        logging.info(mark.format(value))        # on every rank: each rank's log is split at its own markers
        trainer.start_epoch = 0
        trainer.train(epochs=args.epochs)
    comm.set_delay_ns(0)
    comm.set_loss(0.0)
    return trainer


def _delay_sweep(args, trainer, delays_us):
    return _network_sweep(args, trainer, delays_us=delays_us)


def train(args, trainer_cls):
    """One CLI run: returns the trainer (tests and notebooks inspect it)."""
    logging.getLogger().setLevel(args.log)
    train_ds, val_ds, test_ds = _datasets(args)
    net = MotionModel(train_ds.num_features, args.hidden_units, args.stacked_layer, len(MotionDataset.LABELS),
                      cell=getattr(args, "cell", "lstm"))
    extras = {kw: getattr(args, attr, default) for attr, (kw, default) in _OPTIONAL.items()}
    trainer = trainer_cls(model=net, training_set=train_ds, validation_set=val_ds, test_set=test_ds,
                          batch_size=args.batch_size, learning_rate=args.learning_rate,
                          checkpoint_dir=args.checkpoint_directory, **extras)
    delays = [int(v) for v in str(getattr(args, "delay_sweep_us", None) or "").split(",") if v != ""]
    losses = [float(v) for v in str(getattr(args, "loss_sweep_pct", None) or "").split(",") if v != ""]
    if delays or losses:
        return _network_sweep(args, trainer, delays_us=delays, losses_pct=losses)
    logging.info(f"Training model for {args.epochs} epochs...")
    _, train_history, validation_history = trainer.train(epochs=args.epochs)
    if getattr(trainer, "rank", 0) == 0:
        _write_history(args, train_history, validation_history)
    return trainer
