"""Trainer family + sub-command registry + the ``train(args, trainer_cls)`` driver.

Parity: src/motion/trainer/__init__.py — sub-commands ``local`` / ``distributed`` /
``horovod`` (:10-18); ``train`` sets the log level, loads the datasets, drops validation/test
under ``--no-validation``, builds ``MotionModel`` and the trainer, trains, and dumps
``history.json`` (:21-60).  Deviations: ``--validation-fraction`` is honoured
(Q2), ``history.json`` is written by rank 0 only (Q8), the numpy split RNG follows ``--seed``
(Q14).
"""
from __future__ import annotations

import json
import logging

from ..data.dataset import MotionDataset
from ..models.motion import MotionModel
from .base import Trainer
from .ddp import DDPTrainer
from .distributed import DistributedTrainer
from .formatter import TrainingMessageFormatter
from .horovod import HorovodTrainer

__all__ = ["Trainer", "DistributedTrainer", "DDPTrainer", "HorovodTrainer", "TrainingMessageFormatter",
           "add_sub_commands", "train"]


def add_sub_commands(sub_parser):
    create_parser(sub_parser, "local", Trainer)
    create_parser(sub_parser, "distributed", DDPTrainer)
    create_parser(sub_parser, "horovod", HorovodTrainer)


def create_parser(sub_parser, name, trainer_class):
    parser = sub_parser.add_parser(name)
    parser.set_defaults(func=lambda args: train(args, trainer_class))


def train(args, trainer_cls):
    logging.getLogger().setLevel(args.log)
    training_set, validation_set, test_set = MotionDataset.load(
        args.dataset_path, output_path=args.output_path,
        validation_fraction=args.validation_fraction, seed=args.seed)
    logging.info(f"Training set of size {len(training_set)}")
    if args.no_validation:
        validation_set = None
        test_set = None
    else:
        logging.info(f"Validation set of size {len(validation_set)}")
        logging.info(f"Test set of size {len(test_set)}")
    model = MotionModel(input_dim=training_set.num_features, hidden_dim=args.hidden_units,
                        layer_dim=args.stacked_layer, output_dim=len(MotionDataset.LABELS),
                        cell=getattr(args, "cell", "lstm"))
    trainer = trainer_cls(
        model=model, training_set=training_set, validation_set=validation_set, test_set=test_set,
        batch_size=args.batch_size, learning_rate=args.learning_rate, checkpoint_dir=args.checkpoint_directory,
        device=getattr(args, "device", None), deferred_logging=getattr(args, "deferred_logging", False),
        save_every=getattr(args, "save_every", 0), resume=getattr(args, "resume", None))
    logging.info(f"Training model for {args.epochs} epochs...")
    _, train_history, validation_history = trainer.train(epochs=args.epochs)
    if getattr(trainer, "rank", 0) == 0:
        with open(getattr(args, "history_path", None) or "history.json", "w") as file:
            json.dump({"train_history": train_history, "validation_history": validation_history}, file)
    return trainer
