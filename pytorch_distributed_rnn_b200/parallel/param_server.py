"""``parameter-server`` sub-command: asynchronous data parallelism around one shared model.

Parity target: src/motion/param_server/{__init__,master,worker,util}.py.  There, rank 0 hosts
the only ``MotionModel`` behind torch RPC; each worker ships its whole batch to rank 0, the
forward AND backward run on rank 0 (dist-autograd), and a per-worker remote Adam updates the
shared parameters — no barrier between workers (Hogwild-style), shard 0 of the sampler is
never trained, the batch is not divided by the world size, workers never evaluate or
checkpoint.  All of that is kept, except where the model runs:

* rank 0 (the server) owns the shared parameter vector and one Adam state per worker;
* workers keep the compute: pull parameters -> local forward/backward on their own device
  -> push the gradient -> the server applies that worker's Adam step to the shared vector and
  returns the fresh parameters.  One request/response per step instead of the reference's
  four RPC round trips (forward, backward, gradient check, optimizer step; S13-S16).

Transport, two forms:

* GPUs (K10, the product path): ONE-SIDED over NVLink.  The server process only allocates the shared
  parameter vector inside its symmetric heap and then idles; a worker's step is fused forward ->
  fused backward -> ``b200rnn_comm_ps_push`` (reduce its gradient rows, run ITS Adam state, add the
  resulting delta to the server's vector with ``red.global.add`` over NVLink) ->
  ``b200rnn_comm_ps_pull`` (peer loads of the fresh parameters).  No server CPU or kernel is involved
  in a step; the end of training is signalled with the K9 put primitive.
* CPU (plumbing): ``torch.distributed`` point-to-point over gloo (any-source receive on the server)
  carrying one flat fp32 message per direction.  CLI: ``parameter-server --world-size W --rank R
[--master-address localhost] [--master-port 29500]`` — one invocation per node, like the
reference (param_server/__init__.py:11-37).
"""
from __future__ import annotations

import logging
import math
import os
from typing import Dict, Optional

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ..data.dataset import MotionDataset
from ..data.sampler import ShardedEpochSampler
from ..models.motion import MotionModel
from ..trainer.base import Trainer
from ..trainer.formatter import TrainingMessageFormatter
from ..utils.arena import ParamArena

OP_PULL, OP_PUSH, OP_DONE = 0.0, 1.0, 2.0


def add_sub_command(parent_parser):
    parser = parent_parser.add_parser("parameter-server")
    parser.add_argument("--world-size", type=int, required=True,
                        help="Total number of participating processes: the server plus all workers.")
    parser.add_argument("--rank", type=int, required=True, help="Global rank of this process; 0 is the server.")
    parser.add_argument("--master-address", type=str, default="localhost")
    parser.add_argument("--master-port", type=str, default="29500")
    parser.set_defaults(func=execute)


def _init_pg(rank: int, world_size: int, device: torch.device) -> None:
    # request/response messages are staged through host memory: gloo is the only backend with
    # any-source receive, which is what makes the server asynchronous
    if not dist.is_initialized():
        dist.init_process_group("gloo", rank=rank, world_size=world_size)


def _pick_device(rank: int, device: Optional[str]) -> torch.device:
    if device is not None:
        return torch.device(device)
    if torch.cuda.is_available():
        return torch.device("cuda", rank % torch.cuda.device_count())
    return torch.device("cpu")


class ServerState:
    """Shared parameters + one Adam state per worker (the reference gives every worker its own
    ``DistributedOptimizer`` -> its own ``_LocalOptimizer`` on the server; Q11)."""

    def __init__(self, model: MotionModel, lr: float, betas=(0.9, 0.999), eps: float = 1e-8):
        self.arena = ParamArena(model)
        self.lr, self.betas, self.eps = lr, betas, eps
        self.state: Dict[int, dict] = {}

    def apply(self, worker: int, grad: torch.Tensor) -> None:
        p = self.arena.flat[:self.arena.numel]
        st = self.state.setdefault(worker, dict(step=0, m=torch.zeros_like(p), v=torch.zeros_like(p)))
        st["step"] += 1
        b1, b2 = self.betas
        st["m"].mul_(b1).add_(grad, alpha=1 - b1)
        st["v"].mul_(b2).addcmul_(grad, grad, value=1 - b2)
        bc1, bc2 = 1 - b1 ** st["step"], 1 - b2 ** st["step"]
        denom = (st["v"].sqrt() / math.sqrt(bc2)).add_(self.eps)
        p.addcdiv_(st["m"], denom, value=-self.lr / bc1)


def _symm_comm(rank: int, world_size: int, dev: torch.device):
    from .process_group import RankInfo
    from .symm import SymmComm
    return SymmComm(RankInfo(rank, world_size, dev.index or 0, dev, "gloo"), inbox_floats=1 << 16, heap_bytes=64 << 20)


class PSWorkerComm:
    """Engine hook of a GPU worker: push (reduce + own Adam + remote add) then pull."""
    kind = "ps"
    grad_world = 1      # no averaging: every worker applies its own full step (asynchronous SGD)
    always = True

    def __init__(self, comm, heap_off: int, server: int = 0):
        self.comm, self.heap_off, self.server = comm, heap_off, server
        self.world, self.rank, self.device = comm.world, comm.rank, comm.device

    def pull(self, arena) -> None:
        self.comm.ps_pull(arena.flat[:arena.numel], self.server, self.heap_off)

    def allreduce_update(self, rows, n_rows, arena, opt, optdesc, metrics_slot) -> None:
        self.comm.ps_push(rows, n_rows, arena.numel, opt, optdesc, metrics_slot, self.server, self.heap_off)
        self.pull(arena)


def _run_parameter_server_gpu(rank, world_size, dev, input_dim, hidden_dim, layer_dim, output_dim, seed):
    import time
    comm = _symm_comm(rank, world_size, dev)
    if seed is not None:
        torch.manual_seed(seed)
    model = MotionModel(input_dim, hidden_dim, layer_dim, output_dim).to(dev)
    arena = comm.make_arena(model)           # the shared vector lives in the server's symmetric heap
    torch.cuda.synchronize(dev)
    dist.barrier()                           # workers may pull from now on
    print("RPC initialized! Running parameter server... (one-sided NVLink mode: the server only hosts memory)")
    while any(comm.signal_value(w) < 1 for w in range(1, world_size)):
        time.sleep(0.005)
    print("RPC shutdown on parameter server. (workers finished; parameter checksum %.6f)" % float(arena.flat[:arena.numel].double().sum()))
    dist.barrier()
    comm.close()
    dist.destroy_process_group()
    return arena


def run_parameter_server(rank, world_size, input_dim, hidden_dim, layer_dim, output_dim, lr,
                         device: Optional[str] = None, seed: Optional[int] = None):
    dev = _pick_device(rank, device)
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    print("PS master initializing RPC")
    _init_pg(rank, world_size, dev)
    if dev.type == "cuda":
        return _run_parameter_server_gpu(rank, world_size, dev, input_dim, hidden_dim, layer_dim, output_dim, seed)
    if seed is not None:
        torch.manual_seed(seed)
    model = MotionModel(input_dim, hidden_dim, layer_dim, output_dim).to(dev)
    server = ServerState(model, lr)
    P = server.arena.numel
    print("RPC initialized! Running parameter server...")
    # any-source receive: requests are served in arrival order, no barrier between workers
    buf = torch.zeros(P + 1, dtype=torch.float32)
    reply = torch.zeros(P, dtype=torch.float32)
    active = world_size - 1
    updates = 0
    while active > 0:
        w = dist.recv(buf)
        op = float(buf[0])
        if op == OP_DONE:
            active -= 1
            continue
        if op == OP_PUSH:
            server.apply(w, buf[1:].to(dev))
            updates += 1
        reply.copy_(server.arena.flat[:P])
        dist.send(reply, dst=w)
    print(f"RPC shutdown on parameter server. ({updates} asynchronous updates applied)")
    dist.destroy_process_group()
    return server


class ParameterWorkerTrainer(Trainer):
    """Worker: local compute, remote parameters.  Never evaluates (rank != 0) or checkpoints."""

    def __init__(self, rank, world_size, model, training_set, batch_size, learning_rate,
                 validation_set=None, test_set=None, checkpoint_dir=None, device=None):
        self.rank = rank
        self.world_size = world_size
        super().__init__(model=model, training_set=training_set,
                         validation_set=validation_set if rank == 0 else None,
                         test_set=test_set if rank == 0 else None,
                         batch_size=batch_size, learning_rate=learning_rate, checkpoint_dir=checkpoint_dir,
                         sampler=ShardedEpochSampler(len(training_set), num_replicas=world_size, rank=rank),
                         device=device)
        P = self.arena.numel
        if self._ps is not None:
            dist.barrier()                    # the server's arena exists and is initialised
            self._ps.pull(self.arena)
            return
        self._msg = torch.zeros(P + 1, dtype=torch.float32)
        self._reply = torch.zeros(P, dtype=torch.float32)
        self._pull()

    def _get_arena(self):
        if self.device.type == "cuda":
            self._symm = _symm_comm(self.rank, self.world_size, self.device)
            arena = self._symm.make_arena(self._core_model())   # same heap offset as on the server
            self._ps = PSWorkerComm(self._symm, arena.heap_off)
            return arena
        self._ps = None
        return super()._get_arena()

    def _get_comm(self):
        return self._ps

    def _get_engine(self):
        if self._ps is not None:
            return super()._get_engine()   # fused kernels; the optimizer step is the one-sided push
        return None  # CPU: gradients must surface between backward and the (remote) optimizer step

    def _get_formatter(self, epochs):
        return TrainingMessageFormatter(epochs, self.rank)

    def _pull(self):
        self._msg[0] = OP_PULL
        dist.send(self._msg, dst=0)
        dist.recv(self._reply, src=0)
        self.arena.flat[:self.arena.numel].copy_(self._reply)

    def _torch_train_step(self, ds, idx):
        idx = idx.to(ds.features.device)
        data = ds.features[idx].to(self.device)
        labels = ds.labels[idx].to(self.device).long().reshape(-1)
        self.arena.zero_grad()
        output = self.model(data)
        loss = self.loss_fn(output, labels)
        correct = (torch.argmax(output, dim=1) == labels).sum()
        loss.backward()
        P = self.arena.numel
        assert bool(torch.any(self.arena.grad[:P] != 0)), "no gradient reached the parameters"  # worker.py:55-58
        self._msg[0] = OP_PUSH
        self._msg[1:].copy_(self.arena.grad[:P])
        dist.send(self._msg, dst=0)                # push gradient
        dist.recv(self._reply, src=0)              # pull fresh parameters (remote Adam already applied)
        self.arena.flat[:P].copy_(self._reply)
        return torch.stack([loss.detach() * labels.numel(), correct.to(loss.dtype)])

    def finish(self):
        if self._ps is not None:
            torch.cuda.synchronize(self.device)
            self._symm.put(torch.ones(1, device=self.device), dst=0)   # K9: "done" signal to the server
            torch.cuda.synchronize(self.device)
            dist.barrier()
            self._symm.close()
            return
        self._msg[0] = OP_DONE
        dist.send(self._msg, dst=0)

    def _save_checkpoint(self, epoch, loss, best=False):
        return


def run_worker(rank, world_size, epochs, batch_size, learning_rate, input_dim, hidden_dim, layer_dim, output_dim,
               train_set, validation_set, test_set, device: Optional[str] = None, log_level="INFO"):
    logging.getLogger().setLevel(log_level)
    dev = _pick_device(rank, device)
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    print(f"Worker rank {rank} initializing RPC")
    _init_pg(rank, world_size, dev)
    print(f"Worker {rank} done initializing RPC")
    model = MotionModel(input_dim, hidden_dim, layer_dim, output_dim)
    worker = ParameterWorkerTrainer(rank, world_size, model, train_set, batch_size, learning_rate,
                                    validation_set, test_set, device=dev)
    result = worker.train(epochs)
    worker.finish()
    dist.destroy_process_group()
    return result


def execute(args):
    os.environ["MASTER_ADDR"] = args.master_address
    os.environ["MASTER_PORT"] = args.master_port
    world_size = args.world_size
    training_set, validation_set, test_set = MotionDataset.load(
        args.dataset_path, output_path=args.output_path, validation_fraction=args.validation_fraction, seed=args.seed)
    dims = (training_set.num_features, args.hidden_units, args.stacked_layer, len(MotionDataset.LABELS))
    ctx = mp.get_context("spawn")
    if args.rank == 0:
        p = ctx.Process(target=run_parameter_server,
                        args=(0, world_size, *dims, args.learning_rate, args.device, args.seed))
    else:
        torch.set_num_threads(args.num_threads)  # Q3: only PS workers honour --num-threads
        p = ctx.Process(target=run_worker,
                        args=(args.rank, world_size, args.epochs, args.batch_size, args.learning_rate, *dims,
                              training_set, validation_set, test_set, args.device, args.log))
    p.start()
    p.join()
    if p.exitcode != 0:
        raise SystemExit(p.exitcode)
