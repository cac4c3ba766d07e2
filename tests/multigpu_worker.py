"""Worker run under the launcher by tests/test_comm_gpu.py (one process per GPU)."""
import copy
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import FusedAdam, MotionModel, ParamArena  # noqa: E402
from pytorch_distributed_rnn_b200.data.sampler import epoch_indices  # noqa: E402
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine  # noqa: E402
from pytorch_distributed_rnn_b200.parallel import process_group as pg  # noqa: E402
from pytorch_distributed_rnn_b200.parallel.symm import SymmComm  # noqa: E402
import torch.distributed as dist  # noqa: E402


def gather_all(t):
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return out


def main():
    info = pg.init_distributed()
    rank, world, dev = info.rank, info.world_size, info.device
    comm = SymmComm(info, inbox_floats=2 << 20, heap_bytes=128 << 20)
    res = {"rank": rank, "world": world, "topology_all_to_all": comm.topology["all_to_all"],
           "multicast": comm.topology["multicast_supported"]}

    # 1. plain all-reduce vs the analytic sum; bitwise agreement across ranks
    ok = True
    for n in (1, 255, 14150, 1 << 20, (2 << 20) + 77):
        g = torch.Generator(device="cpu").manual_seed(n)
        base = torch.randn(n, generator=g)
        t = (base * (rank + 1)).to(dev)
        comm.allreduce_flat(t, average=False)
        want = base.double() * sum(r + 1 for r in range(world))
        ok &= bool(torch.allclose(t.cpu().double(), want, rtol=1e-5, atol=1e-5))
        ok &= all(torch.equal(x, t) for x in gather_all(t))
    for algo in (4, 1):     # both one-shot forms explicitly on the reference's gradient size
        base = torch.randn(14150, generator=torch.Generator().manual_seed(algo))
        t = (base * (rank + 1)).to(dev)
        comm.allreduce_flat(t, average=False, algo=algo)
        ok &= bool(torch.allclose(t.cpu().double(), base.double() * sum(r + 1 for r in range(world)), rtol=1e-5, atol=1e-5))
        ok &= all(torch.equal(x, t) for x in gather_all(t))
    res["allreduce_ok"] = ok

    # 2. fused all-reduce + Adam (one-shot and two-shot) vs torch Adam on the averaged gradient
    res["nvls_available"] = comm.mc_base is not None
    algos = [(1, "oneshot"), (2, "twoshot"), (4, "ll")] + ([(3, "nvls")] if comm.mc_base is not None else [])
    for algo, name in algos:
        torch.manual_seed(5)
        lin = torch.nn.Linear(301, 97).to(dev)      # 29 294 parameters: several blocks, ragged tail
        ref = copy.deepcopy(lin)
        arena = comm.make_arena(lin)
        opt = FusedAdam(arena, lr=0.01)
        ropt = torch.optim.Adam(ref.parameters(), lr=0.01)
        P = arena.numel
        comm.algo = algo
        good = True
        for step in range(3):
            g = torch.Generator().manual_seed(100 + step)
            base = torch.randn(3, P + 2, generator=g)
            rows = (base * (rank + 1)).to(dev).contiguous()      # 3 local "CTA rows" per rank
            opt.step_count += 1
            slot = torch.zeros(2, device=dev)
            comm.allreduce_update(rows, 3, arena, opt, opt.desc(grad_scale=1.0 / world), slot)
            avg = base[:, :P].sum(0) * sum(r + 1 for r in range(world)) / world
            off = 0
            for p in ref.parameters():
                p.grad = avg[off:off + p.numel()].view_as(p).to(dev)
                off += p.numel()
            ropt.step()
            good &= bool(torch.allclose(slot.cpu(), base[:, P:].sum(0) * (rank + 1), rtol=1e-5))
        flat_ref = torch.cat([p.detach().view(-1) for p in ref.parameters()])
        good &= bool(torch.allclose(arena.flat[:P], flat_ref, atol=2e-6))
        good &= all(torch.equal(x, arena.flat[:P].clone()) for x in gather_all(arena.flat[:P].clone()))
        res[f"update_{name}_ok"] = good
    comm.algo = 0

    # 3. broadcast (heap-resident and staged), put/get, barrier
    t, off = comm.heap_tensor(1000)
    t.fill_(float(rank + 1))
    comm.broadcast_flat(t, root=world - 1)
    u = torch.full((333,), float(rank + 10), device=dev)
    comm.broadcast_flat(u, root=0)
    res["broadcast_ok"] = bool((t == world).all()) and bool((u == 10).all())
    msg = torch.zeros(4, device=dev)
    if rank == 0:
        msg += torch.arange(4, device=dev) + 1
        for d in range(1, world):
            comm.put(msg, d)
    else:
        comm.get(msg, 0)
    comm.barrier()
    res["p2p_ok"] = msg.cpu().tolist() == [1.0, 2.0, 3.0, 4.0]

    # 4. data-parallel engine: per-step loss invariance vs a single-rank run of the same global batch
    torch.manual_seed(123456789)
    model = MotionModel(9, 32, 2, 6).to(dev)
    arena = comm.make_arena(model)
    comm.broadcast_flat(arena.flat[:arena.numel], root=0)
    eng = FusedTrainEngine(model, lr=0.0025, arena=arena, comm=comm)
    torch.manual_seed(123456789)
    solo_model = MotionModel(9, 32, 2, 6).to(dev)
    solo = FusedTrainEngine(solo_model, lr=0.0025)
    g = torch.Generator().manual_seed(1)
    N, GB = 960, 480
    X, Y = torch.randn(N, 128, 9, generator=g).to(dev), torch.randint(0, 6, (N,), generator=g).to(dev)
    inv = True
    for step in range(2):
        idx_l = epoch_indices(N, 1, 0, 0)[step * GB:(step + 1) * GB].to(dev)
        idx_r = epoch_indices(N, world, rank, 0)[step * (GB // world):(step + 1) * (GB // world)].to(dev)
        ml = solo.train_step(X[idx_l], Y[idx_l]).clone()
        mr = eng.train_step(X[idx_r], Y[idx_r]).clone()
        tot = mr.clone()
        dist.all_reduce(tot)
        inv &= abs(float(tot[0]) / GB - float(ml[0]) / GB) < 5e-6 and int(round(float(tot[1]))) == int(round(float(ml[1])))
    # parameters: Adam's first steps are sign-like (m/sqrt(v) = +-1), so an element whose gradient is at
    # rounding-noise level may legitimately differ by ~lr between summation orders; all others agree
    diff = (arena.flat[:arena.numel] - solo.arena.flat[:solo.arena.numel]).abs()
    res["dp_param_mean_abs_diff"] = float(diff.mean())
    res["dp_param_frac_gt_1e-4"] = float((diff > 1e-4).float().mean())
    inv &= float(diff.mean()) < 1e-3
    inv &= all(torch.equal(x, arena.flat[:arena.numel].clone()) for x in gather_all(arena.flat[:arena.numel].clone()))
    res["dp_invariance_ok"] = inv

    # 4b. DDP wrapper on a multi-bucket autograd model: bucket all-reduces run on a side stream while
    #     backward continues; gradients must equal the NCCL-averaged local gradients, identically on all ranks
    from pytorch_distributed_rnn_b200.parallel.ddp import DistributedDataParallel as DDP
    torch.manual_seed(11)
    mlp = torch.nn.Sequential(torch.nn.Linear(512, 512), torch.nn.ReLU(), torch.nn.Linear(512, 512), torch.nn.ReLU(),
                              torch.nn.Linear(512, 10)).to(dev)
    twin = copy.deepcopy(mlp)
    ddp = DDP(mlp, comm=comm, arena=comm.make_arena(mlp))
    ov = len(ddp.buckets) >= 2
    for it in range(3):
        g = torch.Generator().manual_seed(1000 * it + rank)
        xb = torch.randn(64, 512, generator=g).to(dev)
        ddp.arena.zero_grad()
        ddp(xb).square().mean().backward()
        twin.zero_grad()
        twin(xb).square().mean().backward()
        torch.cuda.synchronize()
        for p, q in zip(mlp.parameters(), twin.parameters()):
            want = q.grad.clone()
            dist.all_reduce(want)
            want /= world
            ov &= bool(torch.allclose(p.grad, want, rtol=1e-4, atol=1e-6))
    ov &= all(torch.equal(x, ddp.arena.grad[:ddp.arena.numel].clone()) for x in gather_all(ddp.arena.grad[:ddp.arena.numel].clone()))
    res["ddp_overlap_ok"] = ov
    res["ddp_buckets"] = len(ddp.buckets)

    # 4c. large-hidden engine: per-layer gradient buckets, device-side triggers, slice all-reduce + Adam on the
    #     communication stream — against torch DistributedDataParallel (NCCL) on an fp32 twin of the model
    from pytorch_distributed_rnn_b200.ops.layered_engine import LayeredTrainEngine
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    torch.manual_seed(77)
    big = MotionModel(9, 128, 2, 6).to(dev)
    twin = TorchDDP(copy.deepcopy(big), device_ids=[dev.index])
    b_arena = comm.make_arena(big)
    comm.broadcast_flat(b_arena.flat[:b_arena.numel], root=0)
    leng = LayeredTrainEngine(big, lr=0.0025, arena=b_arena, comm=comm, timeline=True)
    topt = torch.optim.Adam(twin.parameters(), lr=0.0025)
    lay = True
    for step in range(3):
        g = torch.Generator().manual_seed(500 + 10 * step + rank)
        xb, yb = torch.randn(16, 24, 9, generator=g).to(dev), torch.randint(0, 6, (16,), generator=g).to(dev)
        mine = leng.train_step(xb, yb).clone()
        topt.zero_grad()
        tl = torch.nn.functional.cross_entropy(twin(xb), yb)
        tl.backward()
        topt.step()
        lay &= abs(float(mine[0]) / 16 - float(tl.detach())) < 2e-2 * max(1.0, abs(float(tl.detach())))
    torch.cuda.synchronize()
    pd = torch.cat([(p.detach() - q.detach()).abs().view(-1) for p, q in zip(big.parameters(), twin.module.parameters())])
    res["layered_param_mean_abs_diff"] = float(pd.mean())
    lay &= float(pd.mean()) < 0.25 * 0.0025 * 3
    lay &= all(torch.equal(x, b_arena.flat[:b_arena.numel].clone()) for x in gather_all(b_arena.flat[:b_arena.numel].clone()))
    lay &= leng.reducer_status() == 0
    res["layered_timeline"] = leng.last_timeline()
    res["layered_ddp_ok"] = lay

    # 5. buffer-reuse soak with injected delay on one rank (straggler)
    comm.set_delay_ns(20000 if rank == world - 1 else 0)
    soak = True
    for i in range(200):
        v = torch.full((5000,), float(i + rank), device=dev)
        comm.allreduce_flat(v, average=False)
        soak &= float(v[0]) == float(sum(i + r for r in range(world)))
    comm.set_delay_ns(0)
    comm.check()
    res["soak_ok"] = soak
    torch.cuda.synchronize()
    bad = [k for k, v in res.items() if k.endswith("_ok") and v is not True]
    if bad:
        print(f"FAILED_CHECKS rank {rank}: {bad}", file=sys.stderr, flush=True)
    print("RESULT " + json.dumps(res), flush=True)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
