"""In-kernel collectives: world-of-one checks on a single GPU, real multi-rank checks on >= 2 GPUs."""
import copy
import json
import sys
from pathlib import Path

import pytest
import torch

from pytorch_distributed_rnn_b200 import FusedAdam

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def comm1():
    from pytorch_distributed_rnn_b200.parallel import process_group as pg
    from pytorch_distributed_rnn_b200.parallel.symm import SymmComm
    info = pg.init_distributed()
    if info.world_size != 1:
        pytest.skip("single-process fixture")
    c = SymmComm(info, inbox_floats=1 << 20, heap_bytes=64 << 20)
    yield c
    c.close()
    pg.shutdown()


@pytest.mark.gpu
def test_topology_probe():
    from pytorch_distributed_rnn_b200.parallel.bootstrap import probe_topology
    t = probe_topology()
    assert t["n_devices"] == torch.cuda.device_count() and "B200" in t["names"][0] or t["n_devices"] >= 1
    assert t["sm_count"][0] > 0 and t["can_access"][0][0] == 1


@pytest.mark.gpu
@pytest.mark.parametrize("algo", [1, 2, 4])
def test_world1_allreduce_update_equals_fused_update(comm1, algo):
    torch.manual_seed(0)
    lin = torch.nn.Linear(129, 65).cuda()
    ref = copy.deepcopy(lin)
    arena = comm1.make_arena(lin)
    opt = FusedAdam(arena, lr=0.01)
    ropt = torch.optim.Adam(ref.parameters(), lr=0.01)
    P = arena.numel
    comm1.algo = algo
    for step in range(3):
        rows = torch.randn(5, P + 2, device="cuda")
        opt.step_count += 1
        slot = torch.zeros(2, device="cuda")
        comm1.allreduce_update(rows, 5, arena, opt, opt.desc(grad_scale=1.0), slot)
        g = rows[:, :P].sum(0)
        off = 0
        for p in ref.parameters():
            p.grad = g[off:off + p.numel()].view_as(p).clone()
            off += p.numel()
        ropt.step()
        assert torch.allclose(slot, rows[:, P:].sum(0), rtol=1e-5)
    comm1.algo = 0
    comm1.check()
    assert torch.allclose(arena.flat[:P], torch.cat([p.detach().view(-1) for p in ref.parameters()]), atol=2e-6)


@pytest.mark.gpu
def test_world1_allreduce_broadcast_barrier(comm1):
    v = torch.randn(300000, device="cuda")
    w = v.clone()
    comm1.allreduce_flat(v, average=True)
    comm1.barrier()
    comm1.broadcast_flat(v, root=0)
    comm1.check()
    assert torch.equal(v, w)
    for n in (1, 3, 130, 14150, 65536):   # the LL form (auto for small vectors) and the flag form, ragged tails
        for algo in (4, 1):
            a = torch.randn(n, device="cuda")
            b = a.clone()
            comm1.allreduce_flat(a, average=False, algo=algo)
            assert torch.equal(a, b), (n, algo)
    comm1.check()


@pytest.mark.gpu
@pytest.mark.parametrize("wd", [0.0, 0.05])
def test_world1_ps_push_matches_torch_adam_with_weight_decay(comm1, wd):
    """K10 on one GPU (server == worker): three pushes equal three torch Adam steps, weight decay included (the delta
    then depends on the server's current value, which the kernel reads before it adds)."""
    torch.manual_seed(1)
    lin = torch.nn.Linear(37, 11).cuda()
    ref = copy.deepcopy(lin)
    arena = comm1.make_arena(lin)
    opt = FusedAdam(arena, lr=0.01, weight_decay=wd)
    ropt = torch.optim.Adam(ref.parameters(), lr=0.01, weight_decay=wd)
    P = arena.numel
    for step in range(3):
        rows = torch.randn(4, P + 2, device="cuda")
        opt.step_count += 1
        slot = torch.zeros(2, device="cuda")
        comm1.ps_push(rows, 4, P, opt, opt.desc(grad_scale=1.0), slot, 0, arena.heap_off)
        g = rows[:, :P].sum(0)
        off = 0
        for p in ref.parameters():
            p.grad = g[off:off + p.numel()].view_as(p).clone()
            off += p.numel()
        ropt.step()
        assert torch.allclose(slot, rows[:, P:].sum(0), rtol=1e-5)
    comm1.check()
    assert torch.allclose(arena.flat[:P], torch.cat([p.detach().view(-1) for p in ref.parameters()]), atol=2e-6)


@pytest.mark.gpu
def test_injected_loss_costs_retransmission_timeouts_and_keeps_results(comm1):
    """`tc netem loss` analogue: at 100 % every publication pays the full run of 8 retransmission timeouts, at 0 %
    none; the values that arrive are the same either way."""
    a = torch.randn(14150, device="cuda")
    want = a.clone()

    def timed(algo):
        v = a.clone()
        comm1.allreduce_flat(v, average=False, algo=algo)       # warm
        v.copy_(a)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        comm1.allreduce_flat(v, average=False, algo=algo)
        e1.record()
        torch.cuda.synchronize()
        assert torch.equal(v, want)
        return e0.elapsed_time(e1) * 1e3                        # us

    for algo in (4, 1):
        base = timed(algo)
        comm1.set_loss(100.0, rto_ns=50_000)
        lossy = timed(algo)
        comm1.set_loss(0.0)
        assert lossy > base + 0.5 * 8 * 50 and lossy < base + 4 * 8 * 50 + 200, (algo, base, lossy)
    with pytest.raises(ValueError):
        comm1.set_loss(101.0)
    comm1.check()


@pytest.mark.gpu
@pytest.mark.multigpu
def test_watchdog_trips_instead_of_hanging():
    """A rank that never joins an all-reduce: the others' in-kernel waits time out (B200RNN_COMM_TIMEOUT_MS),
    check() raises on them, they exit non-zero, and nobody hangs."""
    import os
    import time
    from pytorch_distributed_rnn_b200.parallel.launch import launch
    n = min(torch.cuda.device_count(), 8)
    env = dict(os.environ, B200RNN_COMM_TIMEOUT_MS="300")
    t0 = time.time()
    res = launch([sys.executable, str(ROOT / "tests" / "watchdog_worker.py")], n, timeout=240, env=env, fail_fast=False)
    assert time.time() - t0 < 200
    for r, cp in enumerate(res):
        if r == n - 1:      # the rank that skipped the collective
            assert cp.returncode == 0, (r, cp.stderr[-2000:])
        else:
            assert cp.returncode == 3, (r, cp.returncode, cp.stdout[-2000:], cp.stderr[-2000:])
            assert "WATCHDOG_TRIPPED" in cp.stdout and f"rank(s) [{n - 1}]" in cp.stdout, cp.stdout[-2000:]


@pytest.mark.gpu
@pytest.mark.multigpu
def test_multirank_collectives_and_dp_invariance():
    from pytorch_distributed_rnn_b200.parallel.launch import launch
    n = min(torch.cuda.device_count(), 8)
    res = launch([sys.executable, str(ROOT / "tests" / "multigpu_worker.py")], n, timeout=600)
    for r, cp in enumerate(res):
        assert cp.returncode == 0, f"rank {r}:\n{cp.stdout[-3000:]}\n{cp.stderr[-3000:]}"
        line = [l for l in cp.stdout.splitlines() if l.startswith("RESULT ")][-1]
        out = json.loads(line[7:])
        bad = {k: v for k, v in out.items() if k.endswith("_ok") and v is not True}
        assert not bad, (r, bad, {k: v for k, v in out.items() if k.startswith("dp_param")})


def test_bucket_plan_native_matches_python():
    import ctypes as C
    from pytorch_distributed_rnn_b200 import _native
    from pytorch_distributed_rnn_b200.parallel.ddp import plan_buckets
    lib = _native.load()
    if lib is None:
        pytest.skip("native library not built")
    sizes = [4 * n for n in (1152, 4096, 128, 128, 4096, 4096, 128, 128, 192, 6, 300000, 7000000, 50)]
    arr = (C.c_int64 * len(sizes))(*sizes)
    out = (C.c_int32 * len(sizes))()
    nb = C.c_int32()
    assert lib.b200rnn_bucket_plan(arr, len(sizes), 1 << 20, 25 << 20, out, C.byref(nb)) == 0
    py = plan_buckets(sizes)
    assert nb.value == len(py)
    for b, ids in enumerate(py):
        assert all(out[i] == b for i in ids)


def test_native_bootstrap_two_processes():
    import subprocess
    import textwrap
    from pytorch_distributed_rnn_b200 import _native
    from pytorch_distributed_rnn_b200.parallel.launch import free_port
    if _native.load() is None:
        pytest.skip("native library not built")
    port = free_port()
    code = textwrap.dedent(f'''
        import sys
        sys.path.insert(0, {str(ROOT)!r})
        from pytorch_distributed_rnn_b200.parallel.bootstrap import NativeBootstrap
        r = int(sys.argv[1]); b = NativeBootstrap(r, 2, "127.0.0.1", {port}, timeout_ms=30000)
        out = b.allgather(b"rank%d" % r); b.barrier(); print(b"|".join(out).decode()); b.close()
    ''')
    ps = [subprocess.Popen([sys.executable, "-c", code, str(r)], stdout=subprocess.PIPE, text=True) for r in range(2)]
    outs = [p.communicate(timeout=90)[0].strip() for p in ps]
    assert outs == ["rank0|rank1", "rank0|rank1"]


@pytest.mark.gpu
@pytest.mark.multigpu
def test_gpu_parameter_server_one_sided(tmp_path):
    """K10: server + 2 workers as separate processes (the reference's launch style), NVLink one-sided."""
    import os
    import re
    import subprocess
    from pytorch_distributed_rnn_b200.data import synthetic
    from pytorch_distributed_rnn_b200.parallel.launch import free_port
    synthetic.write_dataset(tmp_path / "ds", train=384, valid=20, test=30)
    port = free_port()
    procs = []
    for rank in range(3):
        cmd = [sys.executable, "-m", "pytorch_distributed_rnn_b200", "--dataset-path", str(tmp_path / "ds"), "--epochs", "2",
               "--batch-size", "64", "--seed", "1", "--no-validation", "--learning-rate", "0.01",
               "parameter-server", "--world-size", "3", "--rank", str(rank), "--master-address", "127.0.0.1",
               "--master-port", str(port)]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                                      env={**os.environ, "PYTHONPATH": str(ROOT)}))
    outs = []
    try:
        outs = [p.communicate(timeout=300) for p in procs]
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    for p, (o, e) in zip(procs, outs):
        assert p.returncode == 0, o[-2000:] + e[-3000:]
    assert "one-sided NVLink mode" in outs[0][0] and "RPC shutdown on parameter server" in outs[0][0]
    for r in (1, 2):
        losses = [float(m) for m in re.findall(r"Loss: ([\d.]+)", outs[r][1])]
        assert len(losses) == 4 and losses[-1] < losses[0], losses   # 2 epochs x 128/64 batches, learning


@pytest.mark.gpu
@pytest.mark.multigpu
def test_cli_trainers_on_gpus_world_size_invariance(tmp_path):
    """`local` on 1 GPU vs `distributed` / `horovod` on 2 GPUs through the CLI: same mean step loss and
    summed correct count, performance line per rank, DDP checkpoint prefix."""
    import os
    import re
    from pytorch_distributed_rnn_b200.data import synthetic
    from pytorch_distributed_rnn_b200.parallel.launch import launch
    ds = tmp_path / "ds"
    synthetic.write_dataset(ds, train=960, valid=40, test=40)
    step_re = re.compile(r"Rank: (\d+)   Train Batch: (\d+)/(\d+) .*Loss: ([\d.]+)\tAcc: (\d+)/(\d+)")

    def run(n, trainer, extra=()):
        cmd = [sys.executable, "-m", "pytorch_distributed_rnn_b200", "--dataset-path", str(ds), "--epochs", "1",
               "--batch-size", "480", "--seed", "123456789", "--checkpoint-directory", str(tmp_path / f"ck_{trainer}"),
               "--history-path", str(tmp_path / f"h_{trainer}.json"), *extra, trainer]
        res = launch(cmd, n, timeout=300, env={"PYTHONPATH": str(ROOT)})
        steps = {}
        for r, cp in enumerate(res):
            assert cp.returncode == 0, f"{trainer} rank {r}:\n{cp.stdout[-1500:]}\n{cp.stderr[-3000:]}"
            for m in step_re.finditer(cp.stderr):
                steps.setdefault(int(m[2]), []).append((float(m[4]), int(m[5]), int(m[6])))
        return res, steps

    _, local = run(1, "local", ["--no-validation"])
    assert len(local) == 2
    for trainer in ("distributed", "horovod"):
        res, two = run(2, trainer)
        for b, ((l, c, n),) in local.items():
            assert len(two[b]) == 2
            assert abs(sum(x[0] for x in two[b]) / 2 - l) < 5e-6 and sum(x[1] for x in two[b]) == c and sum(x[2] for x in two[b]) == n
        assert all(f"{r}: Memory Usage:" in res[r].stderr for r in range(2))
        assert "Evaluation Epoch: 1/1" in res[0].stderr and "Test Evaluation" in res[0].stderr
        ck = torch.load(tmp_path / f"ck_{trainer}" / "best-model.pt", weights_only=False)
        assert all(k.startswith("module.") for k in ck["model_state"]) == (trainer == "distributed")


@pytest.mark.gpu
@pytest.mark.multigpu
def test_toy_examples_on_gpus():
    from pytorch_distributed_rnn_b200.parallel.launch import launch
    for script in ("example_ddp.py", "example_horovod.py", "example_distributed.py"):
        res = launch([sys.executable, str(ROOT / "examples" / script)], 2, timeout=300)
        for r, cp in enumerate(res):
            assert cp.returncode == 0, f"{script} rank {r}:\n{cp.stdout[-1500:]}\n{cp.stderr[-3000:]}"
        if script != "example_distributed.py":
            finals = [[l for l in cp.stdout.splitlines() if "parameters:" in l][-1].split("parameters:")[1] for cp in res]
            assert finals[0] == finals[1], (script, finals)
        else:
            assert all(f"Rank  {r}  has data  1.0" in res[r].stdout for r in range(2))
