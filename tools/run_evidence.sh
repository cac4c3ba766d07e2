#!/bin/bash
# Evidence run on an N-GPU box: multi-GPU test suite, all-reduce sweep (+ measured algorithm table), and the
# headline / scaled benches for both arms.  Usage: tools/run_evidence.sh N [tests|nosweep]  -> gpurun_out/r2_nN/
N=${1:-2}
OUT=gpurun_out/r2_n$N
mkdir -p $OUT
T="timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ "$2" != "notests" ]; then
  timeout 400 python -m pytest tests -m multigpu -x -q > $OUT/pytest_multigpu.log 2>&1
  tail -n 6 $OUT/pytest_multigpu.log
fi
if [ "$2" != "nosweep" ]; then
  $T bench/allreduce_sweep.py --out $OUT/allreduce_sweep.json --md $OUT/allreduce_sweep.md --write-table $OUT/allreduce_table.json > $OUT/allreduce_sweep.log 2>&1
  $T bench/allreduce_sweep.py --min-bytes 56600 --max-bytes 56600 --rows 64 --iters 30 --out $OUT/allreduce_56k_rows64.json --md $OUT/allreduce_56k_rows64.md > $OUT/allreduce_56k.log 2>&1
  cat $OUT/allreduce_sweep.md $OUT/allreduce_56k_rows64.md; cat $OUT/allreduce_table.json
fi
for impl in reference own; do
  $T bench.py --gpus $N --impl $impl --steps 50 --warmup 10 > $OUT/bench_motion_$impl.json 2> $OUT/bench_motion_$impl.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$OUT/bench_motion_$impl.json") if l.startswith("{")][-1])
    print("motion $impl N=$N:", round(d["ms_per_step"], 4), "ms/step; e2e", round(d["e2e"]["ms_per_step"], 4), "; epoch", d.get("epoch"))
except Exception as e:
    print("motion $impl failed:", e)
PY
done
B200RNN_COMM=nccl $T bench.py --gpus $N --steps 50 --warmup 10 > $OUT/bench_motion_own_nccl.json 2> $OUT/bench_motion_own_nccl.err
python -c "
import json
try:
    d = json.loads([l for l in open('$OUT/bench_motion_own_nccl.json') if l.startswith('{')][-1]); print('motion own (NCCL comm) N=$N:', round(d['ms_per_step'], 4), 'ms/step')
except Exception as e: print('nccl A/B failed', e)"
for impl in reference own; do
  $T bench.py --gpus $N --config scaled --impl $impl --steps 8 --warmup 3 > $OUT/bench_scaled_$impl.json 2> $OUT/bench_scaled_$impl.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$OUT/bench_scaled_$impl.json") if l.startswith("{")][-1])
    print("scaled $impl N=$N:", round(d["ms_per_step"], 3), "ms/step; e2e", round(d["e2e"]["ms_per_step"], 3), d.get("overlap_timeline_us"))
except Exception as e:
    print("scaled $impl failed:", e)
PY
done
if [ "$3" = "grid" ]; then   # the reference's rank-scaling experiment (fabfile.py:48-66), runs packed side by side on the box
  timeout 900 python bench/sweep.py --dataset-path /tmp/b200rnn_motion --results $OUT/results_ranks.json --ranks ${4:-1,2,4} --repeats ${5:-2} --pack $N > $OUT/sweep_ranks.log 2>&1
  python bench/analysis.py $OUT/results_ranks.json --plot $OUT/results_ranks.png > $OUT/results_ranks.md 2>> $OUT/sweep_ranks.log
  tail -n 40 $OUT/results_ranks.md
fi
if [ "$3" = "network" ]; then   # the reference's network experiment (fabfile.py:130-191): injected delay before every flag publication
  timeout 900 python bench/sweep.py --dataset-path /tmp/b200rnn_motion --results $OUT/results_network.json --ranks $N --network --delays ${4:-0,1,2,5,10,25,50,100,200,300,400} > $OUT/sweep_network.log 2>&1
  python bench/analysis.py $OUT/results_network.json --network --plot $OUT/results_network.png > $OUT/results_network.md 2>> $OUT/sweep_network.log
  tail -n 30 $OUT/results_network.md
fi
tail -n 3 $OUT/*.err | grep -v "^$" | tail -n 30
