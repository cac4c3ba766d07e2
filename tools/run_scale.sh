#!/bin/bash
# One multi-GPU session: multi-rank tests, 1/2/4/8 scaling of both bench arms, all-reduce sweeps.
# Usage (8-GPU box): bash tools/run_scale.sh [NGPU]
set -u
N=${1:-8}
OUT=gpurun_out/scale
mkdir -p $OUT
PORT=29600
run() { # n, extra args..., outfile
  local n=$1; shift; local out=$1; shift
  PORT=$((PORT+1))
  if [ "$n" = 1 ]; then timeout 300 python bench.py --gpus 1 "$@" > $OUT/$out.json 2> $OUT/$out.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $OUT/$out.json 2> $OUT/$out.err; fi
  echo "$out rc=$? $(grep -o '"ms_per_step": [0-9.]*' $OUT/$out.json | head -1) $(grep -o '"value": [0-9.]*' $OUT/$out.json | head -1)"
}
timeout 600 python -m pytest tests/test_comm_gpu.py -m gpu -q -x --timeout 500 > $OUT/pytest_comm.log 2>&1; echo "pytest comm rc=$? $(tail -1 $OUT/pytest_comm.log)"
python -m pytorch_distributed_rnn_b200.parallel.launch -n $N -m pytorch_distributed_rnn_b200.parallel.check > $OUT/check.log 2>&1; echo "check rc=$? ok_count=$(grep -c '"ok": true' $OUT/check.log)"
for n in 1 2 4 8; do
  [ $n -le $N ] || continue
  run $n own_n$n --steps 200 --warmup 10
  if [ "${SKIP_REF:-0}" = 0 ] || [ $n = $N ]; then run $n ref_n$n --impl reference --steps 30 --warmup 3; fi
done
B200RNN_COMM=nccl run $N own_n${N}_ncclbaseline --steps 200 --warmup 10
PORT=$((PORT+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench/allreduce_sweep.py --delay-us 0,50 --out $OUT/allreduce_sweep_n$N.json > $OUT/allreduce_sweep_n$N.log 2>&1; echo "sweep n$N rc=$?"
if [ $N -gt 2 ] && [ "${SKIP_N2_SWEEP:-0}" = 0 ]; then
PORT=$((PORT+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench/allreduce_sweep.py --out $OUT/allreduce_sweep_n2.json > $OUT/allreduce_sweep_n2.log 2>&1; echo "sweep n2 rc=$?"
fi
tail -4 $OUT/allreduce_sweep_n$N.log | cut -c1-400
