#!/bin/bash
# Final 2-GPU check of the tree: headline bench (own arm), the multi-rank worker suite, and the network experiment
# (all delays and losses of the reference's lists through both trainers).  Usage: gpurun --gpus 2 -- bash tools/run_final_n2.sh [network]
OUT=gpurun_out/r2_final_n2; mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
if [ "$1" != "network" ]; then
  timeout 70 $T bench.py --gpus 2 --steps 50 --warmup 10 > $OUT/bench_motion_own.json 2> $OUT/bench_motion_own.err; echo "bench rc=$?"; cut -c1-260 $OUT/bench_motion_own.json
  timeout 110 python -m pytorch_distributed_rnn_b200.parallel.launch -n 2 tests/multigpu_worker.py > $OUT/multigpu_worker.log 2>&1; echo "worker rc=$?"; tail -n 4 $OUT/multigpu_worker.log | cut -c1-200
fi
rm -f $OUT/results_network.json
timeout 150 python bench/sweep.py --dataset-path /tmp/b200rnn_motion --results $OUT/results_network.json --ranks 2 --network > $OUT/sweep_network.log 2>&1; echo "sweep rc=$?"
python bench/analysis.py $OUT/results_network.json --network --rule delay --plot $OUT/results_network_delay.png > $OUT/results_network_delay.md 2>> $OUT/sweep_network.log; cat $OUT/results_network_delay.md
python bench/analysis.py $OUT/results_network.json --network --rule loss --plot $OUT/results_network_loss.png > $OUT/results_network_loss.md 2>> $OUT/sweep_network.log; cat $OUT/results_network_loss.md
