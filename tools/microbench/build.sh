#!/bin/bash
# Build the standalone microbenchmarks into build/ (git-ignored; travels to the GPU box with gpurun).
set -e
cd "$(dirname "$0")/../.."
mkdir -p build
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo"
nvcc $FLAGS -o build/tma_stream_bench tools/microbench/tma_stream_bench.cu -lcuda
nvcc $FLAGS -o build/tma_issue_probe tools/microbench/tma_issue_probe.cu
nvcc $FLAGS -o build/tma_timeline tools/microbench/tma_timeline.cu
nvcc $FLAGS -I csrc -I include -o build/mma_issue_probe tools/microbench/mma_issue_probe.cu
ls -la build/tma_stream_bench build/tma_issue_probe build/tma_timeline build/mma_issue_probe
