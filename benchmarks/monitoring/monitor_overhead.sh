#!/bin/bash
# Overhead of the gradient-noise-scale and gradient-variance monitors relative to
# plain S-SGD (reference: /root/reference/benchmarks/monitoring/benchmark.py:75-90,
# run_benchmark.sh: --kf-optimizer {noise-scale,variance} --interval N).
# usage: benchmarks/monitoring/monitor_overhead.sh [np] [model] [batch]
set -e
cd "$(dirname "$0")/../.."
NP=${1:-8}; MODEL=${2:-resnet50}; BS=${3:-128}
for OPT in ssgd gns variance; do
    BASE=fused; [ "$OPT" = variance ] && BASE=sgd
    kungfu_b200/bin/kungfu-run -q -np $NP -allow-nvlink python benchmarks/train_bench.py \
        --model $MODEL --optimizer $OPT --base $BASE --batch-size $BS --steps 20 --warmup 5 2>&1 | grep RESULT
done
