#!/bin/bash
# Loopback elastic soak: the reference's two schedules
# (/root/reference/benchmarks/adaptation/bench-adaptation.sh).
set -e
cd "$(dirname "$0")/.."
for SCHED in 3:2,3:4,3:16,3:1 1:16,1:1,1:16,1:1; do
    NP0=$(echo $SCHED | cut -d, -f1 | cut -d: -f2)   # schedule entries are steps:size
    kungfu_b200/bin/kungfu-run -q -w -builtin-config-port 9199 -np $NP0 -H 127.0.0.1:16 \
        python benchmarks/adaptation/adaptive_trainer.py --schedule $SCHED --model-mb 4
done
