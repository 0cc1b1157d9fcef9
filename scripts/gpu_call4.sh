#!/bin/bash
python -m pytest tests/test_gpu_multi.py -x -q -k "peer_memory_path and not ipc" 2>&1 | tail -25
bash scripts/ab_overlap.sh 2
echo "== traced overlap=1 N=2"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --steps 20 --warmup 5 --skip-e2e --same-box-baseline none --overlap 1 --trace 2>&1 | grep -E "^\[trace\]" | head -30 | cut -c1-200
