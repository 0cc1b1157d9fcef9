#!/bin/bash
# N=2 diagnostics: multi-GPU worker, two independent single-GPU runs side by side, traced overlap A/B
python -m pytest tests/test_gpu_multi.py -x -q -k "peer_memory_path and not ipc" 2>&1 | tail -25
echo "== two independent 1-GPU runs side by side (is the N=2 gap GPU-to-GPU variance?)"
(CUDA_VISIBLE_DEVICES=0 python bench.py --steps 40 --warmup 5 --skip-e2e --same-box-baseline none 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('gpu0 alone-but-concurrent %.3f ms' % d['ms_per_step'])") &
(CUDA_VISIBLE_DEVICES=1 KUNGFU_PORT_BASE=31000 python bench.py --steps 40 --warmup 5 --skip-e2e --same-box-baseline none 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('gpu1 alone-but-concurrent %.3f ms' % d['ms_per_step'])") &
wait
echo "== traced N=2"
for ov in 0 1; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29700+ov)) bench.py --gpus 2 --steps 20 --warmup 5 --skip-e2e --same-box-baseline none --overlap $ov --trace 2>&1 | grep -E "^\[trace\]|^\{" | cut -c1-400
done
echo "== traced N=1"
python bench.py --gpus 1 --steps 20 --warmup 5 --skip-e2e --same-box-baseline none --overlap 0 --trace 2>&1 | grep -E "^\[trace\]" | cut -c1-300
