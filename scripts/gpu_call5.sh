#!/bin/bash
echo "== N=1 top kernels"
KUNGFU_TRACE_TOP=45 CUDA_VISIBLE_DEVICES=0 python bench.py --gpus 1 --steps 20 --warmup 5 --skip-e2e --same-box-baseline none --overlap 0 --trace 2>&1 | grep -E "^\[trace-top\]|^\[trace\] 4" | cut -c1-190
echo "== sampler effect at N=2 (NVML in-process): interval 100 ms vs 5000 ms, overlap 0"
for iv in 100 5000 100 5000; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29700+RANDOM%200)) bench.py --gpus 2 --steps 40 --warmup 5 --skip-e2e --same-box-baseline none --overlap 0 --clock-interval-ms $iv 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('interval $iv: %.3f ms %s' % (d['ms_per_step'], d['clocks']))"
done
echo "== N=1 alone on gpu0, then gpu1 alone"
for g in 0 1; do CUDA_VISIBLE_DEVICES=$g python bench.py --steps 40 --warmup 5 --skip-e2e --same-box-baseline none --overlap 0 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('gpu$g alone: %.3f ms' % d['ms_per_step'])"; done
echo "== gpu0 bench while gpu1 runs an unrelated busy loop (no sampler on gpu1)"
CUDA_VISIBLE_DEVICES=1 python -c "
import torch,time
a=torch.randn(8192,8192,device='cuda',dtype=torch.bfloat16)
t=time.time()
while time.time()-t<45: b=a@a
torch.cuda.synchronize()" &
sleep 5
CUDA_VISIBLE_DEVICES=0 python bench.py --steps 40 --warmup 5 --skip-e2e --same-box-baseline none --overlap 0 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('gpu0 with busy neighbour: %.3f ms' % d['ms_per_step'])"
wait
