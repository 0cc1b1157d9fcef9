#!/bin/bash
# One 8-GPU session: correctness at np=8, overlap A/B, box effects, kernel tables, baselines, other models, elastic growth.
# Everything is written under gpurun_out/n8/.
O=gpurun_out/n8; mkdir -p $O
RUN="kungfu_b200/bin/kungfu-run -q -allow-nvlink -H 127.0.0.1:8"
tr8() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29800 + RANDOM % 100)) "$@"; }
echo "== [1] multi-GPU correctness at np=8 (NVLS required)"; date +%T
python -m pytest tests/test_gpu_multi.py -x -q -k "peer_memory_path and 8" > $O/pytest_np8.log 2>&1; tail -3 $O/pytest_np8.log
echo "== [2] bench N=8: overlap 0 / 1"; date +%T
for ov in 0 1; do tr8 bench.py --gpus 8 --steps 40 --warmup 5 --skip-e2e --same-box-baseline none --overlap $ov 2>/dev/null | grep '^{' > $O/bench8_ov$ov.json; python -c "import json; d=json.load(open('$O/bench8_ov$ov.json')); print('overlap $ov: %.3f ms %.0f img/s check=%s' % (d['ms_per_step'], d['value'], d['check']))"; done
echo "== [3] 8 independent single-GPU runs side by side (box effect without any communication)"; date +%T
for g in 0 1 2 3 4 5 6 7; do (CUDA_VISIBLE_DEVICES=$g python bench.py --steps 40 --warmup 5 --skip-e2e --same-box-baseline none --overlap 0 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('gpu$g independent, 7 busy neighbours: %.3f ms' % d['ms_per_step'])" >> $O/independent8.txt) & done; wait; sort $O/independent8.txt
echo "== [4] kernel tables np=8"; date +%T
$RUN -np 8 python benchmarks/fused_step_bench.py --out $O/fused_kernels.jsonl 2>&1 | grep ROW | cut -c1-220
$RUN -np 8 python benchmarks/allreduce_bench.py --nccl --blocks 16,32 --iters 15 --min-bytes 16384 --json $O/allreduce_np8.json > $O/allreduce_np8.txt 2>&1; tail -25 $O/allreduce_np8.txt
echo "== [5] full bench line N=8 (with e2e, check and the same-box DDP baseline)"; date +%T
tr8 bench.py --gpus 8 --steps 40 --warmup 5 2>/dev/null | grep '^{' > $O/bench8_full.json; cut -c1-1500 $O/bench8_full.json
echo "== [6] stand-ins for the reference's NCCL path, N=8"; date +%T
for m in ref-nccl-fused ref-nccl-unfused; do $RUN -np 8 python benchmarks/baselines.py --mode $m --steps 20 --warmup 5 2>&1 | grep RESULT | tee -a $O/baselines8.txt; done
echo "== [7] other models N=8"; date +%T
$RUN -np 8 python benchmarks/train_bench.py --model densenet121 --optimizer sma --batch-size 128 --steps 30 2>&1 | grep RESULT | tee -a $O/models8.txt
$RUN -np 8 python benchmarks/train_bench.py --model bert_base --optimizer pair --base fused-adamw --batch-size 32 --graph --pair-device-select 1 --fused-ln 1 --fused-grads 1 --fused-linear 1 --pad-vocab 8 --mlm-positions 20 --steps 40 2>&1 | grep RESULT | tee -a $O/models8.txt
echo "== [8] elastic ResNet-50, noise-scale driven growth 1 -> 2 -> 4 -> 8"; date +%T
kungfu_b200/bin/kungfu-run -q -w -builtin-config-port 9100 -np 1 -H 127.0.0.1:8 -allow-nvlink -timeout 300s python examples/elastic_resnet_gns.py --steps 160 --batch-size 128 --check-every 10 --report-every 5 > $O/elastic_1_2_4_8.txt 2>&1; grep -E "^step|resize|finished" $O/elastic_1_2_4_8.txt | cut -c1-260 | tail -45
date +%T
