#!/bin/bash
# 2-GPU dry run of every step of the 8-GPU session (short settings)
O=gpurun_out/n2; mkdir -p $O
RUN="kungfu_b200/bin/kungfu-run -q -allow-nvlink -H 127.0.0.1:8"
set -x
$RUN -np 2 python benchmarks/fused_step_bench.py --iters 10 --out $O/fused_kernels.jsonl 2>&1 | tail -22 | cut -c1-230
$RUN -np 2 python benchmarks/allreduce_bench.py --nccl --blocks 32 --iters 5 --min-bytes 1048576 --max-bytes 16777216 --json $O/ar.json 2>&1 | tail -12
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29810 bench.py --gpus 2 --steps 10 --warmup 3 2>&1 | grep -E '^\{|Error|error' | cut -c1-1600
$RUN -np 2 python benchmarks/baselines.py --mode ref-nccl-fused --steps 5 --warmup 3 2>&1 | tail -2
$RUN -np 2 python benchmarks/train_bench.py --model densenet121 --optimizer sma --batch-size 64 --steps 10 2>&1 | tail -3
$RUN -np 2 python benchmarks/train_bench.py --model bert_base --optimizer pair --base fused-adamw --batch-size 32 --graph --pair-device-select 1 --fused-ln 1 --fused-grads 1 --fused-linear 1 --pad-vocab 8 --mlm-positions 20 --steps 10 2>&1 | tail -3
$RUN -np 2 python benchmarks/train_bench.py --model bert_base --optimizer pair --base adamw --batch-size 32 --mlm-positions 20 --steps 5 2>&1 | tail -3
kungfu_b200/bin/kungfu-run -q -w -builtin-config-port 9100 -np 1 -H 127.0.0.1:8 -allow-nvlink -timeout 200s python examples/elastic_resnet_gns.py --steps 40 --batch-size 64 --check-every 5 --report-every 5 --max-size 2 2>&1 | grep -E "^step|resize|finished|Error|error" | cut -c1-260 | tail -20
