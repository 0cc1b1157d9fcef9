#!/bin/bash
# 4-GPU box: N=4 points of the other models, then N=2 and two N=1 runs side by side on disjoint GPUs.
O=gpurun_out/n4; mkdir -p $O
run() { # run <gpus csv> <np> <portbase> <args...>
  local gpus=$1 np=$2 pb=$3; shift 3
  CUDA_VISIBLE_DEVICES=$gpus kungfu_b200/bin/kungfu-run -q -allow-nvlink -H 127.0.0.1:8 -np $np -port-range $pb-$((pb+60)) -port $((pb+90)) python benchmarks/train_bench.py "$@" 2>&1 | grep RESULT | sed -E 's/\x1b\[[0-9;]*m//g; s/^\[[^]]*\] //'
}
DN="--model densenet121 --optimizer sma --batch-size 128 --steps 30"
DNG="--model densenet121 --optimizer sma --batch-size 128 --steps 30 --graph"
BP="--model bert_base --optimizer pair --base fused-adamw --batch-size 32 --graph --pair-device-select 1 --fused-ln 1 --fused-grads 1 --fused-linear 1 --pad-vocab 8 --mlm-positions 20 --steps 40"
VG="--model vgg16 --optimizer ssgd --batch-size 128 --steps 20"
echo "== N=4"; date +%T
run 0,1,2,3 4 21000 $DN | tee -a $O/curves.txt
run 0,1,2,3 4 21000 $DNG | tee -a $O/curves.txt
run 0,1,2,3 4 21000 $BP | tee -a $O/curves.txt
run 0,1,2,3 4 21000 $VG | tee -a $O/curves.txt
echo "== N=2 (gpus 0,1) || N=1 (gpu 2) || N=1 (gpu 3)"; date +%T
(run 0,1 2 21000 $DNG >> $O/curves.txt) & (run 2 1 22000 $DNG >> $O/curves.txt) & (run 3 1 23000 $DN >> $O/curves.txt) & wait
(run 0,1 2 21000 $BP >> $O/curves.txt) & (run 2 1 22000 $BP >> $O/curves.txt) & (run 3 1 23000 $VG >> $O/curves.txt) & wait
(run 0,1 2 21000 $VG >> $O/curves.txt) & (run 2,3 2 22000 $DN >> $O/curves.txt) & wait
date +%T
cat $O/curves.txt
