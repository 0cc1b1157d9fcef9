#!/bin/bash
# 1 GPU: cooperative BN validation + A/B, BN kernel table, ncu of the emulated communication kernels and the BN kernels
O=gpurun_out/g1; mkdir -p $O
python -m pytest tests/test_gpu_bn_act.py tests/test_gpu_models.py tests/test_gpu_e2e.py tests/test_gpu_optim.py tests/test_gpu_collectives.py -x -q 2>&1 | tail -4
for c in 0 1 0 1; do KUNGFU_BN_COOP=$c python bench.py --steps 40 --warmup 5 --skip-e2e --same-box-baseline none 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('KUNGFU_BN_COOP=$c: %.3f ms  %.0f img/s loss %.4f' % (d['ms_per_step'], d['value'], d['loss']))"; done
for c in 0 1; do echo "== bn_kernel_bench KUNGFU_BN_COOP=$c"; KUNGFU_BN_COOP=$c python benchmarks/bn_kernel_bench.py 2>&1 | tail -12 | tee $O/bn_kernel_bench_coop$c.txt; done
python benchmarks/ncu_comm_single_gpu.py --reps 1 > $O/ncu_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_allreduce_twoshot|k_ssgd_fused|k_sma_fused|k_pair_avg|k_broadcast|k_grad_variance' -c 6 -o $O/comm_emulated -f python benchmarks/ncu_comm_single_gpu.py --reps 1 > $O/ncu_comm.log 2>&1; tail -3 $O/ncu_comm.log
ncu --set full --clock-control none --import-source on -k regex:'k_bn_fwd_coop|k_bn_bwd_coop' -s 40 -c 4 -o $O/bn_coop -f python benchmarks/bn_kernel_bench.py > $O/ncu_bn.log 2>&1; tail -3 $O/ncu_bn.log
ls -la $O
