#!/bin/bash
# 4-GPU box: the whole GPU test suite (np = 2 and 4 multi-GPU tests, elastic 1->2->4->2->1, two loopback "hosts"), then the curves
python -m pytest tests -m gpu -x -q > gpurun_out/gputest_4gpu.log 2>&1; tail -6 gpurun_out/gputest_4gpu.log
bash scripts/gpu_call_n4.sh
