"""Cold-start and resize latency instrumentation: prints wall-clock since the
job / process start at the points the reference logs with `_log_event`
(/root/reference/benchmarks/scaling/benchmark_kungfu_scaling.py,
/root/reference/srcs/python/kungfu/_utils.py:44-50).

    kungfu-run -np 4 python benchmarks/scaling/cold_start.py [--cuda]"""
import os
import sys
import time

T_IMPORT0 = time.time()
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

T_TORCH = time.time()
import kungfu_b200.torch as kf  # noqa: E402

rank = kf.current_rank()  # starts the control plane (first barrier)
T_PEER = time.time()
job_t0 = float(os.environ.get('KUNGFU_JOB_START_TIMESTAMP', T_IMPORT0))
proc_t0 = float(os.environ.get('KUNGFU_PROC_START_TIMESTAMP', T_IMPORT0))


def log_event(name, t):
    print('[rank %d] %-28s +%.3fs since job start, +%.3fs since process start' % (rank, name, t - job_t0, t - proc_t0), flush=True)


log_event('python started', T_IMPORT0)
log_event('torch imported', T_TORCH)
log_event('peer started + first barrier', T_PEER)
if '--cuda' in sys.argv and torch.cuda.is_available():
    torch.cuda.set_device(kf.get_cuda_index())
    x = torch.ones(1 << 20, device='cuda')
    torch.cuda.synchronize()
    log_event('CUDA context ready', time.time())
    kf.ops.all_reduce_fn(x)
    torch.cuda.synchronize()
    log_event('peer memory mapped + 1st all-reduce', time.time())
else:
    kf.ops.all_reduce_fn(torch.ones(1 << 20))
    log_event('first CPU all-reduce', time.time())
kf.run_barrier()
