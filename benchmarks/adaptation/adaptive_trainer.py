"""All-reduce-only trainer that follows a cluster-size schedule through
resize_cluster (config-server driven when kungfu-run has one).

    kungfu-run -w -builtin-config-port 9100 -np 3 -H 127.0.0.1:16 \
        python benchmarks/adaptation/adaptive_trainer.py --schedule 3:2,3:4,3:16,3:1

Reference: /root/reference/benchmarks/adaptation/adaptive_trainer.py:16-19 and
bench-adaptation.sh (schedules `3:2,3:4,3:16,3:1` and `1:16,1:1,1:16,1:1`).
Prints the resize latency of every membership change."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

import kungfu_b200.torch as kf

ap = argparse.ArgumentParser()
ap.add_argument('--schedule', default='3:2,3:4,3:16,3:1')
ap.add_argument('--model-mb', type=float, default=97.6, help='fake gradient size (ResNet-50 fp32 = 97.6 MiB)')
ap.add_argument('--max-step', type=int, default=0)
args = ap.parse_args()


def parse_schedule(config):
    """'steps:size,steps:size,...' (this benchmark's format, unlike the size:steps
    of step_based_schedule) -> [(begin, end, size)], total steps."""
    out, t = [], 0
    for part in config.split(','):
        n, val = map(int, part.split(':'))
        out.append((t, t + n, val))
        t += n
    return out, t


def size_at(schedule, step, default):
    for b, e, v in schedule:
        if b <= step < e:
            return v
    return default


schedule, total = parse_schedule(args.schedule)
max_step = args.max_step or total
grad = torch.ones(int(args.model_mb * (1 << 20) / 4))


def sync_step(step):
    return int(kf.ops.all_reduce_fn(torch.tensor([step], dtype=torch.int64), 'max', name='step')[0])


step = sync_step(0)
while step < max_step:
    np_ = kf.current_cluster_size()
    t0 = time.time()
    out = kf.ops.all_reduce_fn(grad, name='grad')
    dt = time.time() - t0
    assert float(out[0]) == np_
    if kf.current_rank() == 0:
        print('step %d np %d all-reduce %.1f MiB in %.3fs (%.2f GiB/s)' % (step, np_, args.model_mb, dt, args.model_mb / 1024 / dt), flush=True)
    step += 1
    want = size_at(schedule, step, np_)
    if want != np_:
        t0 = time.time()
        changed, detached = kf.resize_cluster(want)
        if detached:
            print('worker %x detached at step %d' % (kf.uid(), step), flush=True)
            sys.exit(0)
        if changed:
            step = sync_step(step)
            if kf.current_rank() == 0:
                print('RESIZE %d -> %d took %.3fs' % (np_, want, time.time() - t0), flush=True)
print('worker %x done' % kf.uid())
