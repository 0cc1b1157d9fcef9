"""1 -> 3 -> 1 -> 3 ...: at every visit to size 1 the survivor reports its open descriptors, threads and
resident memory. Connections to removed peers, their reader threads and their buffers must go away with
the membership that used them (reference: rchannel/client/connection_pool.go:29-51, peer.go:144-166)."""
import os
import sys
import time

import torch

import kungfu_b200.torch as kf
from kungfu_b200.utils import step_based_schedule

schedule, max_step = sys.argv[1], int(sys.argv[2])
step = int(kf.ops.all_reduce_fn(torch.tensor([0], dtype=torch.int64), 'max')[0])
while step < max_step:
    np_ = kf.current_cluster_size()
    kf.ops.all_reduce_fn(torch.ones(100000))
    if np_ > 1:  # p2p connections too
        kf.ops.save_variable(torch.ones(10), 'v')
        kf.run_barrier()
        ok, _ = kf.ops.request_variable((kf.current_rank() + 1) % np_, 'v', torch.ones(10))
        assert ok
    elif kf.current_rank() == 0:
        # the peers that just left are still being disconnected by their reader threads: take the floor of a second
        fds, threads = 1 << 30, 1 << 30
        for _ in range(10):
            fds = min(fds, len(os.listdir('/proc/self/fd')))
            threads = min(threads, len(os.listdir('/proc/self/task')))
            time.sleep(0.1)
        rss = int(open('/proc/self/statm').read().split()[1]) * os.sysconf('SC_PAGE_SIZE') >> 20
        print('LEAK step %d fds %d threads %d rss_mib %d' % (step, fds, threads, rss), flush=True)
    step += 1
    want = step_based_schedule(schedule, step, default=np_, strict=False)
    changed, detached = kf.resize_cluster(want)
    if detached:
        sys.exit(0)
    if changed:
        step = int(kf.ops.all_reduce_fn(torch.tensor([step], dtype=torch.int64), 'max')[0])
print('OK leak worker done at step %d' % step, flush=True)
