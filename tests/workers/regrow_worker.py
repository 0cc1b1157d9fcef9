"""1 -> 2 -> 1 -> 2 ...: the cluster comes back to a size and a rank it had before, with a NEW peer on a
previously used address. The device layer's control exchange (numbered message names) must start over
with every membership, or the survivor and the fresh joiner wait for differently named messages."""
import sys

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf
from kungfu_b200.utils import step_based_schedule

schedule, max_step = sys.argv[1], int(sys.argv[2])
step = int(kf.ops.all_reduce_fn(torch.tensor([0], dtype=torch.int64), 'max')[0])
while step < max_step:
    np_ = kf.current_cluster_size()
    if np_ > 1:
        x = kfpy.control_exchange()
        for k in range(3):  # what PeerComm does when it is created: a few numbered all-gathers
            got = x.all_gather_bytes(bytes([kf.current_rank(), k]))
            assert got == [bytes([r, k]) for r in range(np_)], got
    step += 1
    want = step_based_schedule(schedule, step, default=np_, strict=False)
    changed, detached = kf.resize_cluster(want)
    if detached:
        print('detached at step %d' % step)
        sys.exit(0)
    if changed:
        step = int(kf.ops.all_reduce_fn(torch.tensor([step], dtype=torch.int64), 'max')[0])
print('OK regrow worker rank %d done at step %d' % (kf.current_rank(), step))
