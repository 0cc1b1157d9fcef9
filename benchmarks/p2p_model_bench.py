"""Host-path model exchange: what the prefetching double buffer buys.

    kungfu-run -np 2 python benchmarks/p2p_model_bench.py [--model-mb 97.6] [--compute-ms 60] [--steps 20]

Every step "computes" for --compute-ms (a sleep, as the reference's fake trainers do) and then averages
its model with a peer's. `ModelAveraging` pulls the peer's fused model inside the step; `AsyncModelAveraging`
blends with the newest snapshot that has fully arrived while the runtime's worker threads pull the next one
under the compute (native double buffer, csrc/capi/capi.cpp kf_prefetch_*). Prints the step times and the
share of the pull that left the critical path.

Reference: /root/reference/srcs/cpp/src/tensorflow/ops/cpu/peer_to_peer.cpp:72-258 (ModelAveraging vs
AsyncModelAveraging), /root/reference/tests/go/cmd/kungfu-bench-p2p (request rate).
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import kungfu_b200.torch as kf  # noqa: E402
from kungfu_b200.torch.ops import p2p  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--model-mb', type=float, default=97.6, help='fp32 model size (ResNet-50 = 97.6 MiB)')
ap.add_argument('--compute-ms', type=float, default=60.0)
ap.add_argument('--steps', type=int, default=20)
args = ap.parse_args()

rank, np_ = kf.current_rank(), kf.current_cluster_size()
n = int(args.model_mb * (1 << 20) / 4)
model = [('w', torch.full((n, ), float(rank)))]


def run(op, publish_every=1):
    times = []
    for step in range(args.steps + 2):
        t0 = time.perf_counter()
        time.sleep(args.compute_ms / 1e3)  # forward + backward
        op()
        if step % publish_every == 0:
            op.save()                       # publish the new model for the peers
        times.append(time.perf_counter() - t0)
    return sorted(times[2:])[len(times[2:]) // 2] * 1e3  # median, first two steps are warm-up


blocking = p2p.ModelAveraging(model, 'roundrobin')
blocking.save()
kf.run_barrier()
t_block = run(blocking)
kf.run_barrier()
prefetching = p2p.AsyncModelAveraging(model, 'roundrobin')
t_async = run(prefetching)
fresh = prefetching.fresh
prefetching.close()
kf.run_barrier()
base = args.compute_ms
# what a step costs besides the compute: pull + blend + publish (blocking) vs blend + publish (prefetching)
tb = float(kf.ops.all_reduce_fn(torch.tensor([t_block]), 'max')[0])
ta = float(kf.ops.all_reduce_fn(torch.tensor([t_async]), 'max')[0])
if rank == 0:
    print('RESULT: p2p model exchange np=%d model=%.1f MiB compute=%.0f ms: ModelAveraging %.1f ms/step, '
          'AsyncModelAveraging %.1f ms/step (%d background pulls completed); exchange overhead %.1f -> %.1f ms' %
          (np_, args.model_mb, base, tb, ta, fresh, tb - base, ta - base), flush=True)
