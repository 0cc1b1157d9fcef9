#!/usr/bin/env python
"""A minimal cluster manager for an elastic job, all on one machine.

Two loopback aliases (127.0.0.1, 127.0.0.2) play two hosts. The manager
  1. starts a config server and publishes the first cluster (one worker per host),
  2. starts one EMPTY watch-mode runner per host (`kungfu-run -w -init-version -1`),
  3. starts the job with `kungfu-notify-start`,
  4. then drives the job's size by PUTting new clusters; the workers poll the
     config server each step (`resize_cluster_from_url`) and the runners spawn /
     reap their local workers as the cluster changes.

    python examples/cluster_manager.py [--sizes 2,4,3] [--period 2.0]

Reference: /root/reference/tests/go/cmd/kungfu-cluster-manager-example/ (docker
containers on a virtual network there; loopback aliases here) and
/root/reference/examples/mnist_elastic_docker/README.md.
"""
import argparse
import json
import os
import subprocess
import sys
import time
import urllib.request

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BIN = os.path.join(ROOT, 'kungfu_b200', 'bin')
HOSTS = ['127.0.0.1', '127.0.0.2']
SLOTS = 2


def worker(max_step):
    import torch

    import kungfu_b200.torch as kf
    step = int(kf.ops.all_reduce_fn(torch.tensor([0], dtype=torch.int64), 'max', name='sync-step')[0])
    while step < max_step:
        np_ = kf.current_cluster_size()
        s = kf.ops.all_reduce_fn(torch.ones(1000), name='grad')
        assert float(s[0]) == np_
        step += 1
        time.sleep(0.1)
        changed, detached = kf.ops.resize_cluster_from_url()
        if detached:
            print('rank %d leaves at step %d' % (kf.current_rank(), step), flush=True)
            return
        if changed:
            step = int(kf.ops.all_reduce_fn(torch.tensor([step], dtype=torch.int64), 'max', name='sync-step')[0])
            print('step %d: now rank %d of %d' % (step, kf.current_rank(), kf.current_cluster_size()), flush=True)
    print('rank %d finished %d steps with np=%d' % (kf.current_rank(), step, kf.current_cluster_size()), flush=True)


def cluster(n, runner_port, base):
    """n workers spread round-robin over the hosts."""
    per_host = [0] * len(HOSTS)
    workers = []
    for i in range(n):
        h = i % len(HOSTS)
        workers.append('%s:%d' % (HOSTS[h], base + per_host[h]))
        per_host[h] += 1
    return {'Runners': ['%s:%d' % (h, runner_port) for h in HOSTS], 'Workers': workers}


def put(url, c):
    req = urllib.request.Request(url, data=json.dumps(c).encode(), method='PUT')
    with urllib.request.urlopen(req, timeout=5) as r:
        return r.status


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--worker', action='store_true')
    ap.add_argument('--max-step', type=int, default=80)
    ap.add_argument('--sizes', default='2,4,3')
    ap.add_argument('--period', type=float, default=2.0)
    ap.add_argument('--port-base', type=int, default=26000)
    args = ap.parse_args()
    if args.worker:
        return worker(args.max_step)

    cfg_port, runner_port, base = args.port_base + 90, args.port_base + 80, args.port_base
    url = 'http://127.0.0.1:%d/config' % cfg_port
    sizes = [int(s) for s in args.sizes.split(',')]
    procs = [subprocess.Popen([os.path.join(BIN, 'kungfu-config-server'), '-port', str(cfg_port)],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)]
    try:
        for _ in range(50):
            try:
                put(url, cluster(sizes[0], runner_port, base))
                break
            except Exception:
                time.sleep(0.1)
        hostlist = ','.join('%s:%d' % (h, SLOTS) for h in HOSTS)
        runners = []
        for h in HOSTS:
            runners.append(subprocess.Popen(
                [os.path.join(BIN, 'kungfu-run'), '-q', '-w', '-init-version', '-1', '-self', h, '-H', hostlist,
                 '-port', str(runner_port), '-port-range', '%d-%d' % (base, base + 16), '-config-server', url,
                 '-timeout', '120s', sys.executable, os.path.abspath(__file__), '--worker', '--max-step',
                 str(args.max_step)]))
        procs += runners
        subprocess.check_call([os.path.join(BIN, 'kungfu-notify-start'), '-config-server', url])
        for n in sizes[1:]:
            time.sleep(args.period)
            print('[manager] resize to %d -> HTTP %d' % (n, put(url, cluster(n, runner_port, base))), flush=True)
        rcs = [r.wait(timeout=150) for r in runners]
        print('[manager] runners exited with', rcs, flush=True)
        return 0 if all(rc == 0 for rc in rcs) else 1
    finally:
        for p in procs:
            if p.poll() is None:
                p.terminate()
        for p in procs:
            try:
                p.wait(timeout=10)
            except Exception:
                p.kill()


if __name__ == '__main__':
    sys.exit(main())
