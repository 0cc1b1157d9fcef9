#!/usr/bin/env python
"""A virtual multi-node elastic cluster on one machine, under a random config client.

    python benchmarks/adaptation/virtual_cluster.py [--nodes 3] [--cap 2] [--np 2] [--ttl 10] [--period 0.5]

Loopback aliases 127.0.0.1 .. 127.0.0.<nodes> play the nodes. The script starts

  * one `kungfu-config-server`,
  * one `kungfu-run -w -k -np <np> -H <all nodes> -self <node>` per node (every
    runner derives the same first cluster and starts its local part of it),
  * one `kungfu-config-client -shuffle`, which PUTs a random SUBSET of the slots
    in random order every period for ttl seconds: ranks move, nodes empty out
    and fill up again, sometimes not a single worker survives a change,

then PUTs the full cluster, waits until every slot's worker has finished its
steps, and asks the (kept-alive) runners to exit over their control channel.
Every step of every worker checks an all-reduce against the cluster size.

Reference: /root/reference/benchmarks/adaptation/gen-compose.py:60-123 and
compose-run.sh (docker containers on a virtual network there),
/root/reference/.github/workflows/cluster-test-1.sh,
/root/reference/tests/go/cmd/kungfu-config-client-example/kungfu-config-client-example.go:96-116.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
import urllib.request

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
BIN = os.path.join(ROOT, 'kungfu_b200', 'bin')
WORKER = os.path.join(ROOT, 'examples', 'cluster_manager.py')


def put(url, c):
    req = urllib.request.Request(url, data=json.dumps(c).encode(), method='PUT')
    with urllib.request.urlopen(req, timeout=5) as r:
        return r.status


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--nodes', type=int, default=3)
    ap.add_argument('--cap', type=int, default=2, help='worker slots per node')
    ap.add_argument('--np', type=int, default=2, help='size of the first cluster')
    ap.add_argument('--ttl', type=float, default=10.0, help='seconds of random changes')
    ap.add_argument('--period', type=float, default=0.5)
    ap.add_argument('--seed', type=int, default=1)
    ap.add_argument('--max-step', type=int, default=0, help='steps per worker (default: enough to outlast --ttl)')
    ap.add_argument('--port-base', type=int, default=27000)
    ap.add_argument('--timeout', type=float, default=180.0)
    args = ap.parse_args()

    nodes = ['127.0.0.%d' % (i + 1) for i in range(args.nodes)]
    hostlist = ','.join('%s:%d' % (h, args.cap) for h in nodes)
    base, runner_port, cfg_port = args.port_base, args.port_base + 80, args.port_base + 90
    port_range = '%d-%d' % (base, base + 16)
    url = 'http://127.0.0.1:%d/config' % cfg_port
    total = args.nodes * args.cap
    max_step = args.max_step or int(args.ttl / 0.1) + 60  # a step sleeps 0.1 s

    def cluster(n):
        slots = [(h, j) for h in nodes for j in range(args.cap)][:n]
        return {'Runners': ['%s:%d' % (h, runner_port) for h in nodes], 'Workers': ['%s:%d' % (h, base + j) for h, j in slots]}

    log = tempfile.NamedTemporaryFile('w+', prefix='kf-virtual-cluster-', suffix='.log')
    procs = [subprocess.Popen([os.path.join(BIN, 'kungfu-config-server'), '-port', str(cfg_port)],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)]
    rc = 1
    try:
        for _ in range(50):
            try:
                put(url, cluster(args.np))
                break
            except Exception:
                time.sleep(0.1)
        env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
        runners = []
        for h in nodes:
            runners.append(subprocess.Popen(
                [os.path.join(BIN, 'kungfu-run'), '-q', '-w', '-k', '-np', str(args.np), '-H', hostlist, '-self', h,
                 '-port', str(runner_port), '-port-range', port_range, '-config-server', url,
                 '-timeout', '%ds' % int(args.timeout), sys.executable, WORKER, '--worker', '--max-step', str(max_step)],
                stdout=log, stderr=subprocess.STDOUT, env=env))
        procs += runners
        time.sleep(3)  # first workers import torch
        client = subprocess.run([os.path.join(BIN, 'kungfu-config-client'), '-server', url, '-H', hostlist, '-port', str(runner_port),
                                 '-port-range', port_range, '-period', '%dms' % int(args.period * 1000), '-ttl',
                                 '%ds' % int(args.ttl), '-shuffle', '-seed', str(args.seed)],
                                stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=args.ttl + 30)
        n_put = client.stdout.count('-> 200')
        print('[virtual-cluster] %d random clusters accepted by the config server' % n_put, flush=True)
        print('[virtual-cluster] final cluster of %d -> HTTP %d' % (total, put(url, cluster(total))), flush=True)
        deadline = time.time() + args.timeout
        done = 0
        while time.time() < deadline:
            text = open(log.name).read()
            done = text.count('finished %d steps with np=%d' % (max_step, total))
            if done >= total or any(r.poll() is not None for r in runners):
                break
            time.sleep(0.5)
        text = open(log.name).read()
        print('[virtual-cluster] %d of %d workers finished %d steps; %d membership changes seen by workers, %d workers left' %
              (done, total, max_step, text.count(': now rank '), text.count(' leaves at step ')), flush=True)
        for h in nodes:  # runners were started with -k: ask them to stop
            subprocess.run([os.path.join(BIN, 'kungfu-test-util'), '-kill', '%s:%d' % (h, runner_port)],
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=20)
        rcs = []
        for r in runners:
            try:
                rcs.append(r.wait(timeout=20))
            except subprocess.TimeoutExpired:
                rcs.append(None)
        print('[virtual-cluster] runners exited with', rcs, flush=True)
        rc = 0 if (done == total and n_put >= 3 and all(c == 0 for c in rcs)) else 1
        if rc:
            sys.stdout.write(text[-6000:])
    finally:
        for p in procs:
            if p.poll() is None:
                p.terminate()
        for p in procs:
            try:
                p.wait(timeout=10)
            except Exception:
                p.kill()
    return rc


if __name__ == '__main__':
    sys.exit(main())
