#!/usr/bin/env python
"""Experiment matrix runner: sweep model x optimizer x batch size x cluster size, one
`kungfu-run ... train_bench.py` job per point, results appended to a JSONL file.

    python benchmarks/experiments.py --models resnet50,vgg16 --optimizers ssgd,sma,pair \
        --batch-sizes 64,256 --np 1,2,4,8 --out results.jsonl [--hosts h1:8,h2:8 --user me] [--dry-run]

With --hosts the jobs are started on every host through kungfu-rrun (ssh), otherwise on this
machine. Reference: /root/reference/experiments/ (Go: remote install + train-experiment sweep over
model x optimizer x batch x cluster sizes + scaling experiment with a config server).
"""
import argparse
import itertools
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'kungfu_b200', 'bin')


def commands(args):
    for model, opt, bs, np_ in itertools.product(args.models, args.optimizers, args.batch_sizes, args.np):
        bench = [sys.executable, os.path.join(ROOT, 'benchmarks', 'train_bench.py'), '--model', model, '--optimizer', opt,
                 '--batch-size', str(bs), '--steps', str(args.steps), '--warmup', str(args.warmup)] + args.extra
        if opt != 'pair' and args.graph:
            bench.append('--graph')
        if args.hosts:
            cmd = [os.path.join(BIN, 'kungfu-rrun'), '-H', args.hosts] + (['-u', args.user] if args.user else []) + \
                  ['-np', str(np_), '-allow-nvlink', '-port-range', args.port_range] + bench
        else:
            cmd = [os.path.join(BIN, 'kungfu-run'), '-q', '-np', str(np_), '-allow-nvlink', '-port-range',
                   args.port_range, '-port', str(args.runner_port)] + bench
        yield dict(model=model, optimizer=opt, batch_size=bs, np=np_), cmd


def main():
    ap = argparse.ArgumentParser()
    lst = lambda cast: (lambda s: [cast(x) for x in s.split(',') if x])  # noqa: E731
    ap.add_argument('--models', type=lst(str), default=['resnet50'])
    ap.add_argument('--optimizers', type=lst(str), default=['ssgd'])
    ap.add_argument('--batch-sizes', type=lst(int), default=[256])
    ap.add_argument('--np', type=lst(int), default=[1])
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--graph', action='store_true')
    ap.add_argument('--hosts', default='')
    ap.add_argument('--user', default='')
    ap.add_argument('--port-range', default='31000-31100')
    ap.add_argument('--runner-port', type=int, default=31999)
    ap.add_argument('--timeout', type=float, default=900)
    ap.add_argument('--out', default='experiments.jsonl')
    ap.add_argument('--dry-run', action='store_true')
    ap.add_argument('extra', nargs='*', help='extra train_bench.py flags after --')
    args = ap.parse_args()
    failed = 0
    for point, cmd in commands(args):
        if args.dry_run:
            print(' '.join(cmd))
            continue
        t0 = time.time()
        try:
            res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=args.timeout)
            out, rc = res.stdout, res.returncode
        except subprocess.TimeoutExpired as e:
            out, rc = (e.stdout or b'').decode(errors='replace') if isinstance(e.stdout, bytes) else (e.stdout or ''), 124
        rec = dict(point, rc=rc, wall_s=round(time.time() - t0, 1))
        for line in out.splitlines():
            if 'RESULT:' in line:
                try:
                    rec.update(json.loads(line[line.index('{'):]))
                except ValueError:
                    pass
        if rc != 0:
            failed += 1
            rec['tail'] = out[-500:]
        with open(args.out, 'a') as f:
            f.write(json.dumps(rec) + '\n')
        print(json.dumps(rec)[:300], flush=True)
    return 1 if failed else 0


if __name__ == '__main__':
    sys.exit(main())
