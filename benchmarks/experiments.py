#!/usr/bin/env python
"""Experiment matrix runner: sweep model x optimizer x batch size x cluster size, one
`kungfu-run ... train_bench.py` job per point, results appended to a JSONL file.

    python benchmarks/experiments.py --models resnet50,vgg16 --optimizers ssgd,sma,pair \
        --batch-sizes 64,256 --np 1,2,4,8 --out results.jsonl [--hosts h1:8,h2:8 --user me] [--dry-run]

With --hosts the jobs are started on every host through kungfu-rrun (ssh), otherwise on this
machine. Two more modes, as in the reference's experiment tools:

    --remote-install   pack the tracked tree, push it to every host of --hosts over ssh and `pip install` it
                       there (kungfu-remote-install; it cloned from GitHub, there is no network here)
    --scaling SCHEDULE start a config server, publish the first cluster and run the elastic all-reduce trainer
                       along `steps:size,steps:size,...` (kungfu-run-scaling-experiments), across --hosts or here

Reference: /root/reference/experiments/ (Go: remote install + train-experiment sweep over
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


def remote_install(args):
    """One `ssh host 'rm -rf DIR && mkdir -p DIR && tar -xz -C DIR && cd DIR/kungfu-b200 && pip install'` per host, fed
    with the source tarball on stdin; hosts in parallel."""
    import tempfile
    import threading
    hosts = [h.split(':') for h in args.hosts.split(',') if h]
    targets = [(h[2] if len(h) > 2 else h[0]) for h in hosts]  # ip:slots[:public_addr]
    remote = ('rm -rf {d} && mkdir -p {d} && tar -xz -C {d} && cd {d}/kungfu-b200 && '
              'KUNGFU_BUILD_CUDA={cuda} {py} -m pip install {pip} --no-index --no-build-isolation --no-deps -U .').format(
                  d=args.install_dir, py=args.python, cuda=int(args.build_cuda), pip=args.pip_args)
    cmds = [['ssh', '-o', 'BatchMode=yes', '-o', 'StrictHostKeyChecking=no', (args.user + '@' if args.user else '') + t, remote]
            for t in targets]
    if args.dry_run:
        for c in cmds:
            print('scripts/pack.sh | ' + ' '.join(c[:-1]) + " '" + c[-1] + "'")
        return 0
    with tempfile.TemporaryDirectory() as tmp:
        tarball = subprocess.check_output([os.path.join(ROOT, 'scripts', 'pack.sh'), os.path.join(tmp, 'src.tar.gz')], text=True).strip()
        rcs = [None] * len(cmds)

        def one(i):
            with open(tarball, 'rb') as f:
                r = subprocess.run(cmds[i], stdin=f, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=args.timeout)
            rcs[i] = r.returncode
            print('[%s] install %s\n%s' % (targets[i], 'ok' if r.returncode == 0 else 'FAILED (%d)' % r.returncode,
                                           '' if r.returncode == 0 else r.stdout[-2000:]), flush=True)

        ths = [threading.Thread(target=one, args=(i, )) for i in range(len(cmds))]
        [t.start() for t in ths]
        [t.join() for t in ths]
    return 0 if all(rc == 0 for rc in rcs) else 1


def scaling(args):
    """Elastic all-reduce trainer along a resize schedule under a config server this script owns."""
    import urllib.request
    first = int(args.scaling.split(',')[0].split(':')[1])
    trainer = [sys.executable, os.path.join(ROOT, 'benchmarks', 'adaptation', 'adaptive_trainer.py'), '--schedule', args.scaling,
               '--model-mb', str(args.model_mb)]
    if not args.hosts:
        cmd = [os.path.join(BIN, 'kungfu-run'), '-q', '-w', '-builtin-config-port', str(args.config_port), '-np', str(first), '-H',
               '127.0.0.1:%d' % args.slots, '-port-range', args.port_range, '-port', str(args.runner_port)] + trainer
        if args.dry_run:
            print(' '.join(cmd))
            return 0
        return subprocess.run(cmd, timeout=args.timeout).returncode
    cfg_host = args.hosts.split(',')[0].split(':')[0]
    url = 'http://%s:%d/config' % (cfg_host, args.config_port)
    server = [os.path.join(BIN, 'kungfu-config-server'), '-port', str(args.config_port)]
    init = [os.path.join(BIN, 'kungfu-config-client'), '-server', url, '-H', args.hosts, '-port', str(args.runner_port), '-port-range',
            args.port_range, '-example']
    job = [os.path.join(BIN, 'kungfu-rrun'), '-H', args.hosts] + (['-u', args.user] if args.user else []) + \
          ['-q', '-w', '-np', str(first), '-config-server', url, '-port-range', args.port_range, '-port', str(args.runner_port)] + trainer
    if args.dry_run:
        for c in (server, init, job):
            print(' '.join(c))
        print('curl %s/stop' % url.rsplit('/', 1)[0])
        return 0
    srv = subprocess.Popen(server, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    try:
        # the first cluster: the first `first` slots of the host list (what every runner derives from -np / -H as well)
        full = json.loads(subprocess.check_output(init, text=True))
        full['Workers'] = full['Workers'][:first]
        for _ in range(50):
            if srv.poll() is not None:
                raise RuntimeError('kungfu-config-server exited with %d (is port %d taken?)' % (srv.returncode, args.config_port))
            try:
                req = urllib.request.Request(url, data=json.dumps(full).encode(), method='PUT')
                urllib.request.urlopen(req, timeout=5).read()
                break
            except Exception:
                time.sleep(0.1)
        else:
            raise RuntimeError('the config server at %s never accepted the first cluster' % url)
        return subprocess.run(job, timeout=args.timeout).returncode
    finally:
        try:
            urllib.request.urlopen(url.rsplit('/', 1)[0] + '/stop', timeout=5).read()
        except Exception:
            pass
        try:
            srv.wait(timeout=10)
        except Exception:
            srv.kill()


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
    ap.add_argument('--remote-install', action='store_true')
    ap.add_argument('--install-dir', default='.kungfu')
    ap.add_argument('--python', default='python3')
    ap.add_argument('--pip-args', default='--user', help='where pip installs on the hosts (default --user)')
    ap.add_argument('--build-cuda', action='store_true', help='also compile the sm_100a kernels on the hosts (needs nvcc there)')
    ap.add_argument('--scaling', default='', metavar='SCHEDULE', help='steps:size,steps:size,... e.g. 3:2,3:4,3:8,3:1')
    ap.add_argument('--model-mb', type=float, default=97.6)
    ap.add_argument('--slots', type=int, default=16, help='local slots for --scaling without --hosts')
    ap.add_argument('--config-port', type=int, default=9100)
    ap.add_argument('extra', nargs='*', help='extra train_bench.py flags after --')
    args = ap.parse_args()
    if args.remote_install:
        if not args.hosts:
            ap.error('--remote-install needs --hosts')
        return remote_install(args)
    if args.scaling:
        return scaling(args)
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
