#!/usr/bin/env python
"""Random command lines against kungfu-run / kungfu-config-server.

A launcher must answer garbage with a usage error, never with a crash: every
invocation has to end by itself (exit code >= 0, no signal) within the time
limit. Value flags swallow the next token, so a line like `-logfile -q` creates
a file called "-q": the whole run therefore happens inside a scratch directory
that is removed afterwards, never in the source tree.

    python scripts/fuzz_cli.py [-n 200] [--seed 1]
"""
import argparse
import os
import random
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'kungfu_b200', 'bin')
FLAGS = ['-np', '-H', '-hostfile', '-P', '-u', '-port-range', '-self', '-timeout', '-v', '-nic', '-allow-nvlink',
         '-strategy', '-port', '-debug-port', '-w', '-k', '-init-version', '-config-server', '-t0', '-logfile',
         '-logdir', '-q', '-delay', '-builtin-config-port', '-init', '-ttl', '-endpoint', '--', '-', '']
VALUES = ['0', '1', '3', '-1', '99999999999', 'abc', '127.0.0.1:2', '127.0.0.1:2:x', '25000-25010', '1-0', 'lo', '1s',
          '0.01s', 'RING', 'NOPE', 'http://127.0.0.1:1/x', '\xff', 'a' * 5000, '/dev/null', '/nonexistent/x', '%s%n']


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('-n', type=int, default=200)
    ap.add_argument('--seed', type=int, default=1)
    args = ap.parse_args()
    rng = random.Random(args.seed)
    scratch = tempfile.mkdtemp(prefix='kungfu-fuzz-cli-')
    bad = []
    try:
        for i in range(args.n):
            exe = rng.choice(['kungfu-run', 'kungfu-run', 'kungfu-config-server'])
            argv = [os.path.join(BIN, exe)]
            for _ in range(rng.randrange(0, 8)):
                argv.append(rng.choice(FLAGS if rng.random() < 0.6 else VALUES))
            if exe == 'kungfu-run' and rng.random() < 0.5:
                argv += ['-timeout', '1s', 'true']
            try:
                r = subprocess.run(argv, cwd=scratch, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=20,
                                   env=dict(os.environ, KUNGFU_CONFIG_LOG_LEVEL='ERROR'))
                if r.returncode < 0:
                    bad.append((r.returncode, argv))
            except subprocess.TimeoutExpired:
                # a server that was given usable flags keeps serving: that is not a fault
                if exe != 'kungfu-config-server':
                    bad.append(('timeout', argv))
    finally:
        shutil.rmtree(scratch, ignore_errors=True)
    for rc, argv in bad:
        print('FAILED', rc, ' '.join(repr(a[:60]) for a in argv[1:]))
    print('%d command lines, %d failures' % (args.n, len(bad)))
    return 1 if bad else 0


if __name__ == '__main__':
    sys.exit(main())
