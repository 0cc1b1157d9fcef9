"""The repository root holds source, docs and driver records only: a command-line
fuzzing session once left files named after flag values ("-q", "abc", "\\xff") there."""
import os
import re
import subprocess
import sys

from conftest import ROOT

ALLOWED = re.compile(r'^(\.github|\.gitignore|\.gpurunignore|[A-Z][A-Z_]*(_r\d+)?\.(md|json|jsonl)|Makefile|README\.md|'
                     r'__graft_entry__\.py|bench\.py|setup\.py|benchmarks|csrc|docker|docs|examples|include|kungfu|'
                     r'kungfu_b200|profiles|scripts|tests|baseline)$')


def test_tracked_root_entries_are_the_expected_ones():
    out = subprocess.run(['git', 'ls-files'], cwd=ROOT, stdout=subprocess.PIPE, text=True)
    if out.returncode != 0:
        import pytest
        pytest.skip('not a git checkout')
    tops = sorted({line.split('/')[0] for line in out.stdout.splitlines() if line})
    strange = [t for t in tops if not ALLOWED.match(t)]
    assert not strange, 'unexpected entries at the repository root: %r' % strange


def test_cli_fuzzing_leaves_no_files_behind_and_finds_no_crash():
    before = set(os.listdir(ROOT))
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'scripts', 'fuzz_cli.py'), '-n', '25', '--seed', '7'], cwd=ROOT,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    assert set(os.listdir(ROOT)) == before


def test_no_python_source_reads_a_name_it_never_binds():
    """scripts/check_names.py over every Python file: the device-side modules and the GPU workers only execute on a
    GPU box, so an un-imported or misspelt name in them must be caught here."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'scripts', 'check_names.py')], cwd=ROOT,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:]
