"""In-tree native build: host runtime (g++) and device library (nvcc, sm_100a).

Everything is built into ``kungfu_b200/lib`` and ``kungfu_b200/bin`` so the
artefacts travel with the source tree (no JIT cache, no site-packages install).

    python -m kungfu_b200._build [--force] [--host-only | --cuda-only]

The reference drove CMake from setup.py (/root/reference/setup_tensorflow.py,
/root/reference/CMakeLists.txt); this tree needs neither CMake nor Go.
"""
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, 'kungfu_b200')
LIB_DIR = os.path.join(PKG, 'lib')
BIN_DIR = os.path.join(PKG, 'bin')
OBJ_DIR = os.path.join(ROOT, 'build', 'obj')

NVCC_FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3',
    '-std=c++17', '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr',
]
CXX_FLAGS = ['-O2', '-g', '-std=c++17', '-fPIC', '-Wall', '-pthread', '-mavx2', '-mf16c']

CUDA_SOURCES = [
    'csrc/kernels/collectives.cu',
    'csrc/kernels/optim.cu',
    'csrc/kernels/bn_act.cu',
    'csrc/kernels/pool.cu',
    'csrc/kernels/linear_tc.cu',
    'csrc/kernels/ln_fused.cu',
    'csrc/device/symm.cu',
]
HOST_LIB_DIRS = ['csrc/core', 'csrc/ctrl', 'csrc/elastic', 'csrc/capi']
HOST_BINARIES = {
    'kungfu-run': 'csrc/launch/kungfu_run_main.cpp',
    'kungfu-config-server': 'csrc/launch/config_server_main.cpp',
    'kungfu-rrun': 'csrc/launch/remote_tools.cpp',
    'kungfu-distribute': 'csrc/launch/remote_tools.cpp',
    'kungfu-bench-allreduce': 'csrc/tools/bench_tools.cpp',
    'kungfu-bench-p2p': 'csrc/tools/bench_tools.cpp',
    'kungfu-test-util': 'csrc/tools/ops_tools.cpp',
    'kungfu-notify-start': 'csrc/tools/ops_tools.cpp',
    'kungfu-config-client': 'csrc/tools/ops_tools.cpp',
}


def _nvcc():
    for c in (shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if c and os.path.exists(c):
            return c
    raise RuntimeError('nvcc not found')


def _digest(paths, flags):
    h = hashlib.sha1(' '.join(flags).encode())
    for p in sorted(paths):
        with open(p, 'rb') as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _headers():
    out = []
    for d in ('csrc', 'include'):
        for base, _, files in os.walk(os.path.join(ROOT, d)):
            out += [os.path.join(base, f) for f in files if f.endswith(('.h', '.hpp', '.cuh'))]
    return out


def _run(cmd):
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError('build failed: %s\n%s' % (' '.join(cmd), r.stdout))
    return r.stdout


def _compile(compiler, flags, src, force, hdr_digest):
    os.makedirs(OBJ_DIR, exist_ok=True)
    obj = os.path.join(OBJ_DIR, src.replace('/', '_') + '.o')
    stamp = obj + '.sha1'
    want = _digest([os.path.join(ROOT, src)], flags) + hdr_digest
    if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == want:
        return obj, False
    _run([compiler] + flags + ['-I', os.path.join(ROOT, 'include'), '-c', src, '-o', obj])
    with open(stamp, 'w') as f:
        f.write(want)
    return obj, True


def _compile_all(compiler, flags, sources, force):
    hdr = _digest(_headers(), [])
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        res = list(ex.map(lambda s: _compile(compiler, flags, s, force, hdr), sources))
    return [o for o, _ in res], any(c for _, c in res)


def build_cuda(force=False, verbose=True):
    """Compile every .cu for sm_100a and link kungfu_b200/lib/libkungfu_cuda.so."""
    sources = [s for s in CUDA_SOURCES if os.path.exists(os.path.join(ROOT, s))]
    os.makedirs(LIB_DIR, exist_ok=True)
    out = os.path.join(LIB_DIR, 'libkungfu_cuda.so')
    objs, changed = _compile_all(_nvcc(), NVCC_FLAGS, sources, force)
    if changed or not os.path.exists(out):
        _run([_nvcc(), '-shared', '-o', out] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-lcudart_static', '-ldl', '-lrt', '-lpthread'])
        if verbose:
            print('[kungfu_b200] linked', out)
    return out


def _host_sources():
    out = []
    for d in HOST_LIB_DIRS:
        p = os.path.join(ROOT, d)
        if os.path.isdir(p):
            out += sorted(os.path.join(d, f) for f in os.listdir(p) if f.endswith('.cpp'))
    return out


def build_host(force=False, verbose=True):
    """Build libkungfu_host.so (control plane, planner, elastic runtime) and the CLI tools."""
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(BIN_DIR, exist_ok=True)
    cxx = shutil.which('g++') or 'g++'
    sources = _host_sources()
    if not sources:
        return None
    out = os.path.join(LIB_DIR, 'libkungfu_host.so')
    objs, changed = _compile_all(cxx, CXX_FLAGS + ['-I', os.path.join(ROOT, 'csrc')], sources, force)
    if changed or not os.path.exists(out):
        _run([cxx, '-shared', '-o', out] + objs + ['-pthread', '-ldl'])
        if verbose:
            print('[kungfu_b200] linked', out)
    for name, src in HOST_BINARIES.items():
        if not os.path.exists(os.path.join(ROOT, src)):
            continue
        exe = os.path.join(BIN_DIR, name)
        (obj,), ch = _compile_all(cxx, CXX_FLAGS + ['-I', os.path.join(ROOT, 'csrc')], [src], force)
        # several tools share one source: only the first of them sees it recompiled, so go by the object's age too
        if ch or changed or not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(obj):
            _run([cxx, '-o', exe, obj] + objs + ['-pthread', '-ldl'])
            if verbose:
                print('[kungfu_b200] linked', exe)
    return out


def build_all(force=False, verbose=True):
    build_host(force, verbose)
    build_cuda(force, verbose)


if __name__ == '__main__':
    force = '--force' in sys.argv
    if '--host-only' in sys.argv:
        build_host(force)
    elif '--cuda-only' in sys.argv:
        build_cuda(force)
    else:
        build_all(force)
