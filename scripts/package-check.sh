#!/bin/sh
# Packaging check: a clean source archive must `pip install` offline into a scratch directory (host runtime
# and tools are compiled by setup.py; KUNGFU_BUILD_CUDA=0 skips the kernels so it also runs on boxes without
# nvcc) and a 2-worker job must run from the INSTALLED copy through the installed kungfu-run console script.
# Reference packaging: `pip install .` via setup_tensorflow.py / setup_pytorch.py.
set -e
cd "$(dirname "$0")/.."
tmp=$(mktemp -d)
trap 'rm -rf "$tmp"' EXIT
mkdir "$tmp/src" "$tmp/inst"
git archive HEAD | tar -x -C "$tmp/src"
(cd "$tmp/src" && KUNGFU_BUILD_CUDA=${KUNGFU_BUILD_CUDA:-0} python -m pip install -q --no-index --no-build-isolation --no-deps --target "$tmp/inst" .)
base=$((31500 + $$ % 500))
cd "$tmp"
out=$(PYTHONPATH="$tmp/inst" "$tmp/inst/bin/kungfu-run" -q -np 2 -timeout 100s -port-range "$base-$((base + 8))" -port "$((base + 20))" \
    python -c "import kungfu.torch as kf, torch; assert kf.__file__.startswith('$tmp/inst'); print('PKG', kf.current_rank(), float(kf.ops.all_reduce_fn(torch.ones(3))[0]))" 2>&1)
echo "$out" | grep -c "PKG [01] 2.0" | grep -qx 2 || { echo "$out"; echo "package check FAILED"; exit 1; }
echo "package check ok: installed copy ran a 2-worker all-reduce"
