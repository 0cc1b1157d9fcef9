#!/bin/sh
# Format the tree (reference: /root/reference/scripts/clean-code.sh). Tools are optional.
set -e
cd "$(dirname "$0")/.."
if command -v clang-format >/dev/null 2>&1; then
    find csrc include tests/cpp -name '*.cpp' -o -name '*.hpp' -o -name '*.h' -o -name '*.cu' -o -name '*.cuh' |
        xargs clang-format -i
fi
if python -c 'import yapf' >/dev/null 2>&1; then
    python -m yapf -i -r kungfu_b200 kungfu tests examples benchmarks bench.py __graft_entry__.py
fi
python -m compileall -q kungfu_b200 kungfu tests examples benchmarks >/dev/null
echo "clean-code: done"
