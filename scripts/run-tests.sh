#!/bin/bash
# Build everything and run the CPU test suite; with a GPU also the GPU suite.
# (reference: /root/reference/scripts/tests/run-{integration,op,train,elastic}-tests.sh)
set -e
cd "$(dirname "$0")/.."
python -m kungfu_b200._build
python -m pytest tests -x -q -m "not gpu"
if python -c "import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)"; then
    python -m pytest tests -x -q -m gpu
fi
