#!/bin/bash
# A/B of the overlapped (bucketed) fused step on N GPUs of one box: same box, same session.
N=${1:-2}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps ${STEPS:-40} --warmup 5 --skip-e2e --same-box-baseline none "$@" 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%-44s %.3f ms  %.0f img/s  buckets=%s algo=%s' % (sys.argv[1], d['ms_per_step'], d['value'], d['config'].get('buckets'), d['config'].get('fused_step')))" "$*"; }
[ "$N" = 1 ] && run() { python bench.py --gpus 1 --steps ${STEPS:-40} --warmup 5 --skip-e2e --same-box-baseline none "$@" 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%-44s %.3f ms  %.0f img/s  buckets=%s algo=%s' % (sys.argv[1], d['ms_per_step'], d['value'], d['config'].get('buckets'), d['config'].get('fused_step')))" "$*"; }
run --overlap 0
KUNGFU_OVERLAP_BLOCKS=8 run --overlap 1
KUNGFU_OVERLAP_BLOCKS=32 run --overlap 1
KUNGFU_OVERLAP_BLOCKS=4 KUNGFU_BUCKET_BYTES=$((6<<20)) run --overlap 1
run --overlap 0
