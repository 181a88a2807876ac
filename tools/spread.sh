#!/bin/bash
# Dev tool (under gpurun, ONE GPU): the kernel times every rank of an N-GPU run would see (each rank's shard timed alone).
#   bash tools/spread.sh [N] [tile]
cd "$(dirname "$0")/.."
N=${1:-8}; tile=${2:-32}
for r in $(seq 0 $((N - 1))); do
  HF_BENCH_SHARD_OF=$N HF_BENCH_SHARD_RANK=$r python bench.py --tile $tile --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); print('tile $tile rank $r of $N: fwd %.4f bwd %.4f step %.4f isect %.4f render_kernel %.4f' % (d['roofline']['kernel_ms'], d['roofline']['backward']['kernel_ms'], d['ms_per_step'], d['extras']['intersect_only_ms'], d['extras']['render_step']['roofline']['kernel_ms']))
"; done
