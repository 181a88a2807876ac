#!/bin/bash
# Dev tool (under gpurun): GPU parity tests, then kernel timings with the bundled kernel on and off.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for b in 1 0; do
  HF_B200_BUNDLES=$b python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line)
        print('bundles=$b', 'fwdSI_ms %.3f' % d['roofline']['kernel_ms'], 'isect_ms %.3f' % d['extras']['intersect_only_ms'], 'bwd_ms %.3f' % d['roofline']['backward']['kernel_ms'], 'step_ms %.3f' % d['ms_per_step'], 'value %.0f' % d['value'], 'frac %.3f' % d['roofline']['frac'], 'sm_mhz', d['clocks']['sm_mhz'], 'gradsum', d['extras']['grad_abs_sum'])
    else: print(line, end='')
"
done
