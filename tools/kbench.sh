#!/bin/bash
# Dev tool: kernel timings for every experiment build lib/libhf_b200*.so (run under gpurun).
cd "$(dirname "$0")/.."
for so in mitsuba3-differentiable-heightfield-rendering_b200/lib/libhf_b200*.so; do
  HF_B200_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); r = d['extras']['render_step']
        print('$(basename $so)', 'fwdSI_ms %.3f' % d['roofline']['kernel_ms'], 'isect_ms %.3f' % d['extras']['intersect_only_ms'], 'bwd_ms %.3f' % d['roofline']['backward']['kernel_ms'], 'value %.0f' % d['value'], 'frac %.3f' % d['roofline']['frac'], '| render kernel_ms %.3f step_ms %.3f frac %.3f' % (r['roofline']['kernel_ms'], r['ms_per_step'], r['roofline']['frac']), 'sm_mhz', d['clocks']['sm_mhz'])
"
done
