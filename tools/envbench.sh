#!/bin/bash
# Dev tool (under gpurun): kernel timings of the current build under several environment settings: tools/envbench.sh "A=1" "A=2 B=3" ...
cd "$(dirname "$0")/.."
for e in "$@"; do
  env $e python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); r = d['extras']['render_step']
        print('$e', 'fwdSI_ms %.3f' % d['roofline']['kernel_ms'], 'isect_ms %.3f' % d['extras']['intersect_only_ms'], 'bwd_ms %.3f' % d['roofline']['backward']['kernel_ms'], 'value %.0f' % d['value'], 'frac %.3f' % d['roofline']['frac'], '| render kernel_ms %.3f step_ms %.3f' % (r['roofline']['kernel_ms'], r['ms_per_step']))
"
done
