#!/bin/bash
# Dev tool (under gpurun): C2 / C4 render timings for every lib/libhf_b200*.so
cd "$(dirname "$0")/.."
for so in mitsuba3-differentiable-heightfield-rendering_b200/lib/libhf_b200*.so; do
  echo "== $(basename $so)"
  HF_B200_LIB=$PWD/$so python tools/run_configs.py --cases C4,C2 2>&1 | grep render_forward | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('   %-80s %.3f ms' % (d['case'][:80], d['render_forward_ms']))"
done
