#!/bin/bash
# Dev tool (under gpurun): e2e (host buffers, PCIe inside the timed region) of every lib/libhf_b200*.so on the same box.
cd "$(dirname "$0")/.."
for rep in 1 2; do
for so in mitsuba3-differentiable-heightfield-rendering_b200/lib/libhf_b200*.so; do
  HF_B200_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-cpu "$@" 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); e = d['e2e']
        print('$(basename $so)', 'e2e %.0f Mrays/s %.1f ms' % (e['value'], e['ms_per_step']), 'value %.0f' % d['value'])
"
done
done
