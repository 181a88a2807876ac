#!/bin/bash
# Dev tool (under gpurun): full ncu captures (with source) of the bundled intersection kernel and of the fused render step kernel.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
HF_B200_BUNDLES=1 $BCMD > gpurun_out/p2_plain.log 2>&1 &&
HF_B200_BUNDLES=1 ncu --set full --clock-control none --import-source on -k regex:"hf_intersect_beam_kernel" -s 4 -c 1 -o gpurun_out/p2_beam -f $BCMD > gpurun_out/p2_beam_ncu.log 2>&1
echo "beam rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"hf_render_kernel" -s 3 -c 1 -o gpurun_out/p2_render -f $BCMD > gpurun_out/p2_render_ncu.log 2>&1
echo "render rc=$?"
for k in beam render; do python tools/ncu_summary.py gpurun_out/p2_$k.ncu-rep 60 > gpurun_out/p2_${k}_summary.txt 2>&1; done
ls -la gpurun_out | tail
