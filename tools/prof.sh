#!/bin/bash
# Dev tool (under gpurun): full ncu capture (with source) of one kernel regex $1 at launch index $2 -> gpurun_out/$3.*; env passes through.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:"$1" -s $2 -c 1 -o gpurun_out/$3 -f $BCMD > gpurun_out/$3_ncu.log 2>&1
echo "ncu rc=$?"
python tools/ncu_summary.py gpurun_out/$3.ncu-rep 45 > gpurun_out/$3_summary.txt 2>&1
python tools/ncu_groups.py gpurun_out/$3.ncu-rep >> gpurun_out/$3_summary.txt 2>&1
