#!/bin/bash
# Dev tool (under gpurun): one full ncu capture of kernels matching $1 (regex), launch index $2, output tag $3.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$BCMD > gpurun_out/$3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$1" -s $2 -c 1 -o gpurun_out/$3 -f $BCMD > gpurun_out/$3_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/$3_ncu.log
