#!/bin/bash
# Dev tool (under gpurun): GPU parity tests on the current build, then kernel timings of every lib/libhf_b200*.so (A/B).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
bash tools/kbench.sh "$@"
