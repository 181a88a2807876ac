#!/bin/bash
# Dev tool (run under gpurun): round-2 measurement pass.  gpurun --timeout 2400 -- 'bash tools/gpu_round2.sh r02c'
# GPU parity tests, the C3 bench line, C4 and C5 as BASELINE.json states them, the closing-phase sweep at 1/8 batch, ncu captures.
tag=${1:-r02c}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${tag}_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; cat gpurun_out/${tag}_bench.json
timeout 600 python bench.py --config C4 --heights 2048 --image 512 --steps 3 --no-cpu > gpurun_out/${tag}_c4_small.json 2> gpurun_out/${tag}_c4_small.err; echo "c4 small rc=$?"; tail -3 gpurun_out/${tag}_c4_small.err
timeout 1200 python bench.py --config C4 --steps 5 > gpurun_out/${tag}_c4.json 2> gpurun_out/${tag}_c4.err; echo "c4 rc=$?"; cat gpurun_out/${tag}_c4.json; tail -3 gpurun_out/${tag}_c4.err
timeout 1500 python bench.py --config C5 > gpurun_out/${tag}_c5.json 2> gpurun_out/${tag}_c5.err; echo "c5 rc=$?"; cut -c1-1500 gpurun_out/${tag}_c5.json; tail -3 gpurun_out/${tag}_c5.err
for tm in; do set -- $tm
  HF_BENCH_SHARD_OF=8 HF_B200_TAIL=$1 HF_B200_MIDTAIL=$2 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); print('shard 1/8 tail $1 mid $2: fwdSI_ms %.4f bwd_ms %.4f step_ms %.4f render_kernel_ms %.4f' % (d['roofline']['kernel_ms'], d['roofline']['backward']['kernel_ms'], d['ms_per_step'], d['extras']['render_step']['roofline']['kernel_ms']))
"; done 2>&1 | tee gpurun_out/${tag}_tails.txt
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:hf_ -c 80 --csv --log-file gpurun_out/${tag}_launches.csv $BCMD > gpurun_out/${tag}_ncu_l.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'hf_intersect_fast_kernel|hf_backward_kernel|hf_render_kernel' -s 6 -c 4 -o gpurun_out/${tag}_prof -f $BCMD > gpurun_out/${tag}_ncu_f.log 2>&1; echo "ncu full rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_prof.ncu-rep > gpurun_out/${tag}_ncu_full_summary.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:'hf_render_path_kernel' -s 2 -c 2 -o gpurun_out/${tag}_c4prof -f python bench.py --config C4 --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_c4prof.ncu-rep > gpurun_out/${tag}_ncu_c4_summary.txt 2>&1
ls -la gpurun_out | tail -30
