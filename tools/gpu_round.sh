#!/bin/bash
# Dev tool (run under gpurun): GPU parity tests, the bench line, the ncu launch list and one full capture per hot kernel.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh r01'
# Output lands in gpurun_out/<tag>_*; copy the summaries worth keeping to profiles/.
tag=${1:-r01}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt 2>&1
nproc > gpurun_out/${tag}_nproc.txt; lscpu | head -20 >> gpurun_out/${tag}_nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$BCMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:hf_ -c 60 --csv --log-file gpurun_out/${tag}_launches.csv $BCMD > gpurun_out/${tag}_ncu_l.log 2>&1
echo "ncu launches rc=$?"
$BCMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'hf_intersect_fast_kernel|hf_backward_kernel|hf_render_kernel' -s 6 -c 4 -o gpurun_out/${tag}_prof -f $BCMD > gpurun_out/${tag}_ncu_f.log 2>&1
echo "ncu full rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_prof.ncu-rep > gpurun_out/${tag}_ncu_full_summary.txt 2>&1
ls -la gpurun_out | tail -20
