cd /root/repo
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r02d_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02d_pytest.log
for tm in "24 24" "48 48" "96 96" "12 24" "6 12" "0 0"; do set -- $tm
  HF_BENCH_SHARD_OF=8 HF_B200_RTAIL=$1 HF_B200_RMIDTAIL=$2 HF_B200_TAIL=$(( $1 * 2 )) HF_B200_MIDTAIL=$(( $2 * 2 )) python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); print('shard 1/8 rtail $1 rmid $2 (fast tail x2): fwdSI_ms %.4f bwd_ms %.4f step_ms %.4f render_kernel_ms %.4f render_step_ms %.4f' % (d['roofline']['kernel_ms'], d['roofline']['backward']['kernel_ms'], d['ms_per_step'], d['extras']['render_step']['roofline']['kernel_ms'], d['extras']['render_step']['ms_per_step']))
"; done 2>&1 | tee gpurun_out/r02d_tails.txt
