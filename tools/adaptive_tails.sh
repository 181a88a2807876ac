#!/bin/bash
# Dev tool (under gpurun, ONE GPU): closing-phase lengths of the render kernel under the adaptive schedule, shards of an 8-GPU run.
#   TM="12:12 6:12" RANKS="0 5" bash tools/adaptive_tails.sh        (HF_B200_RTAIL:HF_B200_RMIDTAIL, quarter tiles per resident warp)
cd "$(dirname "$0")/.."
show='
import sys, json
for line in sys.stdin:
    if line.startswith("{"):
        d = json.loads(line); rs = d["extras"]["render_step"]; print(sys.argv[1], "render step %.4f kernel %.4f %s" % (rs["ms_per_step"], rs["roofline"]["kernel_ms"], rs.get("adaptive_schedule")))
'
for tm in ${TM:-12:12 6:12 6:6 2:6}; do
  for r in ${RANKS:-0 5}; do
    HF_B200_RTAIL=${tm%%:*} HF_B200_RMIDTAIL=${tm##*:} HF_BENCH_SHARD_OF=8 HF_BENCH_SHARD_RANK=$r python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "$show" "rank $r rtail:rmid=$tm"
  done
done
