#!/bin/bash
# Dev tool (under gpurun, ONE GPU): adaptive schedule off / on -- the full C3 batch and the shards some ranks of an 8-GPU run would get.
#   SETTINGS="8:16 4:32" RANKS="0 3 5" bash tools/adaptive_ab.sh      (factor:span of hf_pick_heavy_kernel)
cd "$(dirname "$0")/.."
show='
import sys, json
for line in sys.stdin:
    if line.startswith("{"):
        d = json.loads(line); print(sys.argv[1], "fwd %.4f bwd %.4f step %.4f isect %.4f sched %s" % (d["roofline"]["kernel_ms"], d["roofline"]["backward"]["kernel_ms"], d["ms_per_step"], d["extras"]["intersect_only_ms"], d["extras"].get("adaptive_schedule")), "| render step %.4f kernel %.4f %s" % (d["extras"]["render_step"]["ms_per_step"], d["extras"]["render_step"]["roofline"]["kernel_ms"], d["extras"]["render_step"].get("adaptive_schedule")))
'
run() {   # label, env...
  label=$1; shift
  env "$@" python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "$show" "full $label"
  for r in ${RANKS:-0 3 5}; do
    env "$@" HF_BENCH_SHARD_OF=8 HF_BENCH_SHARD_RANK=$r python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "$show" "rank $r of 8 $label"
  done
}
[ -z "$NOFIXED" ] && run "fixed" HF_BENCH_ADAPTIVE=0
for fs in ${SETTINGS:-8:16}; do run "adaptive factor:span=$fs" HF_BENCH_ADAPTIVE=1 HF_B200_HEAVY_FACTOR=${fs%%:*} HF_B200_HEAVY_SPAN=${fs##*:}; done
