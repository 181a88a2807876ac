#!/bin/bash
# Dev tool (under gpurun, ONE GPU): adaptive schedule off / on for the shards of a 4- and a 2-GPU run.
cd "$(dirname "$0")/.."
show='
import sys, json
for line in sys.stdin:
    if line.startswith("{"):
        d = json.loads(line); print(sys.argv[1], "fwd %.4f bwd %.4f step %.4f isect %.4f sched %s" % (d["roofline"]["kernel_ms"], d["roofline"]["backward"]["kernel_ms"], d["ms_per_step"], d["extras"]["intersect_only_ms"], d["extras"].get("adaptive_schedule")))
'
for of in 4 2; do for r in 0 1; do for ad in 0 1; do
  HF_B200_AS_MAX_GPW=100000 HF_BENCH_ADAPTIVE=$ad HF_BENCH_SHARD_OF=$of HF_BENCH_SHARD_RANK=$r python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "$show" "rank $r of $of adaptive=$ad"
done; done; done
