cd /root/repo
show='
import sys, json
for line in sys.stdin:
    if line.startswith("{"):
        d = json.loads(line); print(sys.argv[1], "fwd %.4f bwd %.4f step %.4f isect %.4f sched %s" % (d["roofline"]["kernel_ms"], d["roofline"]["backward"]["kernel_ms"], d["ms_per_step"], d["extras"]["intersect_only_ms"], d["extras"].get("adaptive_schedule")))
'
for tm in 24:24 12:12 8:24 4:8; do
  for r in 0 5; do
    HF_B200_TAIL=${tm%%:*} HF_B200_MIDTAIL=${tm##*:} HF_BENCH_SHARD_OF=8 HF_BENCH_SHARD_RANK=$r python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "$show" "rank $r tail:mid=$tm"
  done
done
