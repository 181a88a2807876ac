#!/bin/bash
# Dev tool (under gpurun --gpus N): bench.py on N GPUs, once per environment setting given as arguments ("A=1 B=2" ...), short summary per run.
# usage: tools/nbench.sh N "ENV1" "ENV2" ...
cd "$(dirname "$0")/.."
N=$1; shift
mkdir -p gpurun_out
i=0
for e in "$@"; do
  i=$((i+1))
  env $e python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+i)) bench.py --gpus $N --steps 5 --warmup 3 --no-cpu > gpurun_out/n${N}_$i.json 2> gpurun_out/n${N}_$i.err
  echo "rc=$? [$e]"
  python - <<PY
import json
for l in open("gpurun_out/n${N}_$i.json"):
    if l.startswith("{"):
        d = json.loads(l); x = d["extras"]; r = d["roofline"]
        print("  value %.0f Mrays/s step %.3f ms | fwd %.3f bwd %.3f allreduce %.3f full %.3f | render step %.3f kernel %.3f | e2e %.0f" % (
            d["value"], d["ms_per_step"], r["kernel_ms"], r["backward"]["kernel_ms"], x.get("allreduce_ms", -1), x["full_step_ms"],
            x["render_step"]["ms_per_step"], x["render_step"]["roofline"]["kernel_ms"], (d.get("e2e") or {}).get("value", 0)))
PY
  tail -2 gpurun_out/n${N}_$i.err | cut -c1-200
done
