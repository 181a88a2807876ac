#!/bin/bash
# Dev tool (under gpurun --gpus N): the C3 bench line at N GPUs for several image-tile sizes.   gpurun --gpus 8 -- "bash tools/multi_c3_tiles.sh 8 r02 '32 16'"
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
N=${1:-8}
for tile in ${3:-32 16}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$tile bench.py --gpus $N --tile $tile --steps 20 --warmup 3 --no-cpu --no-e2e 2> gpurun_out/${2:-r02}_bench_n${N}_t$tile.err | grep "^{" > gpurun_out/${2:-r02}_bench_n${N}_t$tile.json; echo "bench C3 n$N tile $tile rc=$?"
  python tools/show_bench.py gpurun_out/${2:-r02}_bench_n${N}_t$tile.json 2>/dev/null | grep -v "^collective\|^e2e\|^cpu"
done
