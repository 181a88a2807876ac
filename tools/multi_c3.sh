#!/bin/bash
# Dev tool (under gpurun --gpus N): the C3 bench line at N GPUs.   gpurun --gpus 4 -- "bash tools/multi_c3.sh 4 r02"
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
N=${1:-4}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu 2> gpurun_out/${2:-r02}_bench_n$N.err | grep "^{" > gpurun_out/${2:-r02}_bench_n$N.json; echo "bench C3 n$N rc=$?"; tail -3 gpurun_out/${2:-r02}_bench_n$N.err
