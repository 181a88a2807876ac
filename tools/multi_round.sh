#!/bin/bash
# Dev tool (under gpurun --gpus N): the multi-GPU lines of a round -- C3, C4 and the 2^30 column of C5 at N GPUs, then the 2-GPU tests.
#   gpurun --gpus 8 --timeout 1500 -- "bash tools/multi_round.sh 8 r02"
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu > gpurun_out/${2:-r02}_bench_n$N.json 2> gpurun_out/${2:-r02}_bench_n$N.err; echo "bench C3 n$N rc=$?"; tail -3 gpurun_out/${2:-r02}_bench_n$N.err
timeout 900 $TR --master-port 29522 bench.py --config C4 --gpus $N --steps 5 --warmup 3 --no-cpu > gpurun_out/${2:-r02}_c4_n$N.json 2> gpurun_out/${2:-r02}_c4_n$N.err; echo "bench C4 n$N rc=$?"; tail -3 gpurun_out/${2:-r02}_c4_n$N.err
HF_C5_LOG2=30 timeout 900 $TR --master-port 29523 bench.py --config C5 --gpus $N --no-cpu > gpurun_out/${2:-r02}_c5_n$N.json 2> gpurun_out/${2:-r02}_c5_n$N.err; echo "bench C5 n$N rc=$?"; tail -3 gpurun_out/${2:-r02}_c5_n$N.err
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -q 2>&1 | tail -2
