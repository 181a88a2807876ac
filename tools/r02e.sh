cd /root/repo
nvidia-smi topo -m 2>&1 | head -12
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/r02f_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -25 gpurun_out/r02f_pytest_multi.log
HF_BENCH_PEER_ATOMICS=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu > gpurun_out/r02f_bench_n2.json 2> gpurun_out/r02f_bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/r02f_bench_n2.err; cat gpurun_out/r02f_bench_n2.json
