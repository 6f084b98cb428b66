#!/bin/bash
# Round-2, 2 GPUs: the multi-rank parity test and the bench line as the driver launches it.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q -s > gpurun_out/p_multirank.log 2>&1; echo "multirank rc=$?"; tail -n 4 gpurun_out/p_multirank.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/p_bench_2gpu.json 2> gpurun_out/p_bench_2gpu.err; echo "bench2 rc=$?"; tail -c 400 gpurun_out/p_bench_2gpu.json
