#!/bin/bash
# Round-2 GPU check C (2 GPUs): the 2-rank NCCL tests, then the config-2 / config-3 training-step timing with the
# first-bucket all-reduce overlapped (default) and not.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q -s > gpurun_out/multirank.log 2>&1; echo rc=$? >> gpurun_out/multirank.log; tail -n 12 gpurun_out/multirank.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_r2_2gpu.json 2> gpurun_out/bench_r2_2gpu.err; echo bench2 rc=$?; tail -n 3 gpurun_out/bench_r2_2gpu.err
