#!/bin/bash
# Round-2 scaling check on N GPUs of one box: the bench line as the driver launches it.
set -u
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_r2_${N}gpu.json 2> gpurun_out/bench_r2_${N}gpu.err; echo "bench$N rc=$?"; tail -n 3 gpurun_out/bench_r2_${N}gpu.err
