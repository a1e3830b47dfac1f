#!/bin/bash
# Scaling sanity run on one 8-GPU box: bench.py as the driver launches it at N = 8 (and N = 4).
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 1200 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
