#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --no-cpu > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 900 gpurun_out/bench_n1.json; tail -2 gpurun_out/bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --no-cpu > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -c 900 gpurun_out/bench_n2.json; tail -3 gpurun_out/bench_n2.err
