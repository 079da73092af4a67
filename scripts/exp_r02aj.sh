#!/bin/bash
# session AJ (8 GPUs): the 8-GPU bench line of the final build
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02_bench_cfg4_n8.json 2> gpurun_out/r02_bench_cfg4_n8.err; cut -c1-500 gpurun_out/r02_bench_cfg4_n8.json; tail -2 gpurun_out/r02_bench_cfg4_n8.err | cut -c1-200
