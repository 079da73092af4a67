#!/bin/bash
# session AF (2 GPUs): the NCCL trainer test, the 2-GPU bench line, and the 64 Ki-token sweep column on the final build
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_trainer.py -m gpu -q > gpurun_out/r02_pytest_2gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_2gpu.log | cut -c1-200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_bench_cfg4_n2.json 2> gpurun_out/r02_bench_cfg4_n2.err; cut -c1-400 gpurun_out/r02_bench_cfg4_n2.json; tail -2 gpurun_out/r02_bench_cfg4_n2.err | cut -c1-200
timeout 900 python scripts/sweep_kan_layer.py --tokens 65536 > gpurun_out/r02_kan_layer_sweep_T65536_final.jsonl 2>/dev/null; wc -l gpurun_out/r02_kan_layer_sweep_T65536_final.jsonl
