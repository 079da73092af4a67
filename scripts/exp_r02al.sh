#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn5.py -m gpu -q 2>&1 | tail -2 | cut -c1-200
for v in 0 1; do KV_ATTN_D_FROM_O=$v timeout 120 python scripts/bench_msa.py 2>&1 | tail -1 | cut -c1-400 | sed "s/^/D_FROM_O=$v /"; done
