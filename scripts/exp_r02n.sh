#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/debug_stamps.py 32 768 3072 > gpurun_out/r02n_stamps_768x3072.log 2>&1; head -66 gpurun_out/r02n_stamps_768x3072.log | cut -c1-200
timeout 200 python scripts/debug_parity_2blk.py > gpurun_out/r02n_ab_fused.log 2>&1; tail -14 gpurun_out/r02n_ab_fused.log
KV_NO_FUSED_ATTN=1 timeout 200 python scripts/debug_parity_2blk.py > gpurun_out/r02n_ab_old.log 2>&1; tail -14 gpurun_out/r02n_ab_old.log
