#!/bin/bash
# ncu evidence for profiles/: (1) launch list of ONE step of the default bench command, (2) full captures of
# the KAN GEMM kernels and the attention kernels inside that same step.  usage: bash scripts/gpu_profile.sh <tag>
tag=${1:-r01}
mkdir -p gpurun_out
KV_FAST_INIT=1 KV_VERBOSE=1 KV_NCU_RANGE=1 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
  --log-file gpurun_out/${tag}_cfg4_launches.csv python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_launch.log 2>&1
tail -2 gpurun_out/${tag}_ncu_launch.log | cut -c1-200
wc -l gpurun_out/${tag}_cfg4_launches.csv
# one launch of each of the big kernels from block 6 of the same step (skip the first launches: patch embed)
KV_FAST_INIT=1 KV_VERBOSE=1 KV_NCU_RANGE=1 timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on \
  -k regex:'kan_tc_fwd_kernel|effkan8_bwd_input_kernel|kan_tc_bwd_weight_kernel|attn_tc_fwd|attn_tc_bwd' \
  --launch-skip 12 -c 12 -o gpurun_out/${tag}_cfg4_kernels -f python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_full.log 2>&1
tail -2 gpurun_out/${tag}_ncu_full.log | cut -c1-200
