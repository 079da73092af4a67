#!/bin/bash
mkdir -p gpurun_out
for sh in "768 3072" "3072 768"; do set -- $sh
KV_FWD_BN=512 KV_T=66192 KV_STAMPS_SHORT=1 timeout 120 python scripts/debug_stamps.py 32 $1 $2 > gpurun_out/r02u_stamps_$1x$2.log 2>&1; cat gpurun_out/r02u_stamps_$1x$2.log | cut -c1-400
done
