#!/bin/bash
# session AC: where does the forward epilogue's time go (stamps), what bounds the generic weight-gradient kernel
mkdir -p gpurun_out
for sh in "768 3072" "3072 768"; do set -- $sh
KV_FWD_BN=512 KV_T=66192 KV_STAMPS_SHORT=1 timeout 120 python scripts/debug_stamps.py 32 $1 $2 2>&1 | grep -v "^CTA rank 1" | head -4 | cut -c1-300
done
for bits in 0 8; do
  KV_DEBUG_BITS=$bits timeout 120 python scripts/bench_kan_layer.py --grid 10 --inf 768 --outf 768 --T 65536 --what fwdbwd 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('bits $bits', {k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict) and v['tflops']})"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'kan_tc_bwd_weight_kernel' --launch-skip 3 -c 1 \
  -o gpurun_out/r02ac_dw_g10 -f python scripts/bench_kan_layer.py --grid 10 --inf 768 --outf 768 --T 16384 --what fwdbwd --iters 2 > gpurun_out/r02ac_ncu.log 2>&1; tail -2 gpurun_out/r02ac_ncu.log | cut -c1-200
timeout 300 python -m pytest tests/test_gpu_attn_tiled.py -m gpu -q 2>&1 | tail -2
