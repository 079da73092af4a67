#!/bin/bash
# session AG: what paces the weight-gradient kernel (debug bits: 4 no MMA, 8 no basis evaluation, 16 no dY traffic)
mkdir -p gpurun_out
for bits in 0 4 8 16 24 28; do
  KV_DEBUG_BITS=$bits timeout 120 python scripts/bench_kan_layer.py --inf 3072 --outf 768 --T 66192 --what fwdbwd --iters 5 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('3072->768 bits $bits dW', d['kv_kan_bwd_weight'])"
done
for bits in 0 4 8 16 24; do
  KV_DEBUG_BITS=$bits timeout 120 python scripts/bench_msa.py 2>&1 | tail -1 | cut -c1-400 | sed "s/^/msa bits $bits /"
done
