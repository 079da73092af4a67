#!/bin/bash
# session AH: stage timeline of the weight-gradient kernel (stamps), persistent-forward L2 prefetch
mkdir -p gpurun_out
KV_DEBUG_BITS=32 timeout 120 python scripts/bench_kan_layer.py --inf 3072 --outf 768 --T 66192 --what fwdbwd --iters 1 2>&1 | grep "^dW st" | head -16
timeout 120 python scripts/bench_msa.py 2>&1 | tail -1 | cut -c1-400
