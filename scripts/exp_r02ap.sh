#!/bin/bash
mkdir -p gpurun_out
KV_DEBUG_BITS=4096 timeout 120 python scripts/bench_kan_layer.py --family linear --inf 768 --outf 768 --T 65536 --what fwd --iters 1 2>&1 | grep "^ptile\|^epi\|linear" | head -12 | cut -c1-300
