#!/bin/bash
KV_DEBUG_BITS=8192 timeout 120 python scripts/bench_msa.py --iters 1 2>&1 | grep "^dW st" | head -16 | cut -c1-250
