#!/bin/bash
mkdir -p gpurun_out
for shape in "768 3072"; do
  set -- $shape
  for bits in 0 92 348 256 264 280; do
    echo -n "bwd_weight $1->$2 bits=$bits: "
    KV_DEBUG_BITS=$bits timeout 300 python scripts/bench_kan_layer.py --what fwdbwd --T 66192 --inf $1 --outf $2 --iters 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['kv_kan_bwd_weight'])"
  done
done > gpurun_out/r02q_bits.log 2>&1
cat gpurun_out/r02q_bits.log
