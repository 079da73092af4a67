#!/bin/bash
mkdir -p gpurun_out
for shape in "768 3072" "3072 768"; do
  set -- $shape
  for bits in 0 4 8 16 64 12 20 24 28 92; do
    echo -n "bwd_weight $1->$2 bits=$bits: "
    KV_DEBUG_BITS=$bits timeout 300 python scripts/bench_kan_layer.py --what fwdbwd --T 66192 --inf $1 --outf $2 --iters 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['kv_kan_bwd_weight'], d['kv_kan_fwd'])"
  done
done > gpurun_out/r02o_bits.log 2>&1
cat gpurun_out/r02o_bits.log
timeout 900 python -m pytest tests/test_big_parity.py tests/test_gpu_attn5.py -m gpu -q > gpurun_out/r02o_pytest.log 2>&1; tail -8 gpurun_out/r02o_pytest.log | cut -c1-300
