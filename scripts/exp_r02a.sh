#!/bin/bash
# round 2, GPU session A: new parity tests, baseline bench, which stage bounds the forward GEMM (debug bits)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1; tail -5 gpurun_out/r02a_pytest.log
timeout 900 python -m pytest tests/test_big_parity.py -m gpu -q -s > gpurun_out/r02a_bigparity.log 2>&1; grep -E "^big_|passed|failed" gpurun_out/r02a_bigparity.log | tail -30
timeout 600 python bench.py > gpurun_out/r02a_bench_cfg4.json 2> gpurun_out/r02a_bench_cfg4.err; cut -c1-400 gpurun_out/r02a_bench_cfg4.json
for shape in "768 3072" "3072 768"; do
  set -- $shape
  for bits in 0 2 4 66 128; do
    echo "fwd $1->$2 bits=$bits"
    KV_DEBUG_BITS=$bits timeout 300 python scripts/bench_kan_layer.py --what fwd --T 66192 --inf $1 --outf $2 --iters 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:v for k,v in d.items() if k.startswith('kv_kan_fwd')})"
  done
  for bits in 0 256 512 768; do
    echo "fwdbwd $1->$2 bits=$bits"
    KV_DEBUG_BITS=$bits timeout 300 python scripts/bench_kan_layer.py --what fwdbwd --T 66192 --inf $1 --outf $2 --iters 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:v for k,v in d.items() if k.startswith('kv_kan_bwd')})"
  done
done > gpurun_out/r02a_bits.log 2>&1
cat gpurun_out/r02a_bits.log
timeout 200 python scripts/debug_stamps.py 32 3072 768 > gpurun_out/r02a_stamps_3072x768.log 2>&1; head -60 gpurun_out/r02a_stamps_3072x768.log
