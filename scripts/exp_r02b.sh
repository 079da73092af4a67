#!/bin/bash
# round 2, GPU session B: MMA pace microbenchmark, the reworked GEMM kernels per shape, full parity suite, bench
mkdir -p gpurun_out
timeout 120 ./build/mma_pace > gpurun_out/r02b_mma_pace.log 2>&1; cat gpurun_out/r02b_mma_pace.log
for shape in "768 3072" "3072 768"; do
  set -- $shape
  timeout 300 python scripts/bench_kan_layer.py --what fwdbwd --T 66192 --inf $1 --outf $2 --iters 5 2>&1 | tail -1
done > gpurun_out/r02b_layers.log 2>&1
for fam in "effkan 10" "effkan 20" "fastkan 20" "sine 20" "sine 28" "fourier 10" "cheby 20"; do
  set -- $fam
  timeout 300 python scripts/bench_kan_layer.py --family $1 --grid $2 --what fwdbwd --T 65536 --inf 768 --outf 768 --iters 3 2>&1 | tail -1
done >> gpurun_out/r02b_layers.log 2>&1
cat gpurun_out/r02b_layers.log | cut -c1-600
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02b_pytest.log 2>&1; tail -40 gpurun_out/r02b_pytest.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/r02b_bench_cfg4.json 2> gpurun_out/r02b_bench_cfg4.err; cut -c1-600 gpurun_out/r02b_bench_cfg4.json; tail -3 gpurun_out/r02b_bench_cfg4.err
