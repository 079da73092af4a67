#!/bin/bash
mkdir -p gpurun_out
for shape in "768 3072" "3072 768"; do
  set -- $shape
  for bits in 0 4 8 16 12 20 24 28; do
    echo -n "fwd $1->$2 bits=$bits: "
    KV_DEBUG_BITS=$bits timeout 300 python scripts/bench_kan_layer.py --what fwd --T 66192 --inf $1 --outf $2 --iters 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['kv_kan_fwd'])"
  done
done > gpurun_out/r02f_bits.log 2>&1
cat gpurun_out/r02f_bits.log
# full ncu capture of the three GEMM kernels on the 768->3072 layer (memory tables)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'kan_tc_fwd_kernel|effkan8_bwd_input_kernel|kan_tc_bwd_weight_kernel' \
  --launch-skip 9 -c 3 -o gpurun_out/r02f_kan_layer -f python scripts/bench_kan_layer.py --what fwdbwd --T 66192 --iters 3 > gpurun_out/r02f_ncu.log 2>&1
tail -3 gpurun_out/r02f_ncu.log; ls -la gpurun_out/*.ncu-rep
