#!/bin/bash
# session AS: eight epilogue warps in the persistent forward kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py tests/test_big_parity.py tests/test_gpu_attn5.py -m gpu -q > gpurun_out/r02as_pytest.log 2>&1; grep -n "^FAILED" gpurun_out/r02as_pytest.log | head; tail -2 gpurun_out/r02as_pytest.log | cut -c1-200
timeout 120 python scripts/bench_msa.py 2>&1 | tail -1 | cut -c1-400
for shape in "768 768" "768 3072"; do set -- $shape
  timeout 120 python scripts/bench_kan_layer.py --family linear --inf $1 --outf $2 --T 65536 --what fwdbwd 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('linear',d['in'],d['out'],{k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict) and v['tflops']})"
done




