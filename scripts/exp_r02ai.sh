#!/bin/bash
# session AI: weight-gradient kernel with its own X-loader warp and a software-pipelined MMA loop
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_big_parity.py -m gpu -q -x > gpurun_out/r02ai_pytest.log 2>&1; tail -2 gpurun_out/r02ai_pytest.log | cut -c1-200
for shape in "768 3072" "3072 768" "768 768"; do set -- $shape
  timeout 120 python scripts/bench_kan_layer.py --inf $1 --outf $2 --T 66192 --what fwdbwd 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['in'],d['out'],{k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict) and v['tflops']})"
done
timeout 120 python scripts/bench_msa.py 2>&1 | tail -1 | cut -c1-400
