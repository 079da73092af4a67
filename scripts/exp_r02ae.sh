#!/bin/bash
# session AE: kan_tc.cu with -use_fast_math (parity of every bf16 test, then the non-effkan families)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/r02ae_pytest.log 2>&1; grep -n "^FAILED" gpurun_out/r02ae_pytest.log | head -20 | cut -c1-300; tail -2 gpurun_out/r02ae_pytest.log | cut -c1-300
timeout 600 python scripts/sweep_kan_layer.py --tokens 65536 --families cheby,fastkan,sine,fourier 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if not l.startswith('{'): print(l[:200]); continue
    r=json.loads(l)
    if 'fwd_tflops' in r and (r['in'],r['out']) in ((768,768),(768,3072)): print(r['family'],r['in'],r['out'],'G',r['grid'],'fwd',r['fwd_tflops'],'dX',r['bwd_input_tflops'],'dW',r['bwd_weight_tflops'],'fb',r['fwdbwd_tflops_gemm_kernels'])
    elif 'error' in r: print(l[:200])
"
for shape in "768 3072" "3072 768"; do set -- $shape
  timeout 120 python scripts/bench_kan_layer.py --inf $1 --outf $2 --T 66192 --what fwdbwd 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['in'],d['out'],{k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict) and v['tflops']})"
done
