#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/r02ad_pytest.log 2>&1; grep -n "^FAILED" gpurun_out/r02ad_pytest.log | head -20 | cut -c1-300; tail -2 gpurun_out/r02ad_pytest.log | cut -c1-300
timeout 600 python scripts/sweep_kan_layer.py --tokens 65536 --families effkan 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if not l.startswith('{'): print(l[:200]); continue
    r=json.loads(l)
    if 'fwd_tflops' in r: print(r['family'],r['in'],r['out'],'G',r['grid'],'fwd',r['fwd_tflops'],'dX',r['bwd_input_tflops'],'dW',r['bwd_weight_tflops'],'fb',r['fwdbwd_tflops_gemm_kernels'])
    elif 'error' in r: print(l[:200])
"
