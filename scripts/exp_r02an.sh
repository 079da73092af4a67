#!/bin/bash
# session AN: basis splits in the generic input-gradient kernel (more than 32 basis functions on the tensor cores)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_big_parity.py tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/r02an_pytest.log 2>&1; grep -n "^FAILED\|Error" gpurun_out/r02an_pytest.log | head -10 | cut -c1-250; tail -2 gpurun_out/r02an_pytest.log | cut -c1-200
timeout 300 python scripts/sweep_kan_layer.py --tokens 65536 --families fourier --grids 20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if not l.startswith('{'): print(l[:200]); continue
    r=json.loads(l)
    if 'fwd_tflops' in r: print(r['family'],r['in'],r['out'],'G',r['grid'],'M',r['M'],'fwd',r['fwd_tflops'],'dX',r['bwd_input_tflops'],'dW',r['bwd_weight_tflops'],'fb',r['fwdbwd_tflops_gemm_kernels'])
    elif 'error' in r: print(l[:200])
"
