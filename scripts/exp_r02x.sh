#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/r02x_pytest.log 2>&1; tail -5 gpurun_out/r02x_pytest.log | cut -c1-300
timeout 300 python scripts/sweep_kan_layer.py --families effkan --tokens 65536 --grids 10,20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if not l.startswith('{'): print(l[:200]); continue
    r=json.loads(l)
    if 'fwd_tflops' in r: print(r['T'],r['in'],r['out'],'G',r['grid'],'fwd',r['fwd_tflops'],'dX',r['bwd_input_tflops'],'dW',r['bwd_weight_tflops'],'fb',r['fwdbwd_tflops_gemm_kernels'])
    else: print(l[:200])
"
timeout 300 python scripts/sweep_kan_layer.py --families effkan,sine --tokens 1048576 --grids 5 2>&1 | cut -c1-330
timeout 600 python bench.py --config cfg5 > gpurun_out/r02x_bench_cfg5.json 2> gpurun_out/r02x_bench_cfg5.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02x_bench_cfg5.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d['clocks'], d['cpu_baseline'])
for k,v in d['roofline']['by_shape'].items(): print(k,v)
PY
timeout 900 python bench.py --no-cpu > gpurun_out/r02x_bench_cfg4.json 2> gpurun_out/r02x_bench_cfg4.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02x_bench_cfg4.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d['clocks'], d['gpu_launches'])
for k,v in list(d['roofline']['by_shape'].items())[:10]: print(k,v)
PY
