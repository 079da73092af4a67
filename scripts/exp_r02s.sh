#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/r02s_pytest.log 2>&1; grep -n "AssertionError: (" gpurun_out/r02s_pytest.log | cut -c1-300; tail -8 gpurun_out/r02s_pytest.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02s_smoke.log 2>&1; tail -6 gpurun_out/r02s_smoke.log
timeout 900 python bench.py > gpurun_out/r02s_bench_cfg4.json 2> gpurun_out/r02s_bench_cfg4.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02s_bench_cfg4.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d['clocks'], d['cpu_baseline'])
print({k:v for k,v in d['roofline'].items() if k not in ('by_shape','shares')})
for k,v in d['roofline']['by_shape'].items(): print(k,v)
print(d['roofline']['shares'])
PY
tail -3 gpurun_out/r02s_bench_cfg4.err
