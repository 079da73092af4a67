#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; tail -2 gpurun_out/r02_pytest_gpu.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; tail -2 gpurun_out/r02_smoke.log | cut -c1-200
timeout 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_n1.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'], d['gpu_launches'], d['roofline']['frac'], d['roofline']['kernel']); [print(k,v) for k,v in list(d['roofline']['by_shape'].items())[:9]]"
timeout 600 python bench.py --config cfg4_faithful --no-cpu --steps 10 > gpurun_out/r02_bench_cfg4_faithful.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_cfg4_faithful.json')); print('faithful', d['value'], d['ms_per_step'])"
for c in cfg3_fast cfg3_sine; do timeout 300 python bench.py --config $c --no-cpu --steps 10 > gpurun_out/r02_bench_$c.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_$c.json')); print('$c', d['value'], d['ms_per_step'])"; done
timeout 300 python bench.py --config cfg2 --bf16 --no-cpu --steps 10 > gpurun_out/r02_bench_cfg2_bf16.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_cfg2_bf16.json')); print('cfg2_bf16', d['value'], d['ms_per_step'])"
