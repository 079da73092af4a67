#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --config cfg5 > gpurun_out/r02w_bench_cfg5.json 2> gpurun_out/r02w_bench_cfg5.err; cut -c1-1500 gpurun_out/r02w_bench_cfg5.json; tail -3 gpurun_out/r02w_bench_cfg5.err
timeout 900 python bench.py > gpurun_out/r02w_bench_cfg4.json 2> gpurun_out/r02w_bench_cfg4.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02w_bench_cfg4.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d['clocks'], d['gpu_launches'])
for k,v in list(d['roofline']['by_shape'].items())[:14]: print(k,v)
print(d['roofline']['shares'])
PY
timeout 600 python bench.py --config cfg4_faithful --no-cpu > gpurun_out/r02w_bench_cfg4_faithful.json 2> gpurun_out/r02w_bench_cfg4_faithful.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02w_bench_cfg4_faithful.json'))
print('faithful', d['value'], d['ms_per_step'], d['clocks'])
for k,v in list(d['roofline']['by_shape'].items())[:10]: print(k,v)
PY
timeout 2400 python scripts/sweep_kan_layer.py > gpurun_out/r02_kan_layer_sweep.jsonl 2> gpurun_out/r02_kan_layer_sweep.err; wc -l gpurun_out/r02_kan_layer_sweep.jsonl; grep -c error gpurun_out/r02_kan_layer_sweep.jsonl; tail -3 gpurun_out/r02_kan_layer_sweep.err
timeout 600 python bench.py --impl reference --config cfg5 --sweep-cpu > gpurun_out/r02_kan_layer_sweep_cpu.jsonl 2>/dev/null; wc -l gpurun_out/r02_kan_layer_sweep_cpu.jsonl
