#!/bin/bash
# session AM: cfg5 with overlapped end-to-end copies, cfg4_faithful at its new default batch
mkdir -p gpurun_out
timeout 600 python bench.py --config cfg5 > gpurun_out/r02_bench_cfg5.json 2> gpurun_out/r02_bench_cfg5.err; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_cfg5.json')); print(d['value'], d['ms_per_step'], d['e2e'], d['clocks'])"; tail -2 gpurun_out/r02_bench_cfg5.err | cut -c1-300
timeout 600 python bench.py --config cfg4_faithful --no-cpu --steps 10 > gpurun_out/r02_bench_cfg4_faithful.json 2> gpurun_out/r02_bench_cfg4_faithful.err; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_cfg4_faithful.json')); print(d['value'], d['ms_per_step'], d['e2e'], d['clocks'], d['config']['workload'])"
