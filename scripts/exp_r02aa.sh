#!/bin/bash
# session AA: the no-alloc entry point (alone in a fresh process, then after the whole suite), then the config-5 sweep
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02aa_noalloc_alone.log 2>&1; tail -15 gpurun_out/r02aa_noalloc_alone.log | cut -c1-300
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02aa_pytest.log 2>&1; tail -6 gpurun_out/r02aa_pytest.log | cut -c1-300
timeout 2400 python scripts/sweep_kan_layer.py > gpurun_out/r02_kan_layer_sweep.jsonl 2> gpurun_out/r02_kan_layer_sweep.err; wc -l gpurun_out/r02_kan_layer_sweep.jsonl; grep -c error gpurun_out/r02_kan_layer_sweep.jsonl; tail -3 gpurun_out/r02_kan_layer_sweep.err
timeout 600 python bench.py --impl reference --config cfg5 --sweep-cpu > gpurun_out/r02_kan_layer_sweep_cpu.jsonl 2>/dev/null; wc -l gpurun_out/r02_kan_layer_sweep_cpu.jsonl
