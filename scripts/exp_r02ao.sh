#!/bin/bash
# session AO: final validation of the round-2 build (whole GPU suite, smoke) and the full config-5 sweep on it
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; tail -2 gpurun_out/r02_pytest_gpu.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; tail -4 gpurun_out/r02_smoke.log | cut -c1-200
timeout 2400 python scripts/sweep_kan_layer.py > gpurun_out/r02_kan_layer_sweep.jsonl 2> gpurun_out/r02_kan_layer_sweep.err; wc -l gpurun_out/r02_kan_layer_sweep.jsonl; grep -c error gpurun_out/r02_kan_layer_sweep.jsonl
