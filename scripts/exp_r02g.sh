#!/bin/bash
mkdir -p gpurun_out
timeout 60 ./build/bulk_pace > gpurun_out/r02g_bulk_pace.log 2>&1; cat gpurun_out/r02g_bulk_pace.log | cut -c1-200
timeout 300 python -m pytest tests/test_gpu_attn5.py -q -x > gpurun_out/r02g_attn5.log 2>&1; tail -25 gpurun_out/r02g_attn5.log | cut -c1-300
timeout 900 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_attn5.py -q -x -k "5-2-8 or 17-8-24" > gpurun_out/r02g_attn5_memcheck.log 2>&1; tail -15 gpurun_out/r02g_attn5_memcheck.log | cut -c1-300
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02g_pytest.log 2>&1; tail -30 gpurun_out/r02g_pytest.log | cut -c1-400
