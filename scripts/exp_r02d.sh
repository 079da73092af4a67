#!/bin/bash
mkdir -p gpurun_out
timeout 120 ./build/mma_pace > gpurun_out/r02d_mma_pace.log 2>&1; grep "grid 74" gpurun_out/r02d_mma_pace.log | cut -c1-200
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/r02d_pytest.log 2>&1; tail -30 gpurun_out/r02d_pytest.log | cut -c1-300
