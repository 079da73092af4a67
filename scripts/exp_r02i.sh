#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/debug_stamps.py 32 768 3072 > gpurun_out/r02i_stamps_768x3072.log 2>&1; head -64 gpurun_out/r02i_stamps_768x3072.log | cut -c1-220
timeout 200 python scripts/debug_stamps.py 60 768 3072 > gpurun_out/r02i_stamps_768x3072_skel.log 2>&1; head -40 gpurun_out/r02i_stamps_768x3072_skel.log | cut -c1-220
