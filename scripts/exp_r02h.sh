#!/bin/bash
mkdir -p gpurun_out
timeout 120 ./build/mma_pace > gpurun_out/r02h_mma_pace.log 2>&1; cut -c1-200 gpurun_out/r02h_mma_pace.log
for sh in "2 50 2 32" "1 5 2 8" "2 129 3 16" "2 256 1 64" "2 17 8 24"; do
  echo "== $sh"; CUDA_LAUNCH_BLOCKING=1 timeout 120 python scripts/debug_attn5.py $sh 2>&1 | tail -3
done > gpurun_out/r02h_attn_debug.log 2>&1
cat gpurun_out/r02h_attn_debug.log
timeout 200 python scripts/debug_attn5.py 336 197 12 64 20 > gpurun_out/r02h_attn_time.log 2>&1; cat gpurun_out/r02h_attn_time.log
timeout 300 python -m pytest tests/test_gpu_attn5.py -q > gpurun_out/r02h_attn5.log 2>&1; tail -12 gpurun_out/r02h_attn5.log | cut -c1-300
timeout 600 python bench.py --no-cpu --steps 10 > gpurun_out/r02h_bench_cfg4.json 2> gpurun_out/r02h_bench_cfg4.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02h_bench_cfg4.json'))
print(d['value'], d['ms_per_step'], d['e2e'])
for k,v in d['roofline']['by_shape'].items(): print(k,v)
print(d['roofline']['shares'])
PY
tail -3 gpurun_out/r02h_bench_cfg4.err
