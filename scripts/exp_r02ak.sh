#!/bin/bash
# session AK: attention backward with D from the forward output
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn5.py tests/test_gpu_tc.py tests/test_big_parity.py -m gpu -q > gpurun_out/r02ak_pytest.log 2>&1; grep -n "^FAILED" gpurun_out/r02ak_pytest.log | head; tail -2 gpurun_out/r02ak_pytest.log | cut -c1-200
timeout 120 python scripts/bench_msa.py 2>&1 | tail -1 | cut -c1-400
timeout 600 python bench.py --config cfg4_faithful --no-cpu --steps 10 --batch 336 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('faithful b336', d['value'], d['ms_per_step'])"
