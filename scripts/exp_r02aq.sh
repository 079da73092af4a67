#!/bin/bash
# session AQ: coalesced nn.Linear producers
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py tests/test_big_parity.py -m gpu -q > gpurun_out/r02aq_pytest.log 2>&1; grep -n "^FAILED" gpurun_out/r02aq_pytest.log | head; tail -2 gpurun_out/r02aq_pytest.log | cut -c1-200
for shape in "768 768" "768 3072" "3072 768" "192 768"; do set -- $shape
  timeout 120 python scripts/bench_kan_layer.py --family linear --inf $1 --outf $2 --T 65536 --what fwdbwd 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('linear',d['in'],d['out'],{k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict) and v['tflops']})"
done
timeout 600 python bench.py --config cfg4_faithful --no-cpu --steps 10 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('faithful', d['value'], d['ms_per_step'])"
timeout 300 python bench.py --config cfg3_fast --no-cpu --steps 10 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('cfg3_fast', d['value'], d['ms_per_step'])"
