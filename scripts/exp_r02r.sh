#!/bin/bash
mkdir -p gpurun_out
for shape in "768 3072" "3072 768" "768 768"; do
  set -- $shape
  timeout 300 python scripts/bench_kan_layer.py --what fwdbwd --T 66192 --inf $1 --outf $2 --iters 5 2>&1 | tail -1
done > gpurun_out/r02r_layers.log 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/r02r_layers.log'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print(d['family'], d['in'], d['out'], 'M',d['M'], {k:(v['ms'], v['tflops']) for k,v in d.items() if isinstance(v,dict) and k.startswith('kv_kan_') and k not in ('kv_kan_pack','kv_kan_dy_image','kv_kan_unpack_grads')})
PY
timeout 900 python -m pytest tests/test_big_parity.py tests/test_gpu_tc.py -m gpu -q > gpurun_out/r02r_pytest.log 2>&1; grep -n "AssertionError: (" gpurun_out/r02r_pytest.log | cut -c1-400; tail -8 gpurun_out/r02r_pytest.log | cut -c1-300
