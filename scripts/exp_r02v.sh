#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r02v_pytest.log 2>&1; tail -5 gpurun_out/r02v_pytest.log | cut -c1-300
L=gpurun_out/r02v_layer.log; : > $L
for shape in "768 3072" "3072 768" "768 768"; do set -- $shape
  timeout 120 python scripts/bench_kan_layer.py --family linear --inf $1 --outf $2 --T 66192 --what fwdbwd >> $L 2>&1 || echo "FAILED rc=$?" >> $L
done
timeout 120 python scripts/bench_kan_layer.py --inf 64 --outf 192 --T 66192 --what fwd >> $L 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/r02v_layer.log'):
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['family'], d['in'], d['out'], {k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict)})
    else: print(l[:200])
PY
P=gpurun_out/r02v_power.jsonl; : > $P
timeout 60 python scripts/power_probe.py --inf 768 --outf 3072 >> $P 2>&1
timeout 60 python scripts/power_probe.py --inf 3072 --outf 768 >> $P 2>&1
timeout 60 python scripts/power_probe.py --inf 3072 --outf 768 --zeros >> $P 2>&1
timeout 60 python scripts/power_probe.py --inf 3072 --outf 768 --what fwdbwd >> $P 2>&1
timeout 60 python scripts/power_probe.py --family linear --inf 3072 --outf 768 >> $P 2>&1
cat $P | cut -c1-400
