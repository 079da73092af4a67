#!/bin/bash
# session T: persistent forward kernel (BN <= 256) vs the one-tile-per-pair kernel (KV_FWD_BN=512)
mkdir -p gpurun_out
L=gpurun_out/r02t_layer.log; : > $L
for bn in 256 512; do
  for shape in "768 3072 66192" "3072 768 66192" "768 768 66192"; do
    set -- $shape
    echo "BN=$bn $1->$2" >> $L
    KV_FWD_BN=$bn timeout 120 python scripts/bench_kan_layer.py --inf $1 --outf $2 --T $3 --what fwd >> $L 2>&1 || echo "FAILED rc=$?" >> $L
  done
  KV_FWD_BN=$bn timeout 120 python scripts/bench_kan_layer.py --family linear --inf 768 --outf 3072 --T 66192 --what fwdbwd >> $L 2>&1 || echo "FAILED rc=$?" >> $L
done
python - <<'PY'
import json
for l in open('gpurun_out/r02t_layer.log'):
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print({k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict)})
    else: print(l[:200])
PY
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/r02t_pytest.log 2>&1; grep -n "AssertionError: (" gpurun_out/r02t_pytest.log | cut -c1-300; tail -8 gpurun_out/r02t_pytest.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02t_smoke.log 2>&1; tail -3 gpurun_out/r02t_smoke.log
for bn in 256 512; do
KV_FWD_BN=$bn timeout 900 python bench.py > gpurun_out/r02t_bench_cfg4_bn$bn.json 2> gpurun_out/r02t_bench_cfg4_bn$bn.err; BN=$bn python - <<'PY'
import json,os
d=json.load(open('gpurun_out/r02t_bench_cfg4_bn%s.json'%os.environ['BN']))
print(os.environ['BN'], d['value'], d['ms_per_step'], d['e2e'], d['clocks'])
print({k:v for k,v in d['roofline'].items() if k not in ('by_shape','shares')})
for k,v in d['roofline']['by_shape'].items(): print(k,v)
PY
done
