#!/bin/bash
for shape in "768 768" "768 3072" "3072 768"; do set -- $shape
  timeout 120 python scripts/bench_kan_layer.py --family linear --inf $1 --outf $2 --T 65536 --what fwdbwd 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('linear',d['in'],d['out'],{k:(v['ms'],v['tflops']) for k,v in d.items() if isinstance(v,dict) and v['tflops']})"
done
