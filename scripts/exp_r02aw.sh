#!/bin/bash
for n in 16 8; do
  for shape in "768 3072" "3072 768"; do set -- $shape
    KV_BW_NPW=$n timeout 120 python scripts/bench_kan_layer.py --inf $1 --outf $2 --T 66192 --what fwdbwd 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('NPW=$n',d['in'],d['out'],'dW',d['kv_kan_bwd_weight'])"
  done
  KV_BW_NPW=$n timeout 120 python scripts/bench_msa.py 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('NPW=$n msa dW',d['kv_kan_bwd_weight'])"
done
