#!/bin/bash
# Round-2 evidence for profiles/ (one GPU, under gpurun): bench lines of every config, the reference arm, the ncu launch
# list of ONE step of the default bench command, full ncu captures of the GEMM kernels at the bench's layer shapes and of
# the attention / HBM-bound kernels inside a step, (the SASS excerpt profiles/r02_sass_blackwell_ops.txt is made on the build machine: cuobjdump needs no GPU).
# usage: bash scripts/gpu_evidence_r02.sh            (numbers printed by a run under ncu are never bench values)
O=gpurun_out; mkdir -p $O
timeout 1700 python -m pytest tests -m gpu -q > $O/r02_pytest_gpu.log 2>&1; grep -n "^FAILED" $O/r02_pytest_gpu.log | head; tail -1 $O/r02_pytest_gpu.log
timeout 900 python bench.py > $O/r02_bench_n1.json 2> $O/r02_bench_n1.err; cut -c1-400 $O/r02_bench_n1.json
timeout 600 python bench.py --config cfg5 > $O/r02_bench_cfg5.json 2> $O/r02_bench_cfg5.err; cut -c1-300 $O/r02_bench_cfg5.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_bench_reference_n1.json 2> $O/r02_bench_reference_n1.err; cut -c1-300 $O/r02_bench_reference_n1.json
timeout 600 python bench.py --impl reference --config cfg5 --steps 3 > $O/r02_bench_reference_cfg5.json 2>/dev/null; cut -c1-300 $O/r02_bench_reference_cfg5.json
for c in cfg1 cfg2 cfg3_fast cfg3_sine cfg4_faithful; do
  timeout 600 python bench.py --config $c --no-cpu --steps 10 > $O/r02_bench_$c.json 2> $O/r02_bench_$c.err; cut -c1-160 $O/r02_bench_$c.json; echo
done
timeout 300 python bench.py --config cfg1 --cuda-graph --no-cpu --steps 10 > $O/r02_bench_cfg1_graph.json 2>/dev/null
timeout 300 python bench.py --config cfg2 --bf16 --no-cpu --steps 10 > $O/r02_bench_cfg2_bf16.json 2>/dev/null
# (1) launch list of one step of the default bench command
KV_FAST_INIT=1 KV_NCU_RANGE=1 timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
  --log-file $O/r02_cfg4_b336_launches.csv python bench.py --no-cpu --steps 2 --warmup 3 > $O/r02_ncu_launch.log 2>&1
wc -l $O/r02_cfg4_b336_launches.csv
# (2) full captures: the three GEMM kernels at the bench's two MLP shapes (T = 66192) and the q/k/v layer
for shape in "768 3072" "3072 768"; do set -- $shape
  timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:'kan_tc_fwd_kernel|kan_tc_fwd_persistent_kernel|effkan8_bwd_input_kernel|kan_tc_bwd_weight_kernel' --launch-skip 9 -c 3 \
    -o $O/r02_layer_$1x$2 -f python scripts/bench_kan_layer.py --what fwdbwd --T 66192 --inf $1 --outf $2 --iters 3 > $O/r02_ncu_layer_$1x$2.log 2>&1
  # gpurun brings back at most 64 MiB: keep the raw counter table and the one-line-per-launch summary, not the report
  ncu -i $O/r02_layer_$1x$2.ncu-rep --page raw --csv > $O/r02_layer_$1x$2_raw.csv 2>/dev/null
  python scripts/ncu_summary.py $O/r02_layer_$1x$2.ncu-rep "kv_kan_fwd T66192 g1 $1->$2 M9" "kv_kan_bwd_input T66192 g1 $1->$2 M9" "kv_kan_bwd_weight T66192 g1 $1->$2 M9" > $O/r02_layer_$1x$2_summary.txt 2>&1
  cat $O/r02_layer_$1x$2_summary.txt | cut -c1-250
  rm -f $O/r02_layer_$1x$2.ncu-rep
done
# (3) inside one step: attention (tcgen05) forward / backward, the persistent q/k/v forward, LayerNorm, Adam
KV_FAST_INIT=1 KV_NCU_RANGE=1 timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on \
  -k regex:'attn5_fwd_kernel|attn5_bwd_kernel|kan_tc_fwd_persistent_kernel|layernorm_fwd_vec|layernorm_bwd_vec|adam_kernel' --launch-skip 10 -c 8 \
  -o $O/r02_cfg4_step_kernels -f python bench.py --no-cpu --steps 2 --warmup 3 > $O/r02_ncu_step.log 2>&1
ncu -i $O/r02_cfg4_step_kernels.ncu-rep --page raw --csv > $O/r02_cfg4_step_kernels_raw.csv 2>/dev/null
python scripts/ncu_summary.py $O/r02_cfg4_step_kernels.ncu-rep > $O/r02_cfg4_step_kernels_summary.txt 2>&1
cat $O/r02_cfg4_step_kernels_summary.txt | cut -c1-250
rm -f $O/r02_cfg4_step_kernels.ncu-rep
du -sh $O
