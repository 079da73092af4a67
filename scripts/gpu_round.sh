#!/bin/bash
# One GPU session: parity tests, bench lines, ncu launch list and full captures of the three KAN GEMM kernels.
# usage (under gpurun): bash scripts/gpu_round.sh <tag>
tag=${1:-r01}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py > gpurun_out/${tag}_bench_cfg4.json 2> gpurun_out/${tag}_bench_cfg4.err; tail -c 600 gpurun_out/${tag}_bench_cfg4.json
timeout 300 python bench.py --batch 128 --no-cpu --steps 5 > gpurun_out/${tag}_bench_cfg4_b128.json 2>&1; cut -c1-200 gpurun_out/${tag}_bench_cfg4_b128.json
if [ "$2" != "noncu" ]; then
# launch list of one step (a number printed under ncu is never a bench value)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 9000 -c 2400 --csv \
  --log-file gpurun_out/${tag}_cfg4_launches.csv python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'kan_tc_fwd_kernel|effkan8_bwd_input_kernel|kan_tc_bwd_weight_kernel' \
  --launch-skip 9 -c 3 -o gpurun_out/${tag}_kan_layer -f python scripts/bench_kan_layer.py --what fwdbwd --T 12608 --iters 3 > gpurun_out/${tag}_ncu_full.log 2>&1
tail -2 gpurun_out/${tag}_ncu_full.log
fi
