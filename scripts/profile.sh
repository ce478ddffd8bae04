#!/usr/bin/env bash
# Profiling pass (run under gpurun, one GPU): plain bench first, then the ncu launch list of the same command, then one
# `--set full` capture of the dominant kernels.  Numbers printed under ncu are never bench values.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-r01}
ARGS="--steps 1 --warmup 3 --no-cpu-baseline --no-video --no-eager"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-video --no-eager > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err || { echo "bench failed"; tail -5 gpurun_out/bench_$TAG.err; exit 1; }
cat gpurun_out/bench_$TAG.json | cut -c1-600
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 550 -c 600 --csv \
  --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc2_kernel --launch-skip 1 --launch-count 4 \
  -o gpurun_out/gemm_$TAG -f python bench.py $ARGS > gpurun_out/ncu_gemm_$TAG.log 2>&1
echo "gemm full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:flash_attn_tc_kernel --launch-skip 1 --launch-count 1 \
  -o gpurun_out/attn_$TAG -f python bench.py $ARGS > gpurun_out/ncu_attn_$TAG.log 2>&1
echo "attn full rc=$?"
ls -la gpurun_out/
