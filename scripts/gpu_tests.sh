#!/usr/bin/env bash
# Runs the GPU test files one process per group (a faulting kernel poisons its CUDA context), logs to gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
run() { # name, timeout, pytest args...
  local name=$1 to=$2; shift 2
  timeout "$to" python -m pytest "$@" -q -s -m gpu -p no:cacheprovider > "gpurun_out/$name.log" 2>&1
  echo "== $name rc=$? $(tail -1 gpurun_out/$name.log)"
}
run kernels 600 tests/test_gpu_kernels.py
run tc_gemm 300 tests/test_gpu_tc.py::test_gemm_tc_bias
run tc_epi 300 tests/test_gpu_tc.py::test_gemm_tc_epilogues
run tc_conv 300 tests/test_gpu_tc.py::test_conv3x3_tc tests/test_gpu_tc.py::test_output_head_tc_tf32
run tc_head 300 tests/test_gpu_tc.py::test_output_head_fused tests/test_gpu_tc.py::test_output_head_fused_f16
run tc_attn 300 tests/test_gpu_tc.py::test_flash_attention_tc
run fwd_fp32 900 tests/test_gpu_forward.py -k "(fp32 or taps or contract) and not driver"
run fwd_bf16 900 tests/test_gpu_forward.py -k "bf16 and not driver"
run driver 600 tests/test_gpu_forward.py -k "driver"
run video_io 300 tests/test_video_io.py
run fwd_misc 600 tests/test_gpu_forward.py -k "not fp32 and not taps and not contract and not bf16 and not driver"
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "== smoke rc=$? $(tail -1 gpurun_out/smoke.log)"
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "== bench rc=$? $(tail -c 1500 gpurun_out/bench.log)"
