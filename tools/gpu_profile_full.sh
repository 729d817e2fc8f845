#!/bin/bash
# GPU-box: GEMM shape timings + one `ncu --set full` capture covering one train step's main kernels.
mkdir -p gpurun_out
timeout 120 tools/tc_selftest.bin bench 3 > gpurun_out/gemm_bench.log 2>&1; echo "gemm bench rc=$?"
timeout 120 tools/tc_selftest.bin bench 2 >> gpurun_out/gemm_bench.log 2>&1
cat gpurun_out/gemm_bench.log
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k 'regex:pd_fwd|pd_bwd|enc_fwd_kernel|enc_bwd_dw|opt_adam|opt_norm|gemm_tc' -s 40 -c 18 -f -o gpurun_out/prof_step \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/
