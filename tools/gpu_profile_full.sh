#!/bin/bash
# GPU-box: one `ncu --set full` capture covering one train step's main kernels (side stream off so launches serialise)
mkdir -p gpurun_out
export TNT_SIDE_STREAM=0
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on \
  -k 'regex:pd_fwd|pd_bwd|enc_fwd_pipe|enc_bwd_dw_pipe|opt_adam|opt_norm|gemm_pk|tc_pack|att_w1_bwd|softmax_ce' -s 70 -c 36 -f -o gpurun_out/prof_step2 \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_full.log | cut -c1-200; ls -la gpurun_out/*.ncu-rep
