#!/bin/bash
# GPU-box: `ncu --set full` of the HBM-bound kernels of one steady-state train step (side stream off so launches serialise)
mkdir -p gpurun_out
export TNT_SIDE_STREAM=0
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on \
  -k 'regex:enc_fwd_t|enc_bwd_dw_t|enc_transpose|opt_adam|opt_norm|enc_finalize|bn_apply|bn_bwd|softmax_ce|att_w1_bwd' -s 48 -c 14 -f -o gpurun_out/prof_hbm \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_hbm.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_hbm.log | cut -c1-200; ls -la gpurun_out/*.ncu-rep
