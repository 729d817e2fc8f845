#!/bin/bash
# N-GPU box: data-parallel schedule variants (communicators x optimizer order), short benches
N=${1:-8}
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/dpv_$tag.log 2> gpurun_out/dpv_$tag.err
  echo "$tag rc=$? $(tail -1 gpurun_out/dpv_$tag.log | cut -c1-130)"
}
run g2_enc TNT_DP_GROUPS=2 TNT_DP_ORDER=enc
run g2_dec TNT_DP_GROUPS=2 TNT_DP_ORDER=dec
run g1_dec TNT_DP_GROUPS=1 TNT_DP_ORDER=dec
run g1_enc TNT_DP_GROUPS=1 TNT_DP_ORDER=enc
