#!/bin/bash
# N=2 box: environment variants of the data-parallel step (stream priority, hardware connections)
N=${1:-2}
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/dpe_$tag.log 2> gpurun_out/dpe_$tag.err
  echo "$tag rc=$? $(tail -1 gpurun_out/dpe_$tag.log | cut -c40-140)"
}
run base X=1
run conn32 CUDA_DEVICE_MAX_CONNECTIONS=32
run prio_eq TNT_SIDE_PRIO=eq
run prio_eq_conn32 TNT_SIDE_PRIO=eq CUDA_DEVICE_MAX_CONNECTIONS=32
run nodefer TNT_DP_DEFER=0 CUDA_DEVICE_MAX_CONNECTIONS=32
