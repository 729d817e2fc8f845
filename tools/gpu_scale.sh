#!/bin/bash
# multi-GPU box: bench at N = 8, 4, 2, 1 (weak scaling, B=64 per GPU)
mkdir -p gpurun_out
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n$n.log 2> gpurun_out/bench_n$n.err; echo "bench n$n rc=$?"
  tail -1 gpurun_out/bench_n$n.log | cut -c1-160
done
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1.log 2> gpurun_out/bench_n1.err; echo "bench n1 rc=$?"; tail -1 gpurun_out/bench_n1.log | cut -c1-160
