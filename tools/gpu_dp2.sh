#!/bin/bash
# 2-GPU box: data-parallel parity vs the oracle on the global batch, then the bench at N=2
mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py > gpurun_out/dp_check_tiny.log 2>&1; echo "dp tiny rc=$?"; tail -4 gpurun_out/dp_check_tiny.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/dp_check.py full > gpurun_out/dp_check_full.log 2>&1; echo "dp full rc=$?"; tail -12 gpurun_out/dp_check_full.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n2.log 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"
tail -1 gpurun_out/bench_n2.log | cut -c1-900; tail -3 gpurun_out/bench_n2.err
