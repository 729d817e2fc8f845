#!/bin/bash
# N-GPU pass for the encoder-gradient exchange: data-parallel parity vs the oracle, then bench with the peer-memory
# push exchange and (A/B) with the NCCL all-reduce.   tools/gpu_dp_ab.sh TAG N [bench-args...]
TAG=${1:-dpab}; N=${2:-2}; shift; shift
mkdir -p gpurun_out
export TNT_BENCH_WATCHDOG=150
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"
timeout 200 $TR tools/dp_check.py > gpurun_out/${TAG}_dp_check_tiny.log 2>&1; echo "dp_check tiny rc=$?"; grep -E "DP CHECK|worst|Error|error" gpurun_out/${TAG}_dp_check_tiny.log | head -5
timeout 300 $TR tools/dp_check.py full > gpurun_out/${TAG}_dp_check_full.log 2>&1; echo "dp_check full rc=$?"; grep -E "DP CHECK|worst" gpurun_out/${TAG}_dp_check_full.log | head -4
timeout 200 $TR bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/${TAG}_bench_push_n$N.log 2> gpurun_out/${TAG}_bench_push_n$N.err; echo "bench push rc=$?"; tail -2 gpurun_out/${TAG}_bench_push_n$N.err
[ -n "$SKIP_NCCL" ] || TNT_DP_NCCL_GRADS=1 timeout 200 $TR bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/${TAG}_bench_nccl_n$N.log 2> gpurun_out/${TAG}_bench_nccl_n$N.err; echo "bench nccl rc=$?"
for k in push $([ -n "$SKIP_NCCL" ] || echo nccl); do python tools/show_bench.py gpurun_out/${TAG}_bench_${k}_n$N.log | grep -E "value|dp|enc_bwd|opt_" | cut -c1-200; done
