#!/bin/bash
# N-GPU bench with a watchdog: prints where each rank is when it stalls.  tools/gpu_dp_debug.sh TAG N [bench args]
TAG=${1:-dpd}; N=${2:-2}; shift; shift
mkdir -p gpurun_out
export TNT_BENCH_WATCHDOG=170 NCCL_DEBUG=WARN
timeout 260 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/${TAG}_bench_n$N.log 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench n$N rc=$?"
grep -E "^\[bench|File \"/root|File \".*bench.py|line [0-9]+ in|Thread|Error|error" gpurun_out/${TAG}_bench_n$N.err | tail -60
python tools/show_bench.py gpurun_out/${TAG}_bench_n$N.log
