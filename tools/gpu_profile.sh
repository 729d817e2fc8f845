#!/bin/bash
# GPU-box profiling pass: ncu launch list of a short bench run (per-kernel durations; cold-cache, serialised).
mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pre.log 2> gpurun_out/bench_pre.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu.log
