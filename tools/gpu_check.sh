#!/bin/bash
# GPU-box check: tcgen05 GEMM self-test (both descriptor conventions), parity tests, short bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version --format=csv,noheader > gpurun_out/gpu.txt 2>&1
timeout 120 tools/tc_selftest.bin 2 > gpurun_out/selftest_split2.log 2>&1; r0=$?
timeout 120 tools/tc_selftest.bin 3 > gpurun_out/selftest_split3.log 2>&1; r1=$?
echo "selftest split2 rc=$r0 split3 rc=$r1"; tail -4 gpurun_out/selftest_split3.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -2 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
