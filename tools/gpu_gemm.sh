#!/bin/bash
mkdir -p gpurun_out
timeout 120 tools/tc_selftest.bin 3 > gpurun_out/selftest_split3.log 2>&1; echo "selftest3 rc=$?"; tail -18 gpurun_out/selftest_split3.log
timeout 120 tools/tc_selftest.bin 2 > gpurun_out/selftest_split2.log 2>&1; echo "selftest2 rc=$?"; tail -3 gpurun_out/selftest_split2.log
timeout 120 tools/tc_selftest.bin bench 3 > gpurun_out/gemm_bench.log 2>&1; echo "gemm bench rc=$?"
timeout 120 tools/tc_selftest.bin bench 2 >> gpurun_out/gemm_bench.log 2>&1
cat gpurun_out/gemm_bench.log
