#!/bin/bash
# GPU-box quick loop: parity tests, pd trace summary, short bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/pd_trace.py > gpurun_out/pd_trace.log 2>&1; grep -E "kernel loop|t= 7 " gpurun_out/pd_trace.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
l=[x for x in open('gpurun_out/bench.log') if x.startswith('{')]
if l:
    d=json.loads(l[-1]); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','phases_ms')}, d['e2e']['value'], d['greedy']['value'], d['roofline']['frac'])
PY
tail -3 gpurun_out/bench.err
