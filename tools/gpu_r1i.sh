#!/bin/bash
# GPU-box pass r1i: smoke, parity tests, default bench (adds greedy_large), greedy batch sweep
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log | cut -c1-160
timeout 800 python -m pytest tests -m gpu -q --durations=3 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 300 python tools/greedy_bench.py 515 2048 4096 > gpurun_out/greedy_bench.log 2>&1; echo "greedy rc=$?"; cat gpurun_out/greedy_bench.log | cut -c1-200
python - <<'PY'
import json
for f in ("gpurun_out/bench.log",):
    l = [x for x in open(f) if x.startswith("{")]
    if l:
        d = json.loads(l[-1]); print(f, {k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches", "clocks")},
                                     d["e2e"]["value"], d.get("greedy"), d.get("greedy_large"), (d.get("device_dataset") or {}).get("value"), d["roofline"]["frac"])
PY
