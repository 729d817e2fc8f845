#!/bin/bash
# GPU-box pass r1h: parity tests, default bench, launch list of the final step (deferred decoder update, merged packs)
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q --durations=3 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
TNT_OPT_DEFER=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_nodefer.log 2> gpurun_out/bench_nodefer.err; echo "bench nodefer rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/bench.log", "gpurun_out/bench_nodefer.log"):
    l = [x for x in open(f) if x.startswith("{")]
    if l:
        d = json.loads(l[-1]); print(f, {k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches", "clocks")},
                                     d["e2e"]["value"], (d.get("greedy") or {}).get("value"), (d.get("device_dataset") or {}).get("value"), d["roofline"]["frac"], d["roofline"]["traffic"])
PY
