#!/bin/bash
# GPU-box pass r1g: parity tests (incl. dataset / golden / regression), default bench, dataset bench, launch list,
# ncu --set full of the optimizer parts / multi-pack / colsum kernels of a steady-state step (default streams).
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 200 python tools/dataset_bench.py > gpurun_out/dataset_bench.log 2>&1; echo "dataset rc=$?"; tail -1 gpurun_out/dataset_bench.log
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on \
  -k 'regex:opt_adam|opt_norm|tc_pack_multi|colsum' -s 24 -c 8 -f -o gpurun_out/prof_r1g_step \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_step.log 2>&1; echo "ncu step rc=$?"
ls -la gpurun_out/*.ncu-rep
python - <<'PY'
import json
for f in ("gpurun_out/bench.log",):
    l = [x for x in open(f) if x.startswith("{")]
    if l:
        d = json.loads(l[-1]); print(f, {k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches", "phases_ms", "clocks")},
                                     d["e2e"]["value"], (d.get("greedy") or {}).get("value"), d.get("device_dataset"))
PY
