#!/bin/bash
# GPU-box pass r1f: parity tests (incl. dataset), bench with/without the multi-operand pack, dataset bench,
# ncu --set full of the optimizer parts / multi-pack / colsum kernels of a steady-state step and of the dataset kernels.
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
TNT_MULTIPACK=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_nomp.log 2> gpurun_out/bench_nomp.err; echo "bench nomp rc=$?"
timeout 200 python tools/dataset_bench.py > gpurun_out/dataset_bench.log 2>&1; echo "dataset rc=$?"; tail -1 gpurun_out/dataset_bench.log
timeout 400 ncu --set full --clock-control none --import-source on \
  -k 'regex:opt_adam|opt_norm|tc_pack_multi|colsum_kernel' -s 36 -c 12 -f -o gpurun_out/prof_r1f_step \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_step.log 2>&1; echo "ncu step rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k 'regex:ds_' -s 8 -c 4 -f -o gpurun_out/prof_r1f_dataset \
  python tools/dataset_bench.py > gpurun_out/ncu_dataset.log 2>&1; echo "ncu dataset rc=$?"
ls -la gpurun_out/*.ncu-rep
python - <<'PY'
import json
for f in ("gpurun_out/bench.log", "gpurun_out/bench_nomp.log"):
    l = [x for x in open(f) if x.startswith("{")]
    if l:
        d = json.loads(l[-1]); print(f, {k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches", "phases_ms")},
                                     d["e2e"]["value"], (d.get("greedy") or {}).get("value"), d.get("device_dataset"))
PY
