#!/bin/bash
# GPU-box round-end pass: smoke, parity tests, default bench (+ reference arm), ncu launch list, one `ncu --set full`
# capture of the HBM-bound and recurrence kernels of a steady-state step (side stream off so launches serialise).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log | cut -c1-200
timeout 800 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -3 gpurun_out/bench.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
TNT_SIDE_STREAM=0 timeout 500 ncu --set full --clock-control none --import-source on \
  -k 'regex:enc_fwd_t|enc_bwd_dw_t|enc_transpose|opt_adam|opt_norm|pd_fwd|pd_bwd|softmax_ce|att_w1_bwd' -s 44 -c 13 -f -o gpurun_out/prof_r1e \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/bench.log", "gpurun_out/bench_ref.log"):
    l = [x for x in open(f) if x.startswith("{")]
    if l:
        d = json.loads(l[-1]); print(f, {k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches", "phases_ms")},
                                     d["e2e"]["value"], (d.get("greedy") or {}).get("value"), (d.get("roofline") or {}).get("frac"))
PY
