#!/bin/bash
# One GPU-box pass: smoke, parity tests, default bench (+ optional extras named on the command line).
#   tools/gpu_pass.sh [tag] [extra ...]   extras: ref (reference arm), ncu (launch list), c4 (bench --workload C4)
# Every step runs under its own timeout so that a hung kernel cannot eat the box.
TAG=${1:-pass}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/${TAG}_gpu.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log | cut -c1-200
timeout 1500 python -m pytest tests -m gpu -q --durations=8 -x > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/${TAG}_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.log 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench.err
for x in "$@"; do
  case $x in
    ref) timeout 400 python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/${TAG}_bench_ref.log 2>&1; echo "ref rc=$?";;
    c4) timeout 400 python bench.py --workload C4 --no-cpu-baseline > gpurun_out/${TAG}_bench_c4.log 2> gpurun_out/${TAG}_bench_c4.err; echo "c4 rc=$?";;
    ncu) timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?";;
  esac
done
python - <<PY
import json
for f in ("gpurun_out/${TAG}_bench.log", "gpurun_out/${TAG}_bench_c4.log"):
    try:
        l = [x for x in open(f) if x.startswith("{")]
    except OSError:
        continue
    if l:
        d = json.loads(l[-1])
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"],
              "roofline", d["roofline"]["frac"], "greedy", (d.get("greedy") or {}).get("value"),
              (d.get("greedy_large") or {}).get("value"), "c4", (d.get("c4") or {}).get("value"))
        for k in d["kernels"][:24]:
            print(f"   {k['ms']*1e3:8.1f} us  {k['name'][:80]}")
PY
