#!/bin/bash
# N-GPU pass: data-parallel parity vs the oracle on the global batch (tiny + full), then bench at N (and N=1 beside it).
#   tools/gpu_dp.sh TAG N [bench-args...]
TAG=${1:-dp}; N=${2:-2}; shift; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/dp_check.py > gpurun_out/${TAG}_dp_check_tiny.log 2>&1; echo "dp_check tiny rc=$?"; grep -E "DP CHECK|worst|Error|error" gpurun_out/${TAG}_dp_check_tiny.log | head -5
timeout 400 $TR tools/dp_check.py full > gpurun_out/${TAG}_dp_check_full.log 2>&1; echo "dp_check full rc=$?"; grep -E "DP CHECK|worst|step" gpurun_out/${TAG}_dp_check_full.log | head -12
timeout 400 python bench.py --no-cpu-baseline "$@" > gpurun_out/${TAG}_bench_n1.log 2> gpurun_out/${TAG}_bench_n1.err; echo "bench n1 rc=$?"
timeout 400 $TR bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/${TAG}_bench_n$N.log 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench n$N rc=$?"; tail -3 gpurun_out/${TAG}_bench_n$N.err
python - <<PY
import json
r = {}
for n in (1, $N):
    try:
        l = [x for x in open(f"gpurun_out/${TAG}_bench_n{n}.log") if x.startswith("{")]
        d = json.loads(l[-1]); r[n] = d
        print(n, "value", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), d["config"].get("kernel_variants"), d.get("greedy_sharded"))
    except Exception as ex:
        print(n, "no line", ex)
if 1 in r and $N in r:
    print("efficiency", r[$N]["value"] / ($N * r[1]["value"]))
    for k in r[$N]["kernels"][:30]:
        print(f"   {k['ms']*1e3:8.1f} us  {k['name'][:90]}")
PY
