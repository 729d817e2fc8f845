#!/bin/bash
# 4-GPU box: weak-scaling points N=4, 2, 1 of the final tree, back to back on one box
mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
for n in 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
    bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n$n.log 2> gpurun_out/scale_n$n.err; echo "n$n rc=$?"
done
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n1.log 2> gpurun_out/scale_n1.err; echo "n1 rc=$?"
python - <<'PY'
import json
for n in (4, 2, 1):
    l = [x for x in open(f"gpurun_out/scale_n{n}.log") if x.startswith("{")]
    if l:
        d = json.loads(l[-1]); print("N", n, "ms_per_step", round(d["ms_per_step"], 4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["e2e"].get("numa_bound_cpus"), d["clocks"]["samples"])
PY
