#!/bin/bash
# 2-GPU box: N=2 bench A/B (clock poller, multi-pack) to separate this session's changes from box-to-box variance
mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 200)) \
    bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n2_$name.log 2> gpurun_out/bench_n2_$name.err
  echo "$name rc=$?"
  python - "$name" <<'PY'
import json, sys
l = [x for x in open(f"gpurun_out/bench_n2_{sys.argv[1]}.log") if x.startswith("{")]
if l:
    d = json.loads(l[-1]); print(sys.argv[1], "ms_per_step", round(d["ms_per_step"], 4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
PY
}
run a X=1
run nonvml TNT_BENCH_NVML=0
run nomp TNT_MULTIPACK=0 TNT_BENCH_NVML=0
run a2 X=1
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; head -6 gpurun_out/topo.txt
