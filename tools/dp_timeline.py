"""torchrun --nproc-per-node N tools/dp_timeline.py : kernel timeline (CUPTI via torch.profiler) of one
steady-state data-parallel train step on rank 0: name, stream, start and duration in microseconds relative to
the step's first kernel.  Shows where the NCCL kernels sit relative to the compute they should hide behind."""
import datetime, importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from torch.profiler import profile, ProfilerActivity

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=90))
pkg = importlib.import_module("masters-thesis_b200"); syn = pkg.synthetic
B, N, T, V, U = 64, syn.N_VERTICES, 15, 5001, 512
groups, out_groups = syn.select_groups(32)
m = pkg.NIC((groups, out_groups), U, 512, 512, 32, V, T, 0.0, 0.2, 0.2, 0.2, 0.2, 0.2, 0.01, 0.001, 0.00003, 0.00001,
            batch_max=B, seed=42, device=dev)
m.compile(pkg.Adam(1e-4, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1), None)
m.build(N, B)
gen = torch.Generator(device=dev); gen.manual_seed(42 + rank)
xs = [torch.randn(B, N, device=dev, generator=gen) for _ in range(2)]
caps = syn.make_captions(2 * B, T, V, seed=42 + rank).reshape(2, B, T)
cap_d = [torch.from_numpy(c).to(dev) for c in caps]
tgt_d = [torch.from_numpy(syn.make_target_ids(c)).to(dev) for c in caps]
z = torch.zeros(B, U, device=dev)
def step(i):
    j = i % 2
    m._run_step(((xs[j], cap_d[j], z, z), tgt_d[j]), train=True, training=True, read_metrics=False)
for i in range(8):
    step(i)
torch.cuda.synchronize(dev)
if world > 1:
    dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for i in range(4):
        step(i)
    torch.cuda.synchronize(dev)
if rank == 0:
    import tempfile
    path = os.path.join(tempfile.gettempdir(), "tnt_dp_trace.json")
    prof.export_chrome_trace(path)
    tr = json.load(open(path))["traceEvents"]
    ev = [e for e in tr if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e]
    ev.sort(key=lambda e: e["ts"])
    # steady-state step = between the 2nd and 3rd enc_permute_kernel
    starts = [e["ts"] for e in ev if "enc_permute" in e["name"]]
    t0, t1 = starts[2], starts[3]
    print(f"step period {t1 - t0:.1f} us   (columns: start, duration, stream, kernel)")
    main = None
    for e in ev:
        if t0 <= e["ts"] < t1:
            st = e.get("args", {}).get("stream")
            if main is None:
                main = st
            print(f"{e['ts'] - t0:8.1f} {e.get('dur', 0):8.1f}  {'main' if st == main else 's' + str(st):>5s}  {e['name'][:72]}")
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
