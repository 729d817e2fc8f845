"""Kernel timeline (CUPTI via torch.profiler) of one greedy decode of the 515-caption test split."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
pkg = importlib.import_module("masters-thesis_b200"); syn = pkg.synthetic
dev = torch.device("cuda", 0)
Bd, N, T, V, U = int(os.environ.get("GREEDY_B", "515")), syn.N_VERTICES, 15, 5001, 512
groups, out_groups = syn.select_groups(32)
m = pkg.NIC((groups, out_groups), U, 512, 512, 32, V, T, 0.0, 0.2, 0.2, 0.2, 0.2, 0.2, 0.01, 0.001, 0.00003, 0.00001,
            batch_max=Bd, seed=42, device=dev)
m.compile(pkg.Adam(1e-4, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1), None)
m.build(N, Bd)
x = torch.randn(Bd, N, device=dev); z = torch.zeros(Bd, U, device=dev)
start = torch.full((Bd,), 2, dtype=torch.int32, device=dev)
fn = lambda: m.greedy_predict(x, z, z, start, T, U, None, return_raw=False, return_attention=False)
for _ in range(3):
    fn()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    fn(); torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
agg = {}
for e in ev:
    d = e.time_range.end - e.time_range.start
    a = agg.setdefault(e.name[:60], [0, 0.0]); a[0] += 1; a[1] += d
print(f"span {ev[-1].time_range.end - t0:.1f} us, {len(ev)} device activities")
for k, (n, tot) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"{tot:9.1f} us  x{n:4d}  {k}")
print("--- first 70 activities")
for e in ev[:70]:
    print(f"{e.time_range.start - t0:8.1f} {e.time_range.end - e.time_range.start:8.1f}  {e.name[:70]}")
