#!/usr/bin/env python
"""GPU-box: batch assembly from an HBM-resident beta table (tnt_assemble_betas / tnt_assemble_captions) at the
reference's sizes: B=64 rows of 327 684 floats out of a 1024-row table (1.34 GB), ids and one-hot targets.
Prints one JSON line (ms per batch, algorithmic GB/s against MEASURED_PEAKS.json)."""
import importlib, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("masters-thesis_b200")
syn = pkg.synthetic
dev = torch.device("cuda", 0)
B, N, T, V, U = 64, syn.N_VERTICES, 15, 5001, 512
n_keys, n_pairs = 1024, 64 * 64
g = torch.Generator(device=dev); g.manual_seed(0)
table = torch.randn(n_keys, N, device=dev, generator=g)
rs = np.random.RandomState(1)
rows = rs.randint(0, n_keys, n_pairs)
caps = list(syn.make_captions(n_pairs, T, V, seed=2))
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
out = {}
for onehot in (False, True):
    d = pkg.DeviceDataGenerator(table, rows, caps, B, U, T, V, shuffle=True, training=True, onehot=onehot, rng=rs)
    for i in range(4):
        d[i]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 40
    e0.record()
    for i in range(reps):
        d[4 + i]
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    nbytes = 2.0 * B * N * 4 + 3.0 * B * T * 4 + (B * T * V * 4.0 if onehot else 0.0)
    out["onehot" if onehot else "ids"] = {"ms_per_batch": ms, "alg_bytes": nbytes, "GBps": nbytes / ms / 1e6,
                                          "frac_of_hbm_peak": nbytes / ms / 1e6 / peak}
    d.check()
print(json.dumps({"workload": f"B={B} x N={N} fp32 rows from a {n_keys}-row table", "peak_GBps": peak, **out}))
