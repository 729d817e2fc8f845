"""Diagnostic (GPU box): clock64 stamps inside the persistent decoder kernels (TNT_PD_TRACE=1)."""
import sys, os
os.environ["TNT_PD_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import keras_semantics as K
from tests.util import syn, make_model

cfg = K.Config() if len(sys.argv) < 2 else K.Config().no_dropout()
g, _ = syn.select_groups(cfg.G)
groups = [g]
P = K.init_params(cfg, groups, seed=3)
(x, cap, a0, c0), tgt = syn.make_batch(64, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
model = make_model(cfg, groups, P, batch_max=64)
for s in range(3):
    model.train_step(((x, cap, a0, c0), tgt), seed=s)
tr = model.debug_buffer("pd_trace", "i8").reshape(4, 64, 16)[:, :cfg.T]
names_f = ["top", "H ok", "h staged", "q done", "scores", "half ctx out", "h-gemm + A ok", "merged", "ctx-gemm", "cell"]
names_b = ["top", "dz", "d ctx tile", "h-gemm + B1 ok", "att bwd", "B2 ok", "dh gather"]
for kind, base, names in (("fwd", 0, names_f), ("bwd", 2, names_b)):
    for who in (0, 1):
        a = tr[base + who]
        t0 = a[:, 0].min()
        print(f"{kind} CTA {'j=0 (attention)' if who == 0 else 'j=31 (second half-attention)'}: kernel loop {a.max() - t0} clocks")
        order = range(cfg.T) if kind == "fwd" else range(cfg.T - 1, -1, -1)
        prev_top = None
        for t in list(order)[:15]:
            row = a[t]
            d = [int(row[i] - row[0]) if row[i] else -1 for i in range(len(names))]
            step = "" if prev_top is None else f" (+{int(row[0] - prev_top)} since prev top)"
            prev_top = row[0]
            print(f"  t={t:2d} " + " ".join(f"{n}:{v}" for n, v in zip(names[1:], d[1:])) + step)
