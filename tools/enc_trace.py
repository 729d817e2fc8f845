"""Diagnostic (GPU box): per-tile time line (clock64 of CTA 0) of the two tcgen05 encoder kernels."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import keras_semantics as K
from tests.util import syn, make_model

cfg = K.Config().no_dropout()
g, _ = syn.select_groups(cfg.G)
P = K.init_params(cfg, [g], seed=3)
model = make_model(cfg, [g], P, batch_max=64)
(x, cap, a0, c0), tgt = syn.make_batch(64, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
model.loss_and_gradients(((x, cap, a0, c0), tgt), seed=1)
MAXT = 512
buf = torch.zeros(2 * 8 * MAXT, dtype=torch.int64, device="cuda")
lib = model._libh


lib.tnt_debug_encoder_trace(model._ctx, buf.data_ptr())
model.loss_and_gradients(((x, cap, a0, c0), tgt), seed=1)
torch.cuda.synchronize()
lib.tnt_debug_encoder_trace(model._ctx, None)
t = buf.cpu().numpy().reshape(2, MAXT, 8)
names = ["load issued", "poll begin", "tile ready", "conv done", "mma go", "mma committed", "acc seen", "drained"]
for k, kern in enumerate(("forward", "gradient")):
    a = t[k]
    n = int((a[:, 0] != 0).sum())
    t0 = a[0, 0]
    print(f"== {kern}: {n} tiles on CTA 0, {int(a[:n].max() - t0)} cycles from first load to last event")
    print("   tile " + " ".join(f"{s:>13s}" for s in names))
    for j in range(min(n, 24)):
        print(f"   {j:4d} " + " ".join(f"{int(v - t0):13d}" if v else f"{'-':>13s}" for v in a[j]))
    if n > 2:
        d = np.diff(a[:n, 5].astype(np.int64))
        print(f"   mma-commit to mma-commit: median {np.median(d):.0f} cycles; conv (ready->done) median "
              f"{np.median(a[:n, 3] - a[:n, 2]):.0f}; wait for tile (poll begin->ready) median {np.median(a[:n, 2] - a[:n, 1]):.0f}; "
              f"conv done -> mma go median {np.median(a[:n, 4] - a[:n, 3]):.0f}; mma issue median {np.median(a[:n, 5] - a[:n, 4]):.0f}; "
              f"commit -> acc seen median {np.median(a[:n, 6] - a[:n, 5]):.0f}")
