"""Diagnostic (GPU box): time tnt_phase_encode (encoder forward) in isolation."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import keras_semantics as K
from tests.util import syn, make_model
cfg = K.Config()
g, _ = syn.select_groups(cfg.G)
P = K.init_params(cfg, [g], seed=3)
(x, cap, a0, c0), tgt = syn.make_batch(64, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
model = make_model(cfg, [g], P, batch_max=64)
model.train_step(((x, cap, a0, c0), tgt), seed=1)
xb = [torch.randn(64, cfg.N, device="cuda") for _ in range(4)]
lib, ctx, st = model._libh, model._ctx, model._stream()
model.trace(True)
for i in range(12):
    lib.tnt_phase_encode(ctx, xb[i % 4].data_ptr(), 64, 64, 1, 1, st)
torch.cuda.synchronize()
agg = {}
for n, ms in model.trace_read():
    agg.setdefault(n, []).append(ms)
for n, v in agg.items():
    print(f"{os.environ.get('TNT_ENC_DBG','0')} {os.environ.get('TNT_ENC_MMA','1')} {n:28s} {1e3*np.median(v):8.1f} us")
