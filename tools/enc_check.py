"""Diagnostic (GPU box): encoder forward Y against an fp64 restatement at the full C1 parcel table, several batches."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import keras_semantics as K
from tests.util import syn, make_model

cfg = K.Config().no_dropout()
g, _ = syn.select_groups(cfg.G)
P = K.init_params(cfg, [g], seed=3)
model = make_model(cfg, [g], P, batch_max=64)
B, R, G = 64, cfg.R, cfg.G
for seed in (5, 6, 7):
    (x, cap, a0, c0), tgt = syn.make_batch(B, cfg.T, cfg.V, cfg.U, cfg.N, seed=seed)
    model.loss_and_gradients(((x, cap, a0, c0), tgt), seed=seed)
    Y = model.debug_buffer("Y")[:B * R * G].reshape(B, R, G).astype(np.float64)
    x64 = np.asarray(x, dtype=np.float64)
    worst = (0.0, -1)
    errs = np.zeros(R)
    for r in range(R):
        idx = np.asarray(g[r], dtype=np.int64)
        pre = x64[:, idx] @ P["enc0_W"][r].astype(np.float64) + P["enc0_b"][r]
        ref = np.where(pre > 0, pre, 0.2 * pre)
        e = np.abs(Y[:, r, :] - ref).max()
        errs[r] = e
        if e > worst[0]:
            worst = (e, r, len(idx), float(np.abs(pre).max()))
    print(f"seed {seed}: max |Y - ref| = {errs.max():.3e} (parcel, n_r, max|pre|) {worst[1:]}; "
          f"median parcel error {np.median(errs):.3e}; parcels with error > 1e-5: {(errs > 1e-5).sum()}", flush=True)
    order = np.argsort(-errs)[:5]
    print("   worst parcels:", [(int(r), len(g[r]), f"{errs[r]:.2e}") for r in order])
