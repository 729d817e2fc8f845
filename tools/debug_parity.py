"""Diagnostic (not a test): per-variable gradient error and per-buffer activation error vs the oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import keras_semantics as K
from tests.util import pkg, syn, make_model, rel_err, grad_errors, oracle_to_named

def run(cfg, B, full, seed=1):
    if full:
        g, _ = syn.select_groups(cfg.G); groups = [g] * cfg.n_enc
    else:
        groups = [syn.small_groups(cfg.N, cfg.R, seed=1, overlap=True)] * cfg.n_enc
    P = K.init_params(cfg, groups, seed=3)
    rng = np.random.default_rng(11)
    for k in ("att_b1", "att_b2", "inter_b", "out_b"):
        P[k] = (P[k] + 0.05 * rng.standard_normal(P[k].shape)).astype(np.float32)
    (x, cap, a0, c0), tgt = syn.make_batch(B, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
    batch = (x, cap, a0, c0, tgt)
    model = make_model(cfg, groups, P, batch_max=B)
    met_o, grads_o, embsq, stats, out = K.loss_and_grads(cfg, P, groups, batch, seed=seed)
    met_d, grads_d = model.loss_and_gradients(((x, cap, a0, c0), tgt), seed=seed)
    print("oracle ", met_o, "emb_sq", embsq)
    print("device ", met_d)
    T, R, G, A, U, V = cfg.T, cfg.R, cfg.G, cfg.A, cfg.U, cfg.V
    feats = model.debug_buffer("feats")[:B * R * G].reshape(B, R, G)
    print("feats   err", rel_err(feats, out["feats"].detach().numpy()))
    alpha = model.debug_buffer("alpha")[:T * B * R].reshape(T, B, R)
    print("alpha   err", rel_err(alpha, out["attn"].detach().numpy()[..., 0]))
    errs = grad_errors(cfg, grads_o, grads_d)
    agg = {}
    for k, (e, n) in errs.items():
        key = k.split("/dense_")[0] + ("/dense_*/" + k.rsplit("/", 1)[1] if "/dense_" in k else "")
        a = agg.setdefault(key, [0.0, 0.0]); a[0] = max(a[0], e); a[1] = max(a[1], n)
    for k, (e, n) in agg.items():
        print(f"grad {k:<44s} relerr {e:.3e}  |g| {n:.3e}")

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "tiny"
    if which == "tiny":
        run(K.tiny_config().no_dropout(), 4, False)
        run(K.tiny_config(), 4, False, seed=7)
    elif which == "ms2":
        run(K.tiny_config(n_enc=2), 6, False, seed=9)
    else:
        run(K.Config().no_dropout(), 64, True)
        run(K.Config(), 64, True, seed=21)
