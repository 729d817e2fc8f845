"""Diagnostic (GPU box): per-channel parcel bias gradients vs the oracle on the C1 dropout two-step case."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import keras_semantics as K
from tests.util import make_model, grad_errors
from tests.test_gpu_parity import _case

cfg = K.Config()
seed, B = 21, 64
groups, P, batch = _case(cfg, B, full_groups=True)
model = make_model(cfg, groups, P, batch_max=B)
opt = K.new_opt_state(P)
x, cap, a0, c0, tgt = batch
for s in range(2):
    named0 = {k: np.asarray(v, dtype=np.float64).copy() for k, v in model.get_named_weights().items() if k.startswith("enc0/dense_145/")}
    _, gd = model.loss_and_gradients(((x, cap, a0, c0), tgt), seed=seed + s)
    Y0 = model.debug_buffer("Y")[:B * cfg.R * cfg.G].reshape(B, cfg.R, cfg.G).astype(np.float64)[:, 145, :]
    dY0 = model.debug_buffer("dYpre")[:B * cfg.R * cfg.G].reshape(B, cfg.R, cfg.G).astype(np.float64)[:, 145, :]
    idx0 = np.asarray(groups[0][145], dtype=np.int64)
    pre0 = np.asarray(x, dtype=np.float64)[:, idx0] @ named0["enc0/dense_145/kernel"] + named0["enc0/dense_145/bias"]
    ref0 = np.where(pre0 > 0, pre0, 0.2 * pre0)
    print(f"step {s}: parcel 145 Y as used by this step vs fp64 of the weights it used: max err {np.abs(Y0 - ref0).max():.3e}, "
          f"min |pre| {np.abs(pre0).min():.3e}, sign mismatches {(np.sign(Y0) != np.sign(ref0)).sum()}")
    gb_dev = np.asarray(gd["enc0/dense_145/bias"], dtype=np.float64)
    print(f"   bias gradient from the device dYpre buffer (fp64 column sums) vs device bias gradient: {np.abs(dY0.sum(0) - gb_dev).max():.3e}")
    met_o, go = K.train_step(cfg, P, opt, groups, batch, seed=seed + s)
    model.train_step(((x, cap, a0, c0), tgt), seed=seed + s)
    errs = grad_errors(cfg, go, gd)
    from tests.util import oracle_to_named
    gb_or = np.asarray(oracle_to_named(cfg, {**go, "enc0_mean": np.zeros(cfg.G), "enc0_var": np.zeros(cfg.G)})["enc0/dense_145/bias"], dtype=np.float64)
    dd = gb_dev - gb_or
    print("   per-channel bias-gradient difference (device - oracle):", np.array2string(dd, precision=2, max_line_width=200))
    print("   device dYpre |.| per-channel max:", np.array2string(np.abs(dY0).max(0), precision=2, max_line_width=200))
    gch = int(np.argmax(np.abs(dd)))
    print(f"   channel {gch}: dYpre column", np.array2string(dY0[:, gch], precision=2, max_line_width=250))
    print(f"   channel {gch}: Y column", np.array2string(Y0[:, gch], precision=3, max_line_width=250))
    gmax = max(n for _, n in errs.values())
    bias = sorted(((k, v) for k, v in errs.items() if k.endswith("/bias") and k.startswith("enc")), key=lambda kv: -kv[1][0])
    print(f"step {s}: gmax {gmax:.3e}; worst parcel bias gradients (rel err, norm):")
    for k, v in bias[:6]:
        print(f"   {k:28s} {v[0]:.3e} {v[1]:.3e}")
    rel = np.array([v[0] for _, v in bias]); nrm = np.array([v[1] for _, v in bias])
    print(f"   median rel err {np.median(rel):.3e}; median norm {np.median(nrm):.3e}; abs err max {np.max(rel * nrm):.3e} median {np.median(rel * nrm):.3e}")
    kern = sorted(((k, v) for k, v in errs.items() if k.endswith("/kernel") and k.startswith("enc")), key=lambda kv: -kv[1][0])
    print("   worst parcel kernel gradients:", [(k, f"{v[0]:.2e}", f"{v[1]:.2e}") for k, v in kern[:3]])
    named = model.get_named_weights()
    dW = np.array([np.abs(np.asarray(named[f"enc0/dense_{r}/kernel"], dtype=np.float64) - P["enc0_W"][r]).max() for r in range(cfg.R)])
    db = np.array([np.abs(np.asarray(named[f"enc0/dense_{r}/bias"], dtype=np.float64) - P["enc0_b"][r]).max() for r in range(cfg.R)])
    o = np.argsort(-dW)[:4]
    print(f"after step {s}: per-parcel max |W_dev - W_oracle|: median {np.median(dW):.3e}, worst {[(int(r), f'{dW[r]:.2e}') for r in o]}; "
          f"bias: median {np.median(db):.3e} max {db.max():.3e} at {int(db.argmax())}")
    if s == 0:
        W145 = np.asarray(named["enc0/dense_145/kernel"], dtype=np.float64) - P["enc0_W"][145]
        rows = np.abs(W145).max(1)
        print(f"   parcel 145 rows with |diff| > 1e-6: {np.nonzero(rows > 1e-6)[0][:40]} (of {len(rows)})")
        continue
    # the outlier parcel: device weights vs oracle weights, device Y vs fp64 from the DEVICE weights (before this step's update
    # they were the weights the forward above used; after train_step they moved by ~lr, so compare with a fresh forward)
    r = int(bias[0][0].split("dense_")[1].split("/")[0]) if "dense_" in bias[0][0] else 0
    idx = np.asarray(groups[0][r], dtype=np.int64)
    named = model.get_named_weights()
    Wd = np.asarray(named[f"enc0/dense_{r}/kernel" if r else "enc0/dense/kernel"], dtype=np.float64)
    bd = np.asarray(named[f"enc0/dense_{r}/bias" if r else "enc0/dense/bias"], dtype=np.float64)
    print(f"parcel {r}: n_r {len(idx)}; after step {s}: max |W_dev - W_oracle| {np.abs(Wd - P['enc0_W'][r]).max():.3e}, "
          f"max |b_dev - b_oracle| {np.abs(bd - P['enc0_b'][r]).max():.3e}")
    model.loss_and_gradients(((x, cap, a0, c0), tgt), seed=seed + 2)
    Y = model.debug_buffer("Y")[:B * cfg.R * cfg.G].reshape(B, cfg.R, cfg.G).astype(np.float64)[:, r, :]
    pre = np.asarray(x, dtype=np.float64)[:, idx] @ Wd + bd
    ref = np.where(pre > 0, pre, 0.2 * pre)
    d = np.abs(Y - ref)
    bmax, gmax_ = np.unravel_index(np.argmax(d), d.shape)
    print(f"   fresh forward with the device weights: max |Y - ref| {d.max():.3e} at (b={bmax}, g={gmax_}), pre there {pre[bmax, gmax_]:.3e}, "
          f"min |pre| {np.abs(pre).min():.3e}; elements with error > 1e-5: {(d > 1e-5).sum()}")
    pre_o = np.asarray(x, dtype=np.float64)[:, idx] @ P["enc0_W"][r].astype(np.float64) + P["enc0_b"][r]
    print(f"   sign differences device-weights vs oracle-weights pre-activation: {(np.sign(pre) != np.sign(pre_o)).sum()}")
