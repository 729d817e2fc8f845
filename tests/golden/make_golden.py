#!/usr/bin/env python
"""Generates tests/golden/*.json: known-answer vectors of the hot path on small seeded cases.

Where they come from: the reference is TensorFlow/Keras scripts and TensorFlow is not installable in this image
(no network), so these are NOT outputs of the reference itself -- they are outputs of the CPU oracle
(oracle/keras_semantics.py in float64, oracle/data_generator.py), frozen here so that (a) the oracle cannot drift
unnoticed (tests/test_golden.py, CPU) and (b) the CUDA path is checked against committed numbers as well as against
the live oracle (tests/test_gpu_zgolden.py).  Parity with TF itself stays unpinned.

    python tests/golden/make_golden.py        # rewrites the fixtures; inputs are regenerated from seeds by the tests
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import keras_semantics as K          # noqa: E402
from oracle import data_generator as O           # noqa: E402
from tests.util import syn                       # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def case(variant):
    """variant: 'c1' lc_NIC, 'c1_dropout', 'c2' ThinkAndTell loss/attention variant, 'c4' two subjects."""
    kw = {}
    if variant == "c2":
        kw = dict(masked_sparse_ce=True, plain_attention=True)
    if variant == "c4":
        kw = dict(n_enc=2)
    cfg = K.tiny_config(**kw)
    if variant != "c1_dropout":
        cfg = cfg.no_dropout()
    groups = [syn.small_groups(cfg.N, cfg.R, seed=1, overlap=True)] * cfg.n_enc   # ms2: one parcellation, two encoders
    P = K.init_params(cfg, groups, seed=3)
    rng = np.random.default_rng(11)
    for k in ("att_b1", "att_b2", "inter_b", "out_b"):          # keep h0 = 0 off the LeakyReLU kink
        P[k] = (P[k] + 0.05 * rng.standard_normal(P[k].shape)).astype(np.float32)
    B = 4
    (x, cap, a0, c0), tgt = syn.make_batch(B, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
    return cfg, groups, P, (x, cap, a0, c0, tgt)


def run(variant, steps=3, seed=7):
    cfg, groups, P, batch = case(variant)
    out = {"variant": variant, "seed": seed, "steps": []}
    _, grads, _, _, _ = K.loss_and_grads(cfg, P, groups, batch, seed=seed, dtype=torch.float64)
    out["grad_norms_step0"] = {}
    for n in K.trainable_names(cfg):
        g = grads[n]
        v = np.concatenate([np.asarray(a, np.float64).reshape(-1) for a in g]) if isinstance(g, list) else np.asarray(g, np.float64)
        out["grad_norms_step0"][n] = float(np.linalg.norm(v))
    opt = K.new_opt_state(P)
    for s in range(steps):
        met, _ = K.train_step(cfg, P, opt, groups, batch, seed=seed + s, dtype=torch.float64)
        out["steps"].append({k: float(v) for k, v in met.items()})
    out["weights_after"] = {n: float(np.sum(np.concatenate([np.asarray(a, np.float64).reshape(-1) for a in P[n]])
                                            if isinstance(P[n], list) else np.asarray(P[n], np.float64)))
                            for n in K.trainable_names(cfg)}
    if cfg.n_enc == 1:
        x, cap, a0, c0, _ = batch
        start = np.full(x.shape[0], syn.START_ID, np.int32)
        ids = K.greedy_predict(cfg, P, groups, x, a0, c0, start, cfg.T, dtype=torch.float64)[0]
        out["greedy_ids_after"] = np.asarray(ids).reshape(x.shape[0], cfg.T).tolist()
    return out


def full_case(B=16):
    """SURVEY 8(c) known-answer item (2): the FULL C1 configuration (327 684 betas, 345 parcels, LSTM 512, vocab 5001,
    T = 15) with seeded weights; B = 16 keeps the float64 oracle to a few seconds per step."""
    cfg = K.Config().no_dropout()
    g, _ = syn.select_groups(cfg.G)
    groups = [g]
    P = K.init_params(cfg, groups, seed=3)
    rng = np.random.default_rng(11)
    for k in ("att_b1", "att_b2", "inter_b", "out_b"):
        P[k] = (P[k] + 0.05 * rng.standard_normal(P[k].shape)).astype(np.float32)
    (x, cap, a0, c0), tgt = syn.make_batch(B, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
    return cfg, groups, P, (x, cap, a0, c0, tgt)


def run_full(steps=3, seed=7):
    """Metrics of three Adam steps, gradient norms per variable class at step 0 (the 345 parcel kernels / biases are
    summed per class), weight sums afterwards, greedy ids with the relative top-2 margin of every position (a margin
    below 1e-5 is a near-tie: the CUDA path may legitimately pick the other word there and diverge afterwards)."""
    cfg, groups, P, batch = full_case()
    out = {"variant": "full_c1", "seed": seed, "steps": []}
    _, grads, _, _, _ = K.loss_and_grads(cfg, P, groups, batch, seed=seed, dtype=torch.float64)
    out["grad_norms_step0"] = {}
    for n in K.trainable_names(cfg):
        g = grads[n]
        v = np.concatenate([np.asarray(a, np.float64).reshape(-1) for a in g]) if isinstance(g, list) else np.asarray(g, np.float64)
        out["grad_norms_step0"][n] = float(np.linalg.norm(v))
    opt = K.new_opt_state(P)
    for s in range(steps):
        met, _ = K.train_step(cfg, P, opt, groups, batch, seed=seed + s, dtype=torch.float64)
        out["steps"].append({k: float(v) for k, v in met.items()})
    out["weights_after"] = {n: float(np.sum(np.concatenate([np.asarray(a, np.float64).reshape(-1) for a in P[n]])
                                            if isinstance(P[n], list) else np.asarray(P[n], np.float64)))
                            for n in K.trainable_names(cfg)}
    x, cap, a0, c0, _ = batch
    start = np.full(x.shape[0], syn.START_ID, np.int32)
    ids, raw = K.greedy_predict(cfg, P, groups, x, a0, c0, start, cfg.T, dtype=torch.float64)[:2]
    top2 = np.sort(raw, axis=-1)[..., -2:]
    out["greedy_ids_after"] = np.asarray(ids).reshape(x.shape[0], cfg.T).tolist()
    out["greedy_rel_margin"] = ((top2[..., 1] - top2[..., 0]) / top2[..., 1]).tolist()
    return out


def generator_case():
    rng = np.random.default_rng(21)
    betas = rng.standard_normal((9, 40), dtype=np.float32)
    rows = rng.integers(0, 9, 23)
    caps = [[2] + list(map(int, rng.integers(4, 31, rng.integers(0, 9)))) + [3] for _ in range(23)]
    np.random.seed(77)
    g = O.DataGenerator(betas, rows, caps, 4, 6, 7, 31, shuffle=True, training=True)
    out = {"order_epoch0": g.order.tolist(), "batches": []}
    for i in (0, len(g) - 1):
        (x, cap, a0, c0), tgt = g[i]
        out["batches"].append({"index": i, "beta_rows": rows[g.order[i * 4:(i + 1) * 4]].tolist(), "cap": cap.tolist(),
                               "target_ids": tgt.argmax(-1).tolist(), "betas_sum": float(x.astype(np.float64).sum())})
    g.on_epoch_end()
    out["order_epoch1"] = g.order.tolist()
    return out


if __name__ == "__main__":
    for v in ("c1", "c1_dropout", "c2", "c4"):
        with open(os.path.join(HERE, f"tiny_{v}.json"), "w") as f:
            json.dump(run(v), f, indent=1)
    with open(os.path.join(HERE, "full_c1.json"), "w") as f:
        json.dump(run_full(), f, indent=1)
    with open(os.path.join(HERE, "generator.json"), "w") as f:
        json.dump(generator_case(), f, indent=1)
    print("golden fixtures written to", HERE)
