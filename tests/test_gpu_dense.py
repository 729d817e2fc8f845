"""GPU parity of the NIC.py dense voxel encoder (SURVEY 8a row A1'') against oracle/dense_encoder.py, through the
Python mirror / C ABI: features, tape gradients of the four variables, weights and moving statistics after Adam."""
import numpy as np
import pytest
import torch

from oracle import dense_encoder as D
from tests.util import pkg, rel_err

pytestmark = pytest.mark.gpu

NAMES = {"W": "dense_img/kernel", "b": "dense_img/bias", "gamma": "batch_norm/gamma", "beta": "batch_norm/beta"}


def _run(K, E, B, dropout_input, dropout, steps, seed):
    rng = np.random.default_rng(seed)
    P = D.init_params(K, E, seed=seed)
    P["b"] = (0.05 * rng.standard_normal(E)).astype(np.float32)
    P["gamma"] = (1.0 + 0.1 * rng.standard_normal(E)).astype(np.float32)
    P["beta"] = (0.1 * rng.standard_normal(E)).astype(np.float32)
    enc = pkg.DenseEncoder(K, E, dropout_input, dropout, 0.01, batch_max=64)
    enc.compile(pkg.Adam(1e-4, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1))
    enc.set_weights([P["W"], P["b"], P["gamma"], P["beta"], P["mean"], P["var"]])
    opt = D.new_opt_state(P)
    for s in range(steps):
        x = rng.standard_normal((B, K)).astype(np.float32)
        df = (rng.standard_normal((B, E)) / B).astype(np.float32)
        f_o, g_o, st = D.loss_and_grads(P, x, df, seed=seed + s, dropout_input=dropout_input, dropout=dropout,
                                        input_reg=0.0)
        f_d = enc(x, training=True, seed=seed + s).cpu().numpy()
        assert f_d.shape == (B, 1, E)
        assert rel_err(f_d, f_o) < 1e-4, ("features", s, rel_err(f_d, f_o))
        enc.backward(df)
        g_d = enc.gradients()
        for k, n in NAMES.items():
            assert rel_err(g_d[n], g_o[k]) < 1e-3, (n, s, rel_err(g_d[n], g_o[k]))
        enc.apply_gradients()
        g_l2 = dict(g_o); g_l2["W"] = g_o["W"] + np.float32(2 * 0.01) * P["W"]       # the tape adds the L2 gradient
        D.adam_apply(P, opt, g_l2, lr=1e-4, beta1=0.9, beta2=0.98, eps=10.0e-9, clipnorm=0.1)
        D.update_moving_stats(P, st)
        got = enc.get_weights()
        for a, k in zip(got, ("W", "b", "gamma", "beta", "mean", "var")):
            assert np.abs(a - P[k]).max() <= 1e-5, (k, s, np.abs(a - P[k]).max())
    # inference: moving statistics, no dropout
    x = rng.standard_normal((B, K)).astype(np.float32)
    f_o = D.forward({k: torch.as_tensor(v) for k, v in P.items()}, torch.as_tensor(x), False).numpy()
    assert rel_err(enc(x, training=False).cpu().numpy(), f_o) < 1e-4
    return enc


def test_dense_encoder_tiny_ragged():
    _run(K=333, E=128, B=5, dropout_input=0.1, dropout=0.2, steps=3, seed=3)


def test_dense_encoder_visual_cortex_size():
    _run(K=62756, E=512, B=64, dropout_input=0.0, dropout=0.2, steps=2, seed=5)     # NIC.py:27 "(62756,)"


def test_dense_encoder_whole_brain_size():
    enc = _run(K=327684, E=512, B=64, dropout_input=0.0, dropout=0.2, steps=1, seed=7)
    assert enc.launch_count() > 0
