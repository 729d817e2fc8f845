"""Keras-2.x initialisers used by the reference model, by variable name of the flat layout.

  he_normal        LocallyDense / attention W1, W2          lc_NIC.py:87,99
  glorot_uniform   attention V (Dense default), lstm kernel  attention.py:23, lc_NIC.py:118
  orthogonal       lstm recurrent kernel (Keras default)
  zeros + forget 1 lstm bias (unit_forget_bias=True)
  RandomUniform    emb_text (-0.08, 0.08)                    lc_NIC.py:108
  GlorotNormal     time_distributed_{nonlinear,softmax}      lc_NIC.py:144,155
  BatchNorm        gamma 1, beta 0, moving_mean 0, moving_variance 1
"""
import math
import numpy as np

_TRUNC = 0.87962566103423978


def _trunc_normal(rng, shape, std):
    x = rng.standard_normal(shape)
    bad = np.abs(x) > 2.0
    while bad.any():
        x[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(x) > 2.0
    return (x * std).astype(np.float32)


def init_variable(rng, name, rows, cols, units):
    """Initial value (numpy fp32, shape (rows,cols) or (rows,)) of the variable called `name`."""
    leaf = name.split("/")[-1]
    shape = (rows, cols) if cols > 0 else (rows,)
    if leaf in ("bias", "beta", "moving_mean"):
        v = np.zeros(shape, np.float32)
        if name == "lstm/bias":
            v[units:2 * units] = 1.0
        return v
    if leaf in ("gamma", "moving_variance"):
        return np.ones(shape, np.float32)
    if name.startswith("enc") or name in ("attention/W1/kernel", "attention/W2/kernel"):
        return _trunc_normal(rng, shape, math.sqrt(2.0 / rows) / _TRUNC)              # he_normal
    if name in ("attention/V/kernel", "lstm/kernel"):
        lim = math.sqrt(6.0 / (rows + cols))
        return rng.uniform(-lim, lim, shape).astype(np.float32)                        # glorot_uniform
    if name == "lstm/recurrent_kernel":
        a = rng.standard_normal((max(rows, cols), min(rows, cols)))
        q, r = np.linalg.qr(a)
        q = q * np.sign(np.diag(r))
        if rows < cols:
            q = q.T
        return q.reshape(shape).astype(np.float32)                                     # orthogonal
    if name == "emb_text/embeddings":
        return rng.uniform(-0.08, 0.08, shape).astype(np.float32)
    if name.startswith("time_distributed"):
        return _trunc_normal(rng, shape, math.sqrt(2.0 / (rows + cols)) / _TRUNC)     # GlorotNormal
    raise KeyError(name)
