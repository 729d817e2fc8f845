"""TEST INFRASTRUCTURE (CPU oracle) -- not imported by the product.

Restatement of the encoder of AttemptFour/Model/NIC.py (SURVEY 8a row A1''):

    img_input = self.dropout_input(img_input)                      NIC.py:122
    features  = self.dense_in(img_input)        Dense(embedding_dim, LeakyReLU(0.2), GlorotNormal, L2(input_reg))  NIC.py:64-69,125
    features  = self.dropout(features)                              NIC.py:126
    features  = self.expand(features)            (B, 1, E)           NIC.py:127
    features  = self.batch_norm(features)        BatchNormalization() defaults: axis -1, momentum .99, eps 1e-3   NIC.py:128

and of what `tape.gradient` + `Adam(clipnorm).apply_gradients` do to its variables.  Parity unpinned (no TensorFlow
here): Keras semantics restated as in keras_semantics.py; dropout through the shared Philox contract (sites INPUT and
FEAT with index b*K + k and b*E + e)."""
import math

import numpy as np
import torch

from . import philox
from .keras_semantics import BN_EPS, BN_MOMENTUM, _mask, glorot_normal, lrelu


def init_params(K, E, seed=0):
    rng = np.random.default_rng(seed)
    return {"W": glorot_normal(rng, (K, E)), "b": np.zeros(E, np.float32), "gamma": np.ones(E, np.float32),
            "beta": np.zeros(E, np.float32), "mean": np.zeros(E, np.float32), "var": np.ones(E, np.float32)}


TRAINABLE = ("W", "b", "gamma", "beta")


def forward(P, x, training, seed=0, dropout_input=0.0, dropout=0.0, stats_out=None):
    """P: dict of torch tensors; x (B,K).  Returns features (B,1,E)."""
    if training and dropout_input > 0:
        x = x * _mask(tuple(x.shape), dropout_input, philox.SITE_INPUT, 0, seed, True, x.dtype)
    a = lrelu(x @ P["W"] + P["b"])
    if training and dropout > 0:
        a = a * _mask(tuple(a.shape), dropout, philox.SITE_FEAT, 0, seed, True, a.dtype)
    f = a.unsqueeze(1)                                           # (B,1,E)
    if training:
        mean = f.mean(dim=(0, 1))
        var = ((f - mean) ** 2).mean(dim=(0, 1))                 # biased, as tf.nn.moments
        if stats_out is not None:
            stats_out["mean"], stats_out["var"] = mean.detach(), var.detach()
    else:
        mean, var = P["mean"], P["var"]
    inv = P["gamma"] * torch.rsqrt(var + BN_EPS)
    return f * inv + (P["beta"] - mean * inv)


def loss_and_grads(P_np, x, dfeatures, seed=0, dropout_input=0.0, dropout=0.0, input_reg=0.0, dtype=torch.float32):
    """Gradient of <features, dfeatures> + input_reg*||W||^2 -- i.e. what the tape hands the optimizer when the
    rest of the model back-propagates `dfeatures` into the encoder output."""
    P = {k: torch.as_tensor(v, dtype=dtype).clone().requires_grad_(k in TRAINABLE) for k, v in P_np.items()}
    stats = {}
    f = forward(P, torch.as_tensor(x, dtype=dtype), True, seed, dropout_input, dropout, stats)
    total = (f[:, 0, :] * torch.as_tensor(dfeatures, dtype=dtype)).sum() + input_reg * (P["W"] ** 2).sum()
    total.backward()
    grads = {k: P[k].grad.detach().numpy() for k in TRAINABLE}
    return f.detach().numpy(), grads, {k: v.numpy() for k, v in stats.items()}


def adam_apply(P_np, opt, grads, lr=1e-4, beta1=0.9, beta2=0.98, eps=1e-8, clipnorm=0.1):
    """Adam(clipnorm) per variable (main.py:97 hyper-parameters by default).  In place."""
    opt["t"] += 1
    t = opt["t"]
    alpha = lr * math.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
    for n in TRAINABLE:
        g = grads[n].astype(np.float32)
        if clipnorm is not None and clipnorm > 0:
            sq = float((g.astype(np.float64) ** 2).sum())
            g = g * np.float32(clipnorm) / np.float32(max(math.sqrt(sq), clipnorm))
        m, v = opt["m"][n], opt["v"][n]
        m += (g - m) * np.float32(1.0 - beta1)
        v += (g * g - v) * np.float32(1.0 - beta2)
        P_np[n] -= np.float32(alpha) * m / (np.sqrt(v) + np.float32(eps))


def update_moving_stats(P_np, stats):
    for k in ("mean", "var"):
        P_np[k] = P_np[k] * np.float32(BN_MOMENTUM) + stats[k].astype(np.float32) * np.float32(1.0 - BN_MOMENTUM)


def new_opt_state(P_np):
    return {"t": 0, "m": {k: np.zeros_like(P_np[k]) for k in TRAINABLE}, "v": {k: np.zeros_like(P_np[k]) for k in TRAINABLE}}
