"""CPU tests: the oracle's building blocks against INDEPENDENT implementations that ship with this image
(torch.nn / torch.nn.functional / torch.optim, plain numpy).  TensorFlow cannot run here, so parity with Keras stays
unpinned (oracle/__init__.py); these checks make sure the restated blocks are at least not single-sourced: each one is
a second implementation of the same published definition (gate order i,f,g,o; biased batch variance; scatter-add
embedding gradient; softmax cross-entropy; clip-by-norm; Adam)."""
import math
import numpy as np
import torch
import torch.nn.functional as F

from oracle import keras_semantics as K


def _t(a):
    return torch.as_tensor(np.asarray(a), dtype=torch.float64)


def test_lstm_step_matches_torch_lstmcell():
    # Keras LSTMCell: z = x W + h U + b, split i, f, c~, o  ==  torch.nn.LSTMCell's i, f, g, o
    cfg = K.tiny_config().no_dropout()
    rng = np.random.default_rng(0)
    P = {"lstm_W": _t(rng.standard_normal((cfg.G + cfg.E, 4 * cfg.U)) * 0.3),
         "lstm_U": _t(rng.standard_normal((cfg.U, 4 * cfg.U)) * 0.3), "lstm_b": _t(rng.standard_normal(4 * cfg.U) * 0.1)}
    x, h, c = _t(rng.standard_normal((5, cfg.G + cfg.E))), _t(rng.standard_normal((5, cfg.U))), _t(rng.standard_normal((5, cfg.U)))
    h2, c2 = K.lstm_step(cfg, P, x, h, c, 0, False, 0)
    cell = torch.nn.LSTMCell(cfg.G + cfg.E, cfg.U).double()
    with torch.no_grad():
        cell.weight_ih.copy_(P["lstm_W"].T); cell.weight_hh.copy_(P["lstm_U"].T)
        cell.bias_ih.copy_(P["lstm_b"]); cell.bias_hh.zero_()
        h3, c3 = cell(x, (h, c))
    assert torch.allclose(h2, h3, atol=1e-12) and torch.allclose(c2, c3, atol=1e-12)


def test_batchnorm_matches_functional_batch_norm():
    # BatchNormalization(axis=-1) on (B,R,G): per-channel statistics over (B,R), biased variance, eps 1e-3
    cfg = K.tiny_config().no_dropout()
    rng = np.random.default_rng(1)
    groups = [[np.arange(r * 10, r * 10 + 10) for r in range(cfg.R)]]
    P = K.to_torch(K.init_params(cfg, groups, seed=2), torch.float64)
    P["enc0_gamma"] = _t(1 + 0.1 * rng.standard_normal(cfg.G)); P["enc0_beta"] = _t(0.1 * rng.standard_normal(cfg.G))
    x = _t(rng.standard_normal((6, cfg.N)))
    stats = {}
    y, pre = K.encoder(cfg, P, 0, groups, x, True, 0, stats)
    rm, rv = torch.zeros(cfg.G, dtype=torch.float64), torch.ones(cfg.G, dtype=torch.float64)
    ref = F.batch_norm(pre.permute(0, 2, 1), rm, rv, P["enc0_gamma"], P["enc0_beta"], training=True,
                       momentum=1 - K.BN_MOMENTUM, eps=K.BN_EPS).permute(0, 2, 1)
    assert torch.allclose(y, ref, atol=1e-10)
    # moving statistics: Keras moves by the BIASED batch variance, torch by the unbiased one
    n = pre.shape[0] * pre.shape[1]
    Pn = {"enc0_mean": np.zeros(cfg.G, np.float32), "enc0_var": np.ones(cfg.G, np.float32)}
    K.update_moving_stats(cfg, Pn, {k: v.numpy() for k, v in stats.items()})
    assert np.allclose(Pn["enc0_mean"], rm.numpy(), atol=1e-6)
    assert np.allclose((Pn["enc0_var"] - 0.99) * n / (n - 1) + 0.99, rv.numpy(), atol=1e-6)
    # inference path: moving statistics
    P["enc0_mean"], P["enc0_var"] = _t(0.1 * rng.standard_normal(cfg.G)), _t(1 + 0.2 * rng.random(cfg.G))
    y2, pre2 = K.encoder(cfg, P, 0, groups, x, False, 0)
    ref2 = F.batch_norm(pre2.permute(0, 2, 1), P["enc0_mean"], P["enc0_var"], P["enc0_gamma"], P["enc0_beta"],
                        training=False, eps=K.BN_EPS).permute(0, 2, 1)
    assert torch.allclose(y2, ref2, atol=1e-10)


def test_parcel_dense_matches_functional_linear_and_leaky_relu():
    cfg = K.tiny_config().no_dropout()
    rng = np.random.default_rng(3)
    groups = [[rng.choice(cfg.N, size=7 + r, replace=False) for r in range(cfg.R)]]
    P = K.to_torch(K.init_params(cfg, groups, seed=4), torch.float64)
    x = _t(rng.standard_normal((3, cfg.N)))
    _, pre = K.encoder(cfg, P, 0, groups, x, False, 0)
    for r, idx in enumerate(groups[0]):
        ref = F.leaky_relu(F.linear(x[:, torch.as_tensor(idx)], P["enc0_W"][r].T, P["enc0_b"][r]), 0.2)
        assert torch.allclose(pre[:, r], ref, atol=1e-12)


def test_attention_matches_softmax_bmm_formulation():
    cfg = K.tiny_config().no_dropout()
    rng = np.random.default_rng(5)
    P = {k: _t(rng.standard_normal(s) * 0.4) for k, s in
         dict(att_W1=(cfg.G, cfg.A), att_b1=(cfg.A,), att_W2=(cfg.U, cfg.A), att_b2=(cfg.A,), att_V=(cfg.A, 1),
              att_bV=(1,)).items()}
    f, h = _t(rng.standard_normal((4, cfg.R, cfg.G))), _t(rng.standard_normal((4, cfg.U)))
    ctx, w, s = K.attention(cfg, P, h, f, 0, False, 0)
    e = torch.tanh(F.leaky_relu(F.linear(f, P["att_W1"].T, P["att_b1"]), 0.2) +
                   F.leaky_relu(F.linear(h, P["att_W2"].T, P["att_b2"]), 0.2)[:, None, :])
    score = F.linear(e, P["att_V"].T, P["att_bV"])[:, :, 0]
    a = F.softmax(score, dim=1)
    ref_ctx = torch.bmm(a[:, None, :], f)[:, 0]
    assert torch.allclose(w[:, :, 0], a, atol=1e-12) and torch.allclose(ctx, ref_ctx, atol=1e-12)
    assert torch.allclose(s, e, atol=1e-12)


def test_embedding_gradient_is_scatter_add_of_slices():
    # Embedding gradient = IndexedSlices(values (B*T,E), indices); dense form = scatter-add (F.embedding's autograd)
    rng = np.random.default_rng(6)
    V, E = 11, 5
    table = _t(rng.standard_normal((V, E))).requires_grad_(True)
    ids = torch.as_tensor(rng.integers(0, V, size=(4, 6)))
    up = _t(rng.standard_normal((4, 6, E)))
    (F.embedding(ids, table) * up).sum().backward()
    dense = np.zeros((V, E))
    np.add.at(dense, ids.numpy().reshape(-1), up.numpy().reshape(-1, E))
    assert np.allclose(table.grad.numpy(), dense, atol=1e-12)
    t2 = _t(table.detach().numpy()).requires_grad_(True)
    ev = t2[ids.long()]; ev.retain_grad()
    (ev * up).sum().backward()
    assert np.allclose(ev.grad.numpy(), up.numpy()) and np.allclose(t2.grad.numpy(), dense, atol=1e-12)


def test_categorical_crossentropy_matches_nll_of_log_probs():
    # Keras categorical_crossentropy(from_logits=False) away from the clip == NLL of log-probabilities
    rng = np.random.default_rng(7)
    logits = _t(rng.standard_normal((8, 13)))
    y = torch.as_tensor(rng.integers(0, 13, size=8))
    probs = torch.softmax(logits, dim=-1)
    ce = K.categorical_crossentropy(F.one_hot(y, 13).double(), probs)
    assert torch.allclose(ce, F.cross_entropy(logits, y, reduction="none"), atol=1e-10)
    # the clip: a target probability below 1e-7 is clipped (loss = -log 1e-7) and passes no gradient
    lg = torch.zeros(1, 3, dtype=torch.float64); lg[0, 0] = -40.0
    lg.requires_grad_(True)
    c2 = K.categorical_crossentropy(F.one_hot(torch.tensor([0]), 3).double(), torch.softmax(lg, -1))
    assert abs(float(c2) + math.log(1e-7)) < 1e-9
    c2.sum().backward()
    assert float(lg.grad.abs().max()) == 0.0


def test_masked_sparse_ce_matches_cross_entropy_with_mask():
    # ThinkAndTell loss (att_model.py:306-321): sparse CE on logits x (y != 0), batch mean keeps the zeros
    rng = np.random.default_rng(8)
    B, T, V = 5, 4, 9
    logits = _t(rng.standard_normal((B, T, V)))
    tgt = torch.as_tensor(rng.integers(0, V, size=(B, T))); tgt[0, 1] = 0; tgt[3, 2] = 0
    ce, acc = K.masked_sparse_ce_and_accuracy(tgt, logits)
    per = F.cross_entropy(logits.reshape(-1, V), tgt.reshape(-1), reduction="none").reshape(B, T) * (tgt != 0)
    assert abs(float(ce) - float(per.mean(dim=0).sum() / (T + 1))) < 1e-12
    assert abs(float(acc) - float((logits.argmax(-1) == tgt).double().mean())) < 1e-12


def test_clip_by_norm_matches_clip_grad_norm():
    rng = np.random.default_rng(9)
    for scale in (0.01, 3.0):
        g = rng.standard_normal(50) * scale
        s = K.clip_by_norm_scale(float((g ** 2).sum()), 0.1)
        p = torch.nn.Parameter(torch.zeros(50, dtype=torch.float64)); p.grad = _t(g.copy())
        torch.nn.utils.clip_grad_norm_([p], 0.1)
        assert np.allclose(g * s, p.grad.numpy(), rtol=2e-5)      # torch divides by (norm + 1e-6)


def _numpy_adam(w, g, m, v, t, lr, b1, b2, eps):
    """ResourceApplyAdam as TensorFlow documents it (training_ops: lr_t = lr sqrt(1-b2^t)/(1-b1^t))."""
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    lr_t = lr * math.sqrt(1 - b2 ** t) / (1 - b1 ** t)
    return w - lr_t * m / (np.sqrt(v) + eps), m, v


def test_adam_matches_independent_numpy_and_torch_adam():
    cfg = K.tiny_config().no_dropout()
    cfg.apply_clipnorm = False
    groups = [[np.arange(r * 10, r * 10 + 10) for r in range(cfg.R)]]
    P = K.init_params(cfg, groups, seed=5)
    opt = K.new_opt_state(P)
    rng = np.random.default_rng(10)
    w_np = P["lstm_U"].astype(np.float64).copy(); m_np = np.zeros_like(w_np); v_np = np.zeros_like(w_np)
    pt = torch.nn.Parameter(_t(P["lstm_U"]))
    topt = torch.optim.Adam([pt], lr=cfg.lr, betas=(cfg.beta1, cfg.beta2), eps=cfg.adam_eps)
    for t in range(1, 4):
        grads = {n: ([np.zeros_like(a) for a in P[n]] if isinstance(P[n], list) else np.zeros_like(P[n]))
                 for n in K.trainable_names(cfg)}
        g = rng.standard_normal(P["lstm_U"].shape).astype(np.float32)
        grads["lstm_U"] = g
        K.adam_apply(cfg, P, opt, grads, 0.0)
        w_np, m_np, v_np = _numpy_adam(w_np, g.astype(np.float64), m_np, v_np, t, cfg.lr, cfg.beta1, cfg.beta2, cfg.adam_eps)
        pt.grad = _t(g); topt.step()
        assert np.allclose(P["lstm_U"], w_np, atol=2e-7)
        # torch places epsilon differently (eps * sqrt(1-b2^t) in TF's terms): equal while sqrt(v) >> eps
        assert np.allclose(P["lstm_U"], pt.detach().numpy(), atol=2e-7)
