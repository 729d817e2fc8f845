"""CPU tests: the oracle against known answers (Philox KAT, finite differences, hand-checkable cases)."""
import numpy as np
import torch

from oracle import philox, keras_semantics as K
from tests.util import syn


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32-10
    r = philox.philox4x32_10(0, 0, 0, 0, 0, 0)
    assert [int(x) for x in r] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    r = philox.philox4x32_10(0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff)
    assert [int(x) for x in r] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    r = philox.philox4x32_10(0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344, 0xa4093822, 0x299f31d0)
    assert [int(x) for x in r] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_keep_mask_rate_and_scale():
    m = philox.keep_mask((1000, 100), 0.2, philox.SITE_TEXT, 0, 1234)
    assert set(np.unique(m)) == {0.0, 1.25}
    assert abs((m > 0).mean() - 0.8) < 0.01
    assert (philox.keep_mask((7,), 0.0, 1, 0, 1) == 1).all()
    # order independence: element i only depends on (i, site, step, seed)
    m2 = philox.keep_mask((50,), 0.2, philox.SITE_TEXT, 0, 1234)
    assert (m.reshape(-1)[:50] == m2).all()


def _tiny(n_enc=1, dropout=False, **kw):
    cfg = K.tiny_config(n_enc=n_enc, **kw)
    if not dropout:
        cfg = cfg.no_dropout()
    groups = [syn.small_groups(cfg.N, cfg.R, seed=1 + e, overlap=True) for e in range(n_enc)]
    P = K.init_params(cfg, groups, seed=3)
    # h0 = 0 puts LeakyReLU(h @ W2 + b2) exactly on its kink at t = 0 with the zero-initialised bias;
    # move the biases off zero so finite differences are defined
    rng = np.random.default_rng(11)
    for k in ("att_b1", "att_b2", "inter_b", "out_b", "lstm_b"):
        P[k] = (P[k] + 0.05 * rng.standard_normal(P[k].shape)).astype(np.float32)
    B = 4
    (x, cap, a0, c0), tgt = syn.make_batch(B, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
    return cfg, groups, P, (x, cap, a0, c0, tgt)


def _total(cfg, groups, Pd, batch):
    Pt = K.to_torch(Pd, torch.float64)
    xb = K._batch_to_torch(batch, torch.float64)
    o = K.forward(cfg, Pt, groups, xb[:4], True, 0)
    m = K._metrics(cfg, Pt, o, xb[4])
    return float(m["loss"] + m["L2"])


def test_tape_gradient_matches_finite_differences():
    for n_enc in (1, 2):
        cfg, groups, P, batch = _tiny(n_enc)
        _, grads, _, _, _ = K.loss_and_grads(cfg, P, groups, batch, dtype=torch.float64)
        rng = np.random.default_rng(0)
        for name in K.trainable_names(cfg):
            Pd = {k: ([a.astype(np.float64) for a in v] if isinstance(v, list) else v.astype(np.float64))
                  for k, v in P.items()}
            lst = isinstance(Pd[name], list)
            arr = Pd[name][2] if lst else Pd[name]
            g = grads[name][2] if lst else grads[name]
            idx = tuple(int(rng.integers(0, s)) for s in arr.shape)
            if name == "emb":
                idx = (int(batch[1][0, 1]), 3)
            eps = 1e-6
            old = arr[idx]
            arr[idx] = old + eps; lp = _total(cfg, groups, Pd, batch)
            arr[idx] = old - eps; lm = _total(cfg, groups, Pd, batch)
            fd = (lp - lm) / (2 * eps)
            assert abs(fd - g[idx]) <= 1e-6 + 1e-4 * abs(fd), (name, idx, fd, g[idx])


def test_loss_is_unmasked_ce_on_clipped_probs():
    # one row, hand-computed: uniform logits -> p = 1/V, loss = log V for every position incl. pads
    cfg = K.tiny_config().no_dropout()
    probs = torch.full((2, cfg.T, cfg.V), 1.0 / cfg.V)
    tgt = torch.zeros((2, cfg.T), dtype=torch.long)            # all pads: still counted (lc_NIC.py:370)
    ce, acc = K.ce_and_accuracy(tgt, probs, cfg.V)
    assert abs(float(ce) - np.log(cfg.V)) < 1e-6
    assert float(acc) == 1.0                                   # argmax of a constant row is index 0 == pad
    # clipping: p_y = 0 -> loss = -log(1e-7)
    probs = torch.zeros((1, 1, 4)); probs[0, 0, 1] = 1.0
    ce = K.categorical_crossentropy(torch.tensor([[1.0, 0, 0, 0]]), probs[0])
    assert abs(float(ce) + np.log(1e-7)) < 1e-4


def test_attention_metric_sums_over_batch_axis():
    attn = torch.rand(3, 4, 5, 1)
    want = ((1 - attn.squeeze(-1).sum(dim=1)) ** 2).mean()
    assert abs(float(K.attention_mse(attn)) - float(want)) < 1e-7


def test_adam_clipnorm_known_answer():
    cfg = K.tiny_config().no_dropout()
    P = {"x": np.array([1.0, 2.0], np.float32)}
    # direct formula check on one variable
    g = np.array([3.0, 4.0], np.float32)               # norm 5 -> scaled to norm 0.1
    gc = g * 0.1 / 5.0
    m = 0.1 * gc
    v = 0.02 * gc * gc
    alpha = 1e-4 * np.sqrt(1 - 0.98) / (1 - 0.9)
    want = P["x"] - alpha * m / (np.sqrt(v) + 1e-8)
    # through the oracle's implementation
    import types
    names = K.trainable_names
    try:
        K.trainable_names = lambda c: ["x"]
        opt = dict(t=0, m={"x": np.zeros(2, np.float32)}, v={"x": np.zeros(2, np.float32)})
        K.adam_apply(cfg, P, opt, {"x": g}, 0.0)
    finally:
        K.trainable_names = names
    assert np.allclose(P["x"], want, rtol=1e-6)


def test_train_step_reduces_loss_and_updates_moving_stats():
    cfg, groups, P, batch = _tiny(1, dropout=True)
    opt = K.new_opt_state(P)
    cfg.lr = 1e-2
    l0 = K.train_step(cfg, P, opt, groups, batch, seed=1)[0]["loss"]
    for i in range(30):
        l = K.train_step(cfg, P, opt, groups, batch, seed=1)[0]["loss"]
    assert l < l0
    assert not np.allclose(P["enc0_mean"], 0.0)


def test_greedy_shapes_and_first_max_tiebreak():
    cfg, groups, P, batch = _tiny(1)
    x, cap, a0, c0, _ = batch
    out, raw, sc, hid = K.greedy_predict(cfg, P, groups, x, a0, c0, np.full(4, syn.START_ID), 6)
    assert out.shape == (4, 6, 1) and out.dtype == np.int64
    assert raw.shape == (4, 6, cfg.V) and sc.shape == (6, 4, cfg.R, 1) and hid.shape == (6, 4, cfg.R, cfg.A)
    assert (out[:, :, 0] == raw.argmax(-1)).all()
    assert np.allclose(sc.sum(axis=2), 1.0, atol=1e-5)


def test_masked_sparse_ce_variant_known_answers():
    """A11 loss (ThinkAndTell/att_model.py:256-262,306-321): uniform logits -> log V per non-pad target; pads add 0
    to the sum but stay in the batch mean; T step losses are divided by T + 1."""
    V, T = 7, 4
    logits = torch.zeros((3, T, V), dtype=torch.float64)
    y = torch.tensor([[1, 0, 2, 0], [0, 0, 3, 6], [5, 1, 0, 0]])
    ce, acc = K.masked_sparse_ce_and_accuracy(y, logits)
    nonpad_per_step = [2, 1, 2, 1]
    want = sum(n / 3.0 * np.log(V) for n in nonpad_per_step) / (T + 1)
    assert abs(float(ce) - want) < 1e-12
    # random logits against torch's own sparse CE
    logits = torch.randn(3, T, V, dtype=torch.float64)
    ce, _ = K.masked_sparse_ce_and_accuracy(y, logits)
    ref = sum((torch.nn.functional.cross_entropy(logits[:, t], y[:, t], reduction="none") * (y[:, t] != 0)).mean()
              for t in range(T)) / (T + 1)
    assert abs(float(ce) - float(ref)) < 1e-12
    # the variant runs end to end through loss_and_grads and the gradient of a masked row's logits is zero
    cfg, groups, P, batch = _tiny(1)
    cfg = K.tiny_config(masked_sparse_ce=True, plain_attention=True).no_dropout()
    met, grads, _, _, out = K.loss_and_grads(cfg, P, groups, batch, dtype=torch.float64)
    assert np.isfinite(met["loss"]) and all(np.isfinite(g).all() for g in grads.values() if not isinstance(g, list))
