"""GPU parity tests (-m gpu) for the rows added in round 2: step variants (attention loss in the loss, adaptive gradient
clipping, the sharpness-aware step), the GEMM precision modes with their stated bounds, the random / beam decoders, and
the configurations the bench quotes its decode numbers on (B = 515 and B = 4096: the B > 64 encoder path)."""
import numpy as np
import pytest
import torch

from oracle import keras_semantics as K
from tests.util import pkg, syn, make_model, rel_err, grad_errors, oracle_to_named
from tests.test_gpu_parity import _case, _check_train_step, _assert_ids, LOSS_RTOL, GRAD_RTOL

pytestmark = pytest.mark.gpu


# ---- N3: step variants ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", ["batch", "elementwise"])
def test_attention_loss_in_the_loss_tiny(mode):
    _check_train_step(K.tiny_config(attn_loss=mode), 4, seed=5, steps=2, tag=f"attn-{mode}-tiny")


@pytest.mark.parametrize("mode", ["batch", "elementwise"])
def test_attention_loss_in_the_loss_full_config(mode):
    # lc_NIC.py:384 enabled: the persistent recurrence gradient carries dL/d alpha
    _check_train_step(K.Config(attn_loss=mode), 64, seed=41, full_groups=True, tag=f"attn-{mode}-C1")


def test_adaptive_gradient_clipping_tiny_and_full():
    # clip factors chosen so that a good share of the units is actually clipped
    _check_train_step(K.tiny_config(agc=(0.01, 1e-3)), 4, seed=6, steps=3, tag="agc-tiny")
    _check_train_step(K.Config(agc=(0.01, 1e-3)), 64, seed=43, full_groups=True, steps=2, tag="agc-C1")


def _sam_case(cfg, B, full_groups, seed):
    groups, P, batch = _case(cfg, B, full_groups=full_groups)
    model = make_model(cfg, groups, P, batch_max=B)
    opt = K.new_opt_state(P)
    x, cap, a0, c0, tgt = batch
    for s in range(2):
        met_o, _ = K.train_step_sam(cfg, P, opt, groups, batch, seed=seed + 2 * s)
        met_d = model.train_step_sam(((x, cap, a0, c0), tgt), seed=seed + 2 * s)
        for k in ("loss", "L2", "accuracy", "attention"):
            assert abs(met_d[k] - met_o[k]) <= LOSS_RTOL * max(abs(met_o[k]), 1e-3), (s, k, met_d[k], met_o[k])
    named, got = oracle_to_named(cfg, P), model.get_named_weights()
    for k in named:
        if k.startswith("attention/V/bias"):
            continue                                     # mathematically zero gradient: Adam amplifies rounding noise
        w = np.asarray(named[k], dtype=np.float64)
        d = np.abs(np.asarray(got[k], dtype=np.float64).reshape(w.shape) - w).max()
        assert d <= 2e-5, (k, d)


def test_sam_step_tiny():
    _sam_case(K.tiny_config(), 4, False, seed=9)


def test_sam_step_full_config():
    _sam_case(K.Config(), 64, True, seed=51)


# ---- X3: precision modes with their stated bounds (include/tnt.h: enum tnt_precision) ------------------------------
PRECISION_BOUNDS = {"bf16x3": (1e-3, 5e-3), "bf16": (1e-2, 1e-1)}      # (loss, gradient) relative


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
def test_reduced_precision_modes_meet_their_stated_bounds(precision):
    cfg = K.Config()
    groups, P, batch = _case(cfg, 64, full_groups=True)
    model = make_model(cfg, groups, P, batch_max=64, precision=precision)
    x, cap, a0, c0, tgt = batch
    met_d, grads_d = model.loss_and_gradients(((x, cap, a0, c0), tgt), seed=3)
    met_o, grads_o, _, _, _ = K.loss_and_grads(cfg, P, groups, batch, seed=3)
    lb, gb = PRECISION_BOUNDS[precision]
    for k in ("loss", "L2", "accuracy"):
        assert abs(met_d[k] - met_o[k]) <= lb * max(abs(met_o[k]), 1e-3), (precision, k, met_d[k], met_o[k])
    errs = grad_errors(cfg, grads_o, grads_d)
    gmax = max(n for _, n in errs.values())
    sig = {k: v for k, v in errs.items() if v[1] > 1e-6 * gmax and not k.endswith("/bias")}
    worst = max(sig.items(), key=lambda kv: kv[1][0])
    print(f"{precision}: worst gradient error {worst[1][0]:.2e} ({worst[0]}), loss error "
          f"{abs(met_d['loss'] - met_o['loss']) / met_o['loss']:.2e}")
    assert worst[1][0] <= gb, (precision, worst)
    assert precision in model.kernel_variants()


# ---- every kernel variant that stays in the tree meets the oracle --------------------------------------------------
def test_batch_of_96_rows_full_config():
    """B = 96 on one encoder: enc_fwd_mma_kernel (row-major gather, B > 64), the general enc_bwd_dw_kernel and six row
    groups of the persistent recurrence (two cooperative launches)."""
    model = _check_train_step(K.Config(), 96, seed=61, full_groups=True, tag="C1-B96")
    v = model.kernel_variants()
    assert "enc_fwd_mma_kernel" in v and "enc_bwd_dw_kernel" in v and "persistent" in v


def test_kernel_variants_of_the_benchmark_configuration():
    model = _check_train_step(K.Config(), 64, seed=71, full_groups=True, tag="C1-variants")
    v = model.kernel_variants()
    assert "enc_bwd=enc_bwd_dw" in v and "persistent" in v and "bf16x6" in v, v


# ---- the decode configurations the bench quotes --------------------------------------------------------------------
def _greedy_big(B, chunk=512):
    cfg = K.Config().no_dropout()
    groups, P, _ = _case(cfg, 4, full_groups=True)
    rng = np.random.default_rng(2)
    P["enc0_mean"] = (0.1 * rng.standard_normal(cfg.G)).astype(np.float32)
    P["enc0_var"] = (1.0 + 0.2 * rng.random(cfg.G)).astype(np.float32)
    model = make_model(cfg, groups, P, batch_max=B)
    x = np.random.default_rng(5).standard_normal((B, cfg.N)).astype(np.float32)
    z = np.zeros((B, cfg.U), np.float32)
    start = np.full(B, syn.START_ID, np.int32)
    got = model.greedy_predict(x, z, z, start, cfg.T, cfg.U, None)
    assert "enc_fwd_mma_kernel" in model.kernel_variants()
    n_ties = 0
    for s in range(0, B, chunk):                       # the oracle's rows are independent at inference: bound its memory
        e = min(B, s + chunk)
        want = K.greedy_predict(cfg, P, groups, x[s:e], z[s:e], z[s:e], start[s:e], cfg.T)
        sub = (got[0][s:e], got[1][s:e], got[2][:, s:e], got[3][:, s:e])
        n_ties += _assert_ids(want, sub)
        same = (want[0] == sub[0]).all(axis=(1, 2))                    # captions that never hit a near-tie
        assert same.mean() > 0.98
        assert rel_err(sub[1][same], want[1][same]) < 1e-3            # outputs_raw
        assert rel_err(sub[2][:, same], want[2][:, same]) < 1e-3      # attention scores
        assert rel_err(sub[3][:, same], want[3][:, same]) < 1e-3      # attention hidden layer
    return n_ties


def test_greedy_ids_at_the_test_split_batch_515():
    """eval.py:147-193 over the 515-key test split in ONE batch: the configuration bench.py's `greedy` line times."""
    assert _greedy_big(515) <= 8


def test_greedy_ids_at_batch_4096():
    """bench.py's `greedy_large` configuration."""
    assert _greedy_big(4096) <= 64


# ---- N4: random choice and beam search -----------------------------------------------------------------------------
def _assert_sampled(want, got, tol=2e-4):
    # the draw is an inverse CDF: device and oracle probabilities agree to ~1e-5, so a threshold that falls within 2e-4 of a
    # running-sum boundary (relative to the kept mass) may select the neighbouring word -- a near-tie, as for arg-max
    ids_w, ids_g, margins = want[0][:, :, 0], got[0][:, :, 0], want[4]
    first_bad = {}
    for b, t in np.argwhere(ids_w != ids_g):
        first_bad.setdefault(int(b), int(t))
    real = [(b, t, float(margins[t, b])) for b, t in first_bad.items() if margins[t, b] > tol]
    assert not real, f"sampled ids differ beyond near-ties: {real[:5]}"
    return len(first_bad)


@pytest.mark.parametrize("top_k,top_p", [(0, 0.0), (5, 0.0), (0, 0.6), (8, 0.9)])
def test_sample_predict_tiny(top_k, top_p):
    cfg = K.tiny_config().no_dropout()
    groups, P, batch = _case(cfg, 4)
    model = make_model(cfg, groups, P, batch_max=4)
    x, cap, a0, c0, _ = batch
    start = np.full(4, syn.START_ID, np.int32)
    for seed in (1, 2, 3):
        want = K.greedy_predict(cfg, P, groups, x, a0, c0, start, 6, sample=(top_k, top_p, seed))
        got = model.sample_predict(x, a0, c0, start, 6, cfg.U, None, top_k=top_k, top_p=top_p, seed=seed)
        _assert_sampled(want, got)
    # a different seed gives different captions; the same seed the same ones
    a = model.sample_predict(x, a0, c0, start, 6, seed=7)[0]
    b = model.sample_predict(x, a0, c0, start, 6, seed=7)[0]
    c = model.sample_predict(x, a0, c0, start, 6, seed=8)[0]
    assert (a == b).all() and (a != c).any()


def test_sample_predict_full_config():
    cfg = K.Config().no_dropout()
    groups, P, batch = _case(cfg, 32, full_groups=True)
    model = make_model(cfg, groups, P, batch_max=32)
    x, cap, a0, c0, _ = batch
    start = np.full(32, syn.START_ID, np.int32)
    for top_k, top_p in ((0, 0.0), (10, 0.0), (0, 0.5)):
        want = K.greedy_predict(cfg, P, groups, x, a0, c0, start, 15, sample=(top_k, top_p, 11))
        got = model.sample_predict(x, a0, c0, start, 15, cfg.U, None, top_k=top_k, top_p=top_p, seed=11)
        assert _assert_sampled(want, got) <= 2
    g = model.greedy_predict(x, a0, c0, start, 15)[0]
    s1 = model.sample_predict(x, a0, c0, start, 15, top_k=1, seed=5)[0]
    assert (g == s1).all()                                            # top-1 sampling is arg-max


def _check_beam(cfg, B, beam, max_len, full_groups, end_id=None):
    groups, P, batch = _case(cfg, B, full_groups=full_groups)
    model = make_model(cfg, groups, P, batch_max=B * beam)
    x, cap, a0, c0, _ = batch
    start = np.full(B, syn.START_ID, np.int32)
    ids_o, sc_o = K.beam_search(cfg, P, groups, x, start, max_len, beam=beam, end_id=-1 if end_id is None else end_id,
                                a0=a0, c0=c0)
    ids_d, sc_d = model.beam_search(x, start, max_len, cfg.U, beam=beam, end_id=end_id, a0=a0, c0=c0)
    assert ids_d.shape == (B, beam, max_len) and ids_d.dtype == np.int64
    assert np.allclose(sc_d, sc_o, rtol=1e-4, atol=1e-4)
    for b in range(B):
        # candidates closer than 1e-3 in score may swap places: compare beams that are clearly separated
        gaps = np.abs(np.diff(sc_o[b])) > 1e-3
        for w in range(beam):
            sep = (w == 0 or gaps[w - 1]) and (w == beam - 1 or gaps[w])
            if sep:
                assert (ids_d[b, w] == ids_o[b, w]).all(), (b, w, ids_d[b, w], ids_o[b, w])
    return ids_d, sc_d, model, (x, a0, c0, start)


def test_beam_search_tiny():
    ids, sc, model, (x, a0, c0, start) = _check_beam(K.tiny_config().no_dropout(), 4, 3, 6, False)
    g = model.greedy_predict(x, a0, c0, start, 6)[0][:, :, 0]
    one, _ = model.beam_search(x, start, 6, beam=1, a0=a0, c0=c0)
    assert (one[:, 0] == g).all()                                     # width 1 is greedy
    _check_beam(K.tiny_config().no_dropout(), 4, 3, 6, False, end_id=int(ids[0, 0, 1]))


def test_beam_search_full_config():
    _check_beam(K.Config().no_dropout(), 8, 4, 15, True, end_id=syn.END_ID if hasattr(syn, "END_ID") else 2)


# ---- N2: Keras HDF5 checkpoints ------------------------------------------------------------------------------------
def test_keras_h5_checkpoint_roundtrip(tmp_path):
    """ModelCheckpoint(save_weights_only=True) -> model-latest.h5 (main.py:180-190) and
    load_weights(path, by_name=True, skip_mismatch=True) (eval.py:140)."""
    cfg = K.tiny_config()
    groups, P, batch = _case(cfg, 4)
    m1 = make_model(cfg, groups, P, batch_max=4)
    path = str(tmp_path / "model-latest.h5")
    m1.save_weights(path)
    m2 = make_model(cfg, groups, None, batch_max=4, seed=99)
    rep = m2.load_weights(path, by_name=True, skip_mismatch=True)
    assert not rep["missing"] and not rep["skipped"] and not rep["unexpected"]
    for a, b in zip(m1.get_weights(), m2.get_weights()):
        assert np.array_equal(a, b)
    m3 = make_model(cfg, groups, None, batch_max=4, seed=98)
    m3.load_weights(path)                                            # topological loading
    for a, b in zip(m1.get_weights(), m3.get_weights()):
        assert np.array_equal(a, b)
    # a model of another shape: strict loading raises, skip_mismatch reports what it skipped
    cfg2 = K.tiny_config(U=24)
    g2, P2, _ = _case(cfg2, 4)
    m4 = make_model(cfg2, g2, P2, batch_max=4)
    with pytest.raises(ValueError):
        m4.load_weights(path, by_name=True)
    rep = m4.load_weights(path, by_name=True, skip_mismatch=True)
    assert "lstm/kernel" in rep["skipped"] and "enc0/dense_0/kernel" in rep["loaded"]
