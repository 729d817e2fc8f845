"""CPU tests of the oracle restatements added for SURVEY 8f N3 / N4 (step variants, sampling decoders, BLEU):
finite differences, hand-computed cases and defining properties."""
import math
import numpy as np
import torch

from oracle import keras_semantics as K, philox
from tests.util import pkg, syn
from tests.test_oracle import _tiny


def _fd_check(cfg, names, scale=1e-6):
    _, groups, P, batch = _tiny()
    _, grads, _, _, _ = K.loss_and_grads(cfg, P, groups, batch, dtype=torch.float64)
    def total(Pd):
        Pt = K.to_torch(Pd, torch.float64)
        xb = K._batch_to_torch(batch, torch.float64)
        o = K.forward(cfg, Pt, groups, xb[:4], True, 0)
        m = K._metrics(cfg, Pt, o, xb[4])
        extra = m["attention"] if cfg.attn_loss == "batch" else K.attention_mse_elementwise(o["attn"])
        return float(m["loss"] + m["L2"] + extra)
    rng = np.random.default_rng(0)
    for name in names:
        Pd = {k: ([a.astype(np.float64) for a in v] if isinstance(v, list) else v.astype(np.float64)) for k, v in P.items()}
        tgt = Pd[name][1] if isinstance(Pd[name], list) else Pd[name]
        g = grads[name][1] if isinstance(grads[name], list) else grads[name]
        for _ in range(3):
            i = tuple(int(rng.integers(0, s)) for s in tgt.shape)
            old = tgt[i]
            tgt[i] = old + scale; up = total(Pd)
            tgt[i] = old - scale; dn = total(Pd)
            tgt[i] = old
            fd = (up - dn) / (2 * scale)
            assert abs(fd - g[i]) <= 1e-5 * max(1.0, abs(fd)), (name, i, fd, g[i])


def test_attention_loss_gradient_finite_differences():
    # lc_NIC.py:384 enabled / train_step_sam's element-wise form: the extra term reaches the encoder and attention
    for mode in ("batch", "elementwise"):
        cfg = K.tiny_config(attn_loss=mode).no_dropout()
        _fd_check(cfg, ["att_W1", "att_W2", "att_V", "enc0_W", "lstm_U"])


def test_attention_loss_changes_only_the_gradient():
    cfg0, groups, P, batch = _tiny()
    m0, g0, _, _, _ = K.loss_and_grads(cfg0, P, groups, batch)
    cfg1 = K.tiny_config(attn_loss="batch").no_dropout()
    m1, g1, _, _, _ = K.loss_and_grads(cfg1, P, groups, batch)
    assert m0["loss"] == m1["loss"] and m0["attention"] == m1["attention"]
    assert not np.allclose(g0["att_V"], g1["att_V"])
    assert np.allclose(g0["out_W"], g1["out_W"])          # the head does not see the attention loss


def test_agc_hand_case():
    # agc.py:20-38 on a 2-D kernel: per-column norms; a column whose gradient norm exceeds clip * max(||w||, eps) is
    # rescaled onto that bound, the others are untouched
    cfg = K.tiny_config(agc=(0.5, 1e-3)).no_dropout()
    cfg.n_enc = 1
    names = K.trainable_names(cfg)
    P = {n: np.zeros((2, 2), np.float32) for n in names}
    g = {n: np.zeros((2, 2), np.float32) for n in names}
    for n in ("enc0_W", "enc0_b"):
        P[n] = [np.zeros((2, 2), np.float32)]; g[n] = [np.zeros((2, 2), np.float32)]
    P["lstm_U"] = np.array([[3.0, 0.0], [4.0, 0.0]], np.float32)     # column norms 5, 0 -> bounds 2.5, 0.0005
    g["lstm_U"] = np.array([[6.0, 0.0003], [8.0, 0.0]], np.float32)  # column norms 10, 0.0003
    P["emb"] = np.ones((3, 2), np.float32)                           # column norms sqrt(3) -> bound 0.866
    vals = np.array([[3.0, 0.1], [4.0, 0.1]])                        # un-aggregated slices: column norms 5, 0.141
    g["emb"] = np.array([[7.0, 0.2], [0, 0], [0, 0]], np.float32)    # both slices hit row 0
    out, emb_sq = K.adaptive_clip_grad(cfg, P, g, vals)
    assert np.allclose(out["lstm_U"], [[1.5, 0.0003], [2.0, 0.0]], atol=1e-6)
    f = math.sqrt(3) * 0.5 / 5.0
    assert np.allclose(out["emb"], [[7.0 * f, 0.2], [0, 0], [0, 0]], atol=1e-6)
    assert abs(emb_sq - ((3 * f) ** 2 + (4 * f) ** 2 + 0.02)) < 1e-9


def test_sam_step_is_two_passes_and_restores_weights():
    cfg, groups, P, batch = _tiny()
    P1 = {k: ([a.copy() for a in v] if isinstance(v, list) else v.copy()) for k, v in P.items()}
    P2 = {k: ([a.copy() for a in v] if isinstance(v, list) else v.copy()) for k, v in P.items()}
    o1, o2 = K.new_opt_state(P1), K.new_opt_state(P2)
    m_sam, _ = K.train_step_sam(cfg, P1, o1, groups, batch, seed=3, rho=0.05)
    m_std, _ = K.train_step(cfg, P2, o2, groups, batch, seed=3)
    assert m_sam["attention"] == 0.0 and o1["t"] == 1
    # rho -> 0 turns the second pass into the plain step at the same weights (no dropout: the seed does not matter)
    P3 = {k: ([a.copy() for a in v] if isinstance(v, list) else v.copy()) for k, v in P.items()}
    o3 = K.new_opt_state(P3)
    K.train_step_sam(cfg, P3, o3, groups, batch, seed=3, rho=0.0)
    for n in ("lstm_U", "att_W2", "out_W"):
        assert np.allclose(P3[n], P2[n], atol=1e-7)
        assert not np.allclose(P1[n], P2[n], atol=1e-9)       # rho = 0.05 moves the step
    # weights differ from the start only by an Adam step (|dw| <= ~lr): the climb was undone
    assert np.abs(P1["lstm_U"] - P["lstm_U"]).max() < 5 * cfg.lr


def test_choose_words_properties():
    rng = np.random.default_rng(0)
    p = rng.random((6, 50)).astype(np.float32); p /= p.sum(axis=1, keepdims=True)
    ids, _ = K.choose_words(p, step=2, top_k=1, seed=7)
    assert (ids == p.argmax(axis=1)).all()                              # top-1 == greedy
    ids, _ = K.choose_words(p, step=2, top_p=1e-6, seed=7)
    assert (ids == p.argmax(axis=1)).all()                              # the nucleus always holds the mode
    for k in (3, 10):
        ids, _ = K.choose_words(p, step=1, top_k=k, seed=11)
        topk = np.argsort(-p, axis=1, kind="stable")[:, :k]
        assert all(ids[b] in topk[b] for b in range(6))
    # nucleus: the chosen word lies in the shortest most-likely prefix whose mass exceeds top_p
    ids, _ = K.choose_words(p, step=0, top_p=0.3, seed=5)
    for b in range(6):
        order = np.argsort(-p[b], kind="stable"); cs = np.cumsum(p[b][order].astype(np.float64))
        n = int(np.nonzero(cs > np.float32(0.3))[0][0]) + 1
        assert ids[b] in order[:n]
    # the draw is a function of (row, step, seed) only, and follows the distribution
    q = np.tile(np.array([[0.5, 0.3, 0.2]], np.float32), (4000, 1))
    ids, _ = K.choose_words(q, step=0, seed=123)
    ids2, _ = K.choose_words(q[:10], step=0, seed=123)
    assert (ids[:10] == ids2).all()
    freq = np.bincount(ids, minlength=3) / 4000.0
    assert np.abs(freq - [0.5, 0.3, 0.2]).max() < 0.03
    u = philox.sample_uniform(4000, 0, 123)
    assert 0.0 < u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.02


def _greedy_setup():
    cfg = K.tiny_config().no_dropout()
    groups = [syn.small_groups(cfg.N, cfg.R, seed=1, overlap=True)]
    P = K.init_params(cfg, groups, seed=3)
    (x, cap, a0, c0), _ = syn.make_batch(4, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
    return cfg, groups, P, x, a0, c0, np.full(4, syn.START_ID, np.int32)


def test_beam_search_width_one_is_greedy_and_scores_are_ordered():
    cfg, groups, P, x, a0, c0, start = _greedy_setup()
    g = K.greedy_predict(cfg, P, groups, x, a0, c0, start, 6)[0][:, :, 0]
    ids, sc = K.beam_search(cfg, P, groups, x, start, 6, beam=1, a0=a0, c0=c0)
    assert (ids[:, 0] == g).all()
    ids3, sc3 = K.beam_search(cfg, P, groups, x, start, 6, beam=3, a0=a0, c0=c0)
    assert ids3.shape == (4, 3, 6) and (np.diff(sc3, axis=1) <= 1e-6).all()
    assert (sc3[:, 0] >= sc[:, 0] - 1e-5).all()                          # a wider beam never scores worse
    # frozen beams: after end_id only pad follows and the score stops moving
    end = int(ids3[0, 0, 1])
    ids_e, _ = K.beam_search(cfg, P, groups, x, start, 6, beam=3, end_id=end, a0=a0, c0=c0)
    for b in range(4):
        for w in range(3):
            hit = np.nonzero(ids_e[b, w] == end)[0]
            if hit.size:
                assert (ids_e[b, w, hit[0] + 1:] == 0).all()


def test_sample_predict_oracle_shapes_and_reproducibility():
    cfg, groups, P, x, a0, c0, start = _greedy_setup()
    a = K.greedy_predict(cfg, P, groups, x, a0, c0, start, 5, sample=(0, 0.0, 42))
    b = K.greedy_predict(cfg, P, groups, x, a0, c0, start, 5, sample=(0, 0.0, 42))
    c = K.greedy_predict(cfg, P, groups, x, a0, c0, start, 5, sample=(0, 0.0, 43))
    assert (a[0] == b[0]).all() and a[0].shape == (4, 5, 1) and a[4].shape == (5, 4)
    assert (a[0] != c[0]).any()


def test_bleu_known_answers():
    # the worked examples of nltk.translate.bleu_score's documentation (Papineni et al. 2002)
    M = pkg.metrics
    hyp1 = "It is a guide to action which ensures that the military always obeys the commands of the party".split()
    hyp2 = "It is to insure the troops forever hearing the activity guidebook that party direct".split()
    ref1 = "It is a guide to action that ensures that the military will forever heed Party commands".split()
    ref2 = "It is the guiding principle which guarantees the military forces always being under the command of the Party".split()
    ref3 = "It is the practical guide for the army always to heed the directions of the party".split()
    assert abs(M.sentence_bleu([ref1, ref2, ref3], hyp1) - 0.5045666840058485) < 1e-12
    assert M.modified_precision([ref1, ref2, ref3], hyp1, 1) == (17, 18)
    assert M.modified_precision([ref1, ref2, ref3], hyp2, 1) == (8, 14)
    the = "the the the the the the the".split()
    assert M.modified_precision(["the cat is on the mat".split(), "there is a cat on the mat".split()], the, 1) == (2, 7)
    assert M.closest_ref_length([list(range(13)), list(range(15))], 14) == 13
    assert M.brevity_penalty(13, 12) == math.exp(1 - 13 / 12) and M.brevity_penalty(13, 14) == 1.0
    assert M.sentence_bleu([ref1], "completely unrelated words here".split()) == 0.0
    # method0: a missing higher-order match collapses the geometric mean to ~0, individual unigram score survives
    short = "It is a".split()
    s = M.compute_bleu(short, [ref1])
    assert len(s) == 8 and s[0] == s[4] > 0 and s[3] < 1e-100
    iw = {1: "<start>", 2: "a", 3: "dog", 4: "<end>"}
    assert M.ids_to_words([2, 3, 4, 2, 0], iw) == ["a", "dog"]
