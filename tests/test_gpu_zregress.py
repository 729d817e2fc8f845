"""GPU (-m gpu) regression tests for faults found on the box."""
import numpy as np
import pytest

from oracle import keras_semantics as K
from tests.util import pkg, syn, make_model

pytestmark = pytest.mark.gpu


def test_train_step_after_a_larger_context_was_created():
    """The packed-operand arena is shared by all contexts of a process; creating a larger-batch model (bench.py's
    515-caption greedy model) re-reserves it.  Its capacity used to be an arbitrary byte count, which misaligned
    every side-stream operand (allocated downwards from the end) of the next train step: 'misaligned address'."""
    cfg = K.Config().no_dropout()
    g, _ = syn.select_groups(cfg.G)
    groups = [g]
    P = K.init_params(cfg, groups, seed=3)
    (x, cap, a0, c0), tgt = syn.make_batch(64, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
    data = ((x, cap, a0, c0), tgt)
    a = make_model(cfg, groups, P, batch_max=64)
    twin = make_model(cfg, groups, P, batch_max=64)
    first = a.train_step(data, seed=1)
    assert first == pytest.approx(twin.train_step(data, seed=1), rel=1e-6)
    big = make_model(cfg, groups, P, batch_max=515)                # forces a larger arena
    start = np.full(515, syn.START_ID, np.int32)
    xb = np.tile(x, (9, 1))[:515]
    z = np.zeros((515, cfg.U), np.float32)
    ids = big.greedy_predict(xb, z, z, start, cfg.T, cfg.U, None, return_raw=False, return_attention=False)[0]
    assert ids.shape == (515, cfg.T, 1)
    second = a.train_step(data, seed=2)                            # used to fault here
    want = twin.train_step(data, seed=2)
    for k in want:
        assert abs(second[k] - want[k]) <= 1e-6 * max(abs(want[k]), 1e-6), (k, second[k], want[k])
