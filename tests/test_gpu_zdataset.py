"""GPU tests (-m gpu) of batch assembly from HBM-resident tables: `DeviceDataGenerator` /
`MultiSubjectDeviceDataGenerator` through the C ABI (tnt_assemble_betas / tnt_assemble_captions) against the numpy
restatement of the reference generators.  Index and byte work: BIT-EXACT."""
import importlib

import numpy as np
import pytest
import torch

from oracle import data_generator as O
from oracle import keras_semantics as K
from tests.util import pkg, syn, make_model

ds = importlib.import_module("masters-thesis_b200.dataset")
pytestmark = pytest.mark.gpu


def _tables(n_keys, N, n_pairs, T, V, seed=0):
    rng = np.random.default_rng(seed)
    betas = rng.standard_normal((n_keys, N), dtype=np.float32)
    rows = rng.integers(0, n_keys, n_pairs)
    caps = [[syn.START_ID] + list(rng.integers(4, V, rng.integers(0, T + 4))) + [syn.END_ID] for _ in range(n_pairs)]
    keys = np.array([f"KID{r}" for r in rows])
    return betas, rows, caps, keys


def _same(batch_d, batch_o, onehot):
    (x, cap, a0, c0), tgt = batch_d[0], batch_d[1]
    (xo, capo, a0o, c0o), tgto = batch_o[0], batch_o[1]
    assert x.is_cuda and cap.dtype == torch.int32
    assert np.array_equal(x.cpu().numpy(), xo)
    assert np.array_equal(cap.cpu().numpy(), capo)
    assert np.array_equal(a0.cpu().numpy(), a0o) and np.array_equal(c0.cpu().numpy(), c0o)
    if onehot:
        assert tgt.dtype == torch.float32 and np.array_equal(tgt.cpu().numpy(), tgto)
    else:
        assert np.array_equal(tgt.cpu().numpy(), tgto.argmax(-1))


@pytest.mark.parametrize("N,onehot", [(203, True), (200, False), (4096 * 3 + 4, True)])
def test_batches_bit_exact_ragged_widths_shuffled(N, onehot):
    # N = 203: scalar path (width not a multiple of 4); 200 / 12292: float4 path incl. a partial last chunk
    T, V, B = 7, 37, 5
    betas, rows, caps, keys = _tables(17, N, 43, T, V, seed=N)
    np.random.seed(123)
    o = O.DataGenerator(betas, rows, caps, B, 16, T, V, keys=keys, shuffle=True, training=False)
    np.random.seed(123)
    d = ds.DeviceDataGenerator(betas, rows, caps, B, 16, T, V, keys=keys, shuffle=True, training=False, onehot=onehot)
    assert len(d) == len(o) == 8
    for epoch in range(2):
        for i in range(len(o) + 1):                     # eval.py:155 also asks for the ragged remainder
            bd, bo = d[i], o[i]
            _same(bd, bo, onehot)
            assert list(bd[2]) == list(bo[2])
        np.random.seed(1000 + epoch); o.on_epoch_end()      # both draw from NumPy's global stream: reseed alike
        np.random.seed(1000 + epoch); d.on_epoch_end()
    e = d[len(o) + 5]                                   # past the end: empty batch, nothing launched
    assert e[0][0].shape == (0, N) and e[1].shape[0] == 0
    d.check()
    assert d.launch_count() > 0


def test_full_width_batch_bit_exact_and_bandwidth_shape():
    N, B, T, V = syn.N_VERTICES, 64, 15, 5001
    n_keys = 160                                         # 210 MB table
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    table = torch.randn(n_keys, N, device="cuda", generator=g)
    rng = np.random.default_rng(2)
    rows = rng.integers(0, n_keys, 640)
    caps = syn.make_captions(640, T, V, seed=3)
    d = ds.DeviceDataGenerator(table, rows, [list(c) for c in caps], B, 512, T, V, shuffle=True, training=True,
                               rng=np.random.RandomState(5))
    order = d._plans[0].order.copy()
    for i in (0, 3, len(d) - 1):
        (x, cap, a0, c0), tgt = d[i]
        sel = order[i * B:(i + 1) * B]
        want = table[torch.from_numpy(rows[sel]).cuda()]
        assert torch.equal(x, want)                      # size-independent property: gather == index_select
        assert np.array_equal(cap.cpu().numpy(), caps[sel])
        assert np.array_equal(tgt.cpu().numpy(), syn.make_target_ids(caps[sel]))
        assert x.shape == (B, N) and a0.shape == (B, 512)
    # three rotating slots: a batch stays valid while the next two are assembled
    b0 = d[0][0][0]; keep = b0.clone(); d[1]; d[2]
    assert torch.equal(b0, keep)
    d.check()


def test_multisubject_halves_and_data_parallel_shards():
    N, T, V, B = 64, 6, 29, 4
    tA = _tables(9, N, 40, T, V, seed=1)
    tB = _tables(11, N, 40, T, V, seed=2)
    np.random.seed(9)
    o = O.MultiSubjectDataGenerator((tA[0], tB[0]), (tA[1], tB[1]), (tA[2], tB[2]), B, 8, T, V, shuffle=True, training=True)
    np.random.seed(9)
    d = ds.MultiSubjectDeviceDataGenerator((tA[0], tB[0]), (tA[1], tB[1]), (tA[2], tB[2]), B, 8, T, V, shuffle=True,
                                           training=True, onehot=True)
    assert len(d) == len(o) == 10
    for i in range(len(o)):
        _same(d[i], o[i], True)
        assert d[i][0][0].shape == (2 * B, N)
    # world 2: rank r's batch i = rows [r*B, (r+1)*B) of each subject's global batch i (batch_size 2B in the oracle)
    og = O.MultiSubjectDataGenerator((tA[0], tB[0]), (tA[1], tB[1]), (tA[2], tB[2]), 2 * B, 8, T, V, shuffle=False, training=True)
    for r in range(2):
        dr = ds.MultiSubjectDeviceDataGenerator((tA[0], tB[0]), (tA[1], tB[1]), (tA[2], tB[2]), B, 8, T, V, shuffle=False,
                                                training=True, rank=r, world=2)
        assert len(dr) == len(og) == 5
        (x, cap, _, _), tgt = dr[2]
        (xo, capo, _, _), tgto = og[2]
        pick = np.r_[r * B:(r + 1) * B, 2 * B + r * B:2 * B + (r + 1) * B]
        assert np.array_equal(x.cpu().numpy(), xo[pick]) and np.array_equal(cap.cpu().numpy(), capo[pick])
        assert np.array_equal(tgt.cpu().numpy(), tgto.argmax(-1)[pick])


def test_bad_indices_are_flagged_not_read():
    betas, rows, caps, keys = _tables(5, 32, 12, 5, 11)
    with pytest.raises(IndexError):
        ds.DeviceDataGenerator(betas, np.r_[rows[:-1], 5], caps, 4, 8, 5, 11)
    d = ds.DeviceDataGenerator(betas, rows, caps, 4, 8, 5, 11, shuffle=False, onehot=True)
    d._subs[0].rows_epoch[1] = 99                        # corrupt the device-side index behind the host check
    x = d[0][0][0]
    assert not x[1].any() and x[0].any()                 # the row is zero-filled, its neighbours are intact
    with pytest.raises(IndexError):
        d.check()
    d2 = ds.DeviceDataGenerator(betas, rows, [[2, 10, 11, 3]] * 12, 4, 8, 5, 11, shuffle=False, onehot=True)
    d2[0]
    with pytest.raises(IndexError):                      # token id 11 has no one-hot column at vocab_size 11
        d2.check()


def test_fit_from_device_generator_equals_host_batches():
    cfg = K.tiny_config()
    groups = [syn.small_groups(cfg.N, cfg.R, seed=1, overlap=True)]
    P = K.init_params(cfg, groups, seed=3)
    betas, rows, caps, keys = _tables(13, cfg.N, 24, cfg.T, cfg.V, seed=4)
    B = 4
    host = O.DataGenerator(betas, rows, caps, B, cfg.U, cfg.T, cfg.V, shuffle=False, training=True)
    logs_h = []
    m1 = make_model(cfg, groups, P, batch_max=B)
    for i in range(len(host)):
        logs_h.append(m1.train_step(host[i], seed=100 + i))
    m2 = make_model(cfg, groups, P, batch_max=B)
    dev = ds.DeviceDataGenerator(betas, rows, caps, B, cfg.U, cfg.T, cfg.V, shuffle=False, training=True)
    for i in range(len(dev)):
        got = m2.train_step(dev[i], seed=100 + i)
        for k, v in logs_h[i].items():
            assert abs(got[k] - v) <= 1e-6 * max(abs(v), 1e-6), (i, k, got[k], v)
    # and through NIC.fit (prefetch of batch i+1 while batch i computes): same weights as the plain loop
    m3 = make_model(cfg, groups, P, batch_max=B)
    m3._seed, m3._step = m2._seed, 0
    m4 = make_model(cfg, groups, P, batch_max=B)
    dev2 = ds.DeviceDataGenerator(betas, rows, caps, B, cfg.U, cfg.T, cfg.V, shuffle=False, training=True, onehot=True)
    m3.fit(dev2, epochs=1)
    for i in range(len(host)):
        m4.train_step(host[i])
    for a, b in zip(m3.get_weights(), m4.get_weights()):
        assert np.allclose(a, b, rtol=0, atol=1e-7)
