"""CPU tests of the batch generators: the oracle restatement of
data_generator_guse.py / data_generator_multisub.py on hand-made cases, and the product's host-side logic
(`pad_sequences`, `BatchPlan`: batch bounds, data-parallel sharding, shuffle order) against it.  No GPU calls."""
import importlib

import numpy as np
import pytest

from oracle import data_generator as O

ds = importlib.import_module("masters-thesis_b200.dataset")


def test_pad_sequences_matches_oracle_and_keras_contract():
    seqs = [[2, 7, 9, 3], [], [2] + list(range(4, 30)) + [3], [5], list(range(1, 16))]
    want = O.pad_sequences(seqs, 15)
    got = ds.pad_sequences(seqs, 15)
    assert got.dtype == np.int32 and got.shape == (5, 15)
    assert (got == want).all()
    assert list(got[0]) == [2, 7, 9, 3] + [0] * 11                     # padding='post'
    assert list(got[2]) == [2] + list(range(4, 18))                   # truncating='post': the first 15 ids stay
    assert (got[1] == 0).all()


def test_oracle_generator_known_answer():
    betas = np.arange(5 * 4, dtype=np.float32).reshape(5, 4)
    rows = [3, 0, 4, 1, 1, 2]
    caps = [[2, 5, 6, 3], [2, 7, 3], [2, 8, 9, 9, 3], [2, 3], [2, 6, 3], [2, 5, 3]]
    g = O.DataGenerator(betas, rows, caps, batch_size=2, units=3, max_len=4, vocab_size=10, keys=list("abcdef"),
                        shuffle=False, training=False)
    assert len(g) == 3
    (x, cap, a0, c0), tgt, keys = g[1]
    assert (x == betas[[4, 1]]).all()
    assert cap.tolist() == [[2, 8, 9, 9], [2, 3, 0, 0]]               # truncated / zero padded
    assert tgt.shape == (2, 4, 10) and tgt.dtype == np.float32
    assert tgt.argmax(-1).tolist() == [[8, 9, 9, 0], [3, 0, 0, 0]]    # shifted left, trailing 0 = class 0
    assert (tgt.sum(-1) == 1).all()
    assert a0.shape == (2, 3) and not a0.any() and list(keys) == ["c", "d"]
    assert len(O.DataGenerator(betas, rows, caps, 2, 3, 4, 10, training=True)[0]) == 2


def test_oracle_multisubject_layout():
    bA = np.full((3, 2), 1.0, np.float32) * np.arange(3)[:, None]
    bB = 10 + bA
    g = O.MultiSubjectDataGenerator((bA, bB), ([0, 1, 2, 0], [2, 1, 0, 2]), ([[2, 3]] * 4, [[2, 4, 3]] * 4),
                                    batch_size=2, units=3, max_len=3, vocab_size=6, shuffle=False, training=True)
    (x, cap, a0, c0), tgt = g[1]
    assert x[:, 0].tolist() == [2.0, 0.0, 10.0, 12.0]                 # subject A's rows first, then subject B's
    assert cap.tolist() == [[2, 3, 0]] * 2 + [[2, 4, 3]] * 2
    assert a0.shape == (4, 3)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_batch_plan_shards_the_global_batch(world):
    n, B = 103, 5
    plans = [ds.BatchPlan(n, B, r, world, shuffle=False) for r in range(world)]
    assert all(len(p) == n // (B * world) for p in plans)
    for i in range(len(plans[0])):
        got = np.concatenate([p.select(i) for p in plans])
        assert (got == np.arange(i * B * world, (i + 1) * B * world)).all()
    # the batch one past the last full one is the (possibly empty) remainder
    lo, hi = plans[0].bounds(len(plans[0]))
    assert 0 <= hi - lo <= B and hi <= n
    assert plans[0].bounds(10 ** 6) == (n, n)
    with pytest.raises(IndexError):
        plans[0].bounds(-1)


def test_batch_plan_shuffle_equals_reference_shuffle_of_the_pairs():
    # np.random.shuffle of the index array draws exactly what data_generator_guse.py:67-71 draws on `self.pairs`
    pairs = np.array([[f"k{i}", f"cap {i}"] for i in range(50)])
    np.random.seed(7)
    ref = pairs.copy()
    np.random.shuffle(ref)
    np.random.shuffle(ref)
    np.random.seed(7)
    p = ds.BatchPlan(50, 8)
    p.reshuffle(); p.reshuffle()
    assert (pairs[p.order] == ref).all()
    np.random.seed(7)
    o = O.DataGenerator(np.zeros((1, 1), np.float32), [0] * 50, [[2, 3]] * 50, 8, 2, 3, 5, shuffle=True)
    o.on_epoch_end()
    assert (o.order == p.order).all()
    q = ds.BatchPlan(50, 8, rng=np.random.RandomState(3)); q.reshuffle()
    r = ds.BatchPlan(50, 8, rng=np.random.RandomState(3)); r.reshuffle()
    assert (q.order == r.order).all() and sorted(q.order) == list(range(50))


def test_device_generator_refuses_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises((RuntimeError, AssertionError)):
        ds.DeviceDataGenerator(np.zeros((2, 8), np.float32), [0, 1], [[2, 3], [2, 3]], 1, 4, 3, 5, device="cpu")
