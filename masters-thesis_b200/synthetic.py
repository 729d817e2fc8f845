"""Synthetic NSD-shaped inputs for the hot path.

The real inputs are external files the reference loads from hard-coded paths
(AttemptFour/DataLoaders/load_avg_betas.py:35-40, data_generator_guse.py:141-155); this module
produces data of the same shape and contract:

  * parcel groups  -- what `loader.select_groups(group_size, remove=[...])` returns
                      (load_avg_betas.py:96-114, main.py:115): a list of int index arrays into the
                      327 684-vertex beta vector plus `[group_size] * R`.
  * betas          -- (B, 327684) float32, z-scored per vertex upstream => standard normal.
  * captions       -- `pad_sequences(texts_to_sequences(cap), maxlen=15, truncating='post',
                      padding='post')` (data_generator_guse.py:157-158): <start> w1..wL <end> 0 0 ...
  * target         -- caption shifted left by one with trailing 0 (data_generator_guse.py:160-162);
                      `to_categorical` gives the one-hot the reference's train_step receives.
  * states         -- zeros (B, units) (data_generator_guse.py:166).
"""
import numpy as np

N_VERTICES = 327684            # AttemptFour/config.yaml:46
REMOVED_PARCELS = [142, 17, 133, 315, 1, 197, 158, 192, 135, 153, 137, 140, 92, 183, 125]  # main.py:115
START_ID, END_ID, FIRST_WORD = 2, 3, 4


def make_groups(n_vertices=N_VERTICES, parcels_per_hemi=180, unlabeled_frac=0.085, min_size=64,
                max_block=64, seed=42, remove=None, hemi_offset=True):
    """Glasser-like parcellation: two hemispheres, each vertex labelled 0 (dropped, mirrors
    `groups_lh[1:]`, load_avg_betas.py:81) or one of `parcels_per_hemi` parcels with log-normal sizes,
    labels laid out over the vertex order in contiguous patches of 1..max_block vertices.

    hemi_offset=False reproduces the reference's indexing literally: both hemispheres' index arrays
    are built from `range(len(glasser_xh))` (load_avg_betas.py:64-72) and therefore BOTH address the
    first half of the beta vector (overlapping groups).  The default offsets the second hemisphere
    so every labelled vertex is read once (the intended, more bandwidth-hungry case).
    """
    rng = np.random.default_rng(seed)
    half = n_vertices // 2
    groups = []
    for hemi in range(2):
        n_lab = int(round(half * (1.0 - unlabeled_frac)))
        w = rng.lognormal(0.0, 0.8, parcels_per_hemi)
        sizes = np.maximum(min_size, np.floor(w / w.sum() * n_lab)).astype(np.int64)
        # trim/extend the largest parcel so the sizes sum to n_lab
        sizes[np.argmax(sizes)] += n_lab - sizes.sum()
        assert sizes.min() > 0 and sizes.sum() == n_lab
        # lay labels out in patches
        labels = np.zeros(half, dtype=np.int32)
        remaining = sizes.copy()
        remaining0 = half - n_lab
        pos = 0
        weights_lab = np.concatenate([[remaining0], remaining]).astype(np.float64)
        while pos < half:
            tot = weights_lab.sum()
            lab = rng.choice(len(weights_lab), p=weights_lab / tot)
            blk = int(min(rng.integers(1, max_block + 1), weights_lab[lab], half - pos))
            labels[pos:pos + blk] = lab
            weights_lab[lab] -= blk
            pos += blk
        base = hemi * half if hemi_offset else 0
        idx_all = np.arange(half, dtype=np.int32)
        for p in range(1, parcels_per_hemi + 1):
            groups.append((idx_all[labels == p] + base).astype(np.int32))
    if remove:
        rm = set(remove)
        groups = [g for i, g in enumerate(groups) if i not in rm]
    return groups


def select_groups(out_dim, remove=REMOVED_PARCELS, **kw):
    """Drop-in for loader.select_groups (load_avg_betas.py:102-114)."""
    g = make_groups(remove=remove, **kw)
    return g, [out_dim] * len(g)


def small_groups(n_vertices, n_groups, seed=0, overlap=False, min_size=3):
    """Ragged little parcellation for tests (sizes not multiples of anything)."""
    rng = np.random.default_rng(seed)
    perm = rng.permutation(n_vertices).astype(np.int32)
    n_used = int(n_vertices * 0.9)
    cuts = np.sort(rng.choice(np.arange(min_size, n_used - min_size), n_groups - 1, replace=False))
    parts = np.split(perm[:n_used], cuts)
    groups = [np.sort(p).astype(np.int32) for p in parts]
    groups = [g if len(g) >= 1 else perm[:min_size].copy() for g in groups]
    if overlap:
        groups[-1] = np.sort(np.concatenate([groups[-1], groups[0][: max(1, len(groups[0]) // 2)]])).astype(np.int32)
    return groups


def make_captions(n, T=15, vocab_size=5001, seed=42):
    """Token ids (n,T) int32: <start>, L~clip(Poisson(10.5),5,13) Zipf(1.1)-distributed words, <end>, pads."""
    rng = np.random.default_rng(seed + 1)
    cap = np.zeros((n, T), dtype=np.int32)
    n_words = max(1, vocab_size - FIRST_WORD)
    ranks = np.arange(1, n_words + 1, dtype=np.float64)
    p = ranks ** -1.1
    p /= p.sum()
    L = np.clip(rng.poisson(10.5, n), 5, 13)
    for i in range(n):
        seq = [START_ID] + list(FIRST_WORD + rng.choice(n_words, size=L[i], p=p)) + [END_ID]
        seq = seq[:T]
        cap[i, :len(seq)] = seq
    return np.minimum(cap, vocab_size - 1).astype(np.int32)


def make_target_ids(cap):
    """Shift left by one, trailing 0 (data_generator_guse.py:160-162)."""
    tgt = np.zeros_like(cap)
    tgt[:, :-1] = cap[:, 1:]
    return tgt


def to_categorical(ids, vocab_size):
    out = np.zeros(ids.shape + (vocab_size,), dtype=np.float32)
    np.put_along_axis(out, ids[..., None].astype(np.int64), 1.0, axis=-1)
    return out


def make_betas(n, n_vertices=N_VERTICES, seed=42):
    rng = np.random.default_rng(seed + 2)
    return rng.standard_normal((n, n_vertices), dtype=np.float32)


def make_batch(B, T=15, vocab_size=5001, units=512, n_vertices=N_VERTICES, seed=42, onehot=False):
    """((betas, cap, a0, c0), target) exactly as DataGenerator.__getitem__ returns it
    (data_generator_guse.py:168-169); target is ids unless onehot=True."""
    betas = make_betas(B, n_vertices, seed)
    cap = make_captions(B, T, vocab_size, seed)
    tgt = make_target_ids(cap)
    z = np.zeros((B, units), dtype=np.float32)
    return (betas, cap, z, z.copy()), (to_categorical(tgt, vocab_size) if onehot else tgt)
