"""Philox4x32-10 counter RNG in numpy (oracle side of the dropout-mask contract).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference draws dropout masks from TensorFlow's stateful RNG
(`tf.keras.layers.Dropout`, e.g. AttemptFour/Model/lc_NIC.py:51-55, layers.py:51,
attention.py:36, and the `dropout=` argument of the Keras LSTM at lc_NIC.py:118-124).
Those streams cannot be replayed, so parity-with-dropout is defined through a shared,
order-independent counter RNG instead: element `i` of dropout site `s` at decode step
`t` is kept iff  philox(counter=(i//4, s, t, 0), key=(seed_lo, seed_hi))[i % 4] >= thr,
thr = floor(rate * 2**32).  The CUDA kernels implement the same function
(csrc/philox.cuh); tests compare the two bit for bit.
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)

# dropout site ids (shared with csrc/philox.cuh)
SITE_INPUT = 1     # dropout_input on betas           index = b*N + v
SITE_FEAT = 2      # dropout_features after BN         index = (b*R + r)*G + g
SITE_TEXT = 3      # dropout_text on embeddings        index = (b*T + t)*E + e
SITE_ATTN = 4      # dropout_attn on tanh layer        index = (b*R + r)*A + a   (step t)
SITE_LSTM_IN = 5   # Keras LSTM(dropout=) input mask   index = b*(G+E) + j        (step t)
SITE_LSTM_OUT = 6  # dropout_lstm on h_t               index = b*U + u            (step t)
SITE_OUT = 7       # dropout_output on dense_inter     index = (b*T + t)*H + h
SITE_FEAT2 = 8     # ms2_NIC's second feature dropout  index = (b*R + r)*G + g
SITE_SAMPLE = 9    # random word choice (tnt_sample)   counter = (row b, site, step t, 0), first word of the draw


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  All inputs broadcastable uint32-valued arrays."""
    c0 = np.asarray(c0, dtype=np.uint64)
    c1 = np.asarray(c1, dtype=np.uint64) + np.zeros_like(c0)
    c2 = np.asarray(c2, dtype=np.uint64) + np.zeros_like(c0)
    c3 = np.asarray(c3, dtype=np.uint64) + np.zeros_like(c0)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        n0 = hi1 ^ c1 ^ np.uint64(k0)
        n2 = hi0 ^ c3 ^ np.uint64(k1)
        c0, c1, c2, c3 = n0, lo1, n2, lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def threshold(rate):
    """uint32 threshold: keep iff r >= thr.  rate in [0,1)."""
    # rate is rounded to fp32 first: the C ABI carries dropout rates as float (include/tnt.h)
    return int(np.floor(float(np.float32(rate)) * 4294967296.0)) & 0xFFFFFFFF if rate > 0 else 0


def random_u32(n, site, step, seed):
    """The first n uint32 draws of (site, step) under `seed` (linear element order)."""
    nblk = (n + 3) // 4
    blk = np.arange(nblk, dtype=np.uint64)
    r = philox4x32_10(blk, site, step, 0, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return np.stack(r, axis=1).reshape(-1)[:n]


def keep_mask(shape, rate, site, step, seed, dtype=np.float32):
    """Inverted-dropout multiplier: 0 or 1/(1-rate), Keras Dropout semantics."""
    n = int(np.prod(shape))
    if rate <= 0:
        return np.ones(shape, dtype=dtype)
    r = random_u32(n, site, step, seed)
    keep = r >= np.uint32(threshold(rate))
    scale = dtype(np.float32(1.0) / (np.float32(1.0) - np.float32(rate)))   # fp32, as the kernels compute it
    return (keep.astype(dtype) * scale).reshape(shape)


def sample_uniform(n_rows, step, seed):
    """u in (0,1) for rows 0..n_rows-1 of decode step `step`: ((first uint32 of the row's draw) + 0.5) / 2^32
    (csrc/sampling.cu: sample_kernel)."""
    r = philox4x32_10(np.arange(n_rows, dtype=np.uint64), SITE_SAMPLE, step, 0, seed & 0xFFFFFFFF,
                      (seed >> 32) & 0xFFFFFFFF)
    return (r[0].astype(np.float64) + 0.5) / 4294967296.0
