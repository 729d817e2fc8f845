"""Device-resident drop-ins for the reference's batch generators.

`DeviceDataGenerator` mirrors `DataGenerator(keras.utils.Sequence)` of
AttemptFour/DataLoaders/data_generator_guse.py:25-171 (`__len__`, `__getitem__`, `on_epoch_end`, the
`((betas, cap_vector, init_state, init_state), target[, keys])` batch tuple) and
`MultiSubjectDeviceDataGenerator` the two-subject one of data_generator_multisub.py:26-120 -- but a subject's
beta matrix (n_keys, 327 684) fp32 and the tokenised, padded caption table stay in HBM (13 GB per 10 000-image
subject; 180 GB per GPU), and a batch is assembled by two index-gather kernels of libtnt.so
(`tnt_assemble_betas`, `tnt_assemble_captions`, include/tnt.h) instead of 64 `np.load` calls, `pad_sequences`
and a 19 MB `to_categorical` on the host every step.  The tensors returned are CUDA tensors, which
`NIC.train_step / test_step / greedy_predict` take zero-copy.

Differences from the reference, on purpose:
  * the target defaults to token ids (B,T) int32 (what the kernels consume); `onehot=True` materialises the
    reference's (B,T,V) float32 one-hot on the device instead;
  * the pairs are index arrays into the tables (`pair_rows[i]` = beta row of pair i), not tuples of strings;
    shuffling permutes an index array with `np.random.shuffle` exactly like data_generator_guse.py:67-71 permutes
    `self.pairs` (same draws, same order for the same NumPy seed);
  * data parallel: rank r of `world` takes rows [r*B, (r+1)*B) of every world*B-row global batch; all
    ranks must shuffle identically (seed NumPy alike, or pass the same `rng`);
  * a returned batch lives in one of `slots` rotating device buffers: it stays valid until `slots - 1` further
    `__getitem__` calls (fit's one-batch lookahead needs 2).
There is no CPU path: the class needs libtnt.so and a CUDA device.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib


def pad_sequences(seqs, maxlen):
    """pad_sequences(seqs, maxlen, truncating='post', padding='post') of data_generator_guse.py:158: int32
    (n, maxlen); longer id lists keep their first maxlen entries, shorter ones are zero-filled at the end."""
    out = np.zeros((len(seqs), int(maxlen)), dtype=np.int32)
    for i, s in enumerate(seqs):
        a = np.asarray(s, dtype=np.int32).reshape(-1)[:maxlen]
        out[i, :a.size] = a
    return out


class BatchPlan:
    """Host-side index bookkeeping of one subject: which pairs form batch `index` on rank `rank` (pure NumPy)."""

    def __init__(self, n_pairs, batch_size, rank=0, world=1, shuffle=True, rng=None):
        if batch_size <= 0 or world <= 0 or not (0 <= rank < world):
            raise ValueError("bad batch_size / rank / world")
        self.order = np.arange(int(n_pairs), dtype=np.int64)
        self.batch_size, self.rank, self.world = int(batch_size), int(rank), int(world)
        self.shuffle, self.rng = bool(shuffle), rng

    def __len__(self):
        return len(self.order) // (self.batch_size * self.world)        # data_generator_guse.py:63-65

    def reshuffle(self):
        if self.shuffle:                                                 # data_generator_guse.py:67-71
            (self.rng if self.rng is not None else np.random).shuffle(self.order)

    def bounds(self, index):
        """[start, stop) into `order` of this rank's rows of global batch `index` (the batch one past the last full
        one is the ragged remainder eval.py:155 asks for; it may be empty)."""
        if index < 0:
            raise IndexError(index)
        lo = (index * self.world + self.rank) * self.batch_size
        lo = min(lo, len(self.order))
        return lo, min(lo + self.batch_size, len(self.order))

    def select(self, index):
        lo, hi = self.bounds(index)
        return self.order[lo:hi]


class _Subject:
    def __init__(self, betas, pair_rows, cap_table, device):
        t = betas if isinstance(betas, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(betas, dtype=np.float32))
        if t.dim() != 2:
            t = t.reshape(t.shape[0], -1)                               # image control: (n, 196, 512) -> (n, 100352)
        self.table = t.to(device=device, dtype=torch.float32).contiguous()
        self.pair_rows = np.asarray(pair_rows, dtype=np.int64).reshape(-1)
        if self.pair_rows.size and (self.pair_rows.min() < 0 or self.pair_rows.max() >= self.table.shape[0]):
            raise IndexError("pair_rows refers to a beta row outside the table")
        if cap_table.shape[0] != self.pair_rows.size:
            raise ValueError("one caption per pair expected")
        self.caps = torch.from_numpy(cap_table).to(device)
        self.rows_epoch = None      # (n_pairs,) int32 device: beta row of the pair at each position of `order`
        self.caps_epoch = None      # (n_pairs,) int32 device: caption row (= pair index) at each position


class _DeviceGeneratorBase:
    def __init__(self, subjects, batch_size, units, max_len, vocab_size, keys, shuffle, training, onehot, device, rank,
                 world, rng, slots):
        self._lib = _lib.load()                                          # raises if libtnt.so is missing: no fallback
        self.device = torch.device(device if device is not None else "cuda")
        if self.device.type != "cuda":
            raise RuntimeError("DeviceDataGenerator keeps its tables in HBM: a CUDA device is required")
        self.batch_size, self.units, self.max_len, self.vocab_size = int(batch_size), int(units), int(max_len), int(vocab_size)
        self.shuffle, self.training, self.onehot = bool(shuffle), bool(training), bool(onehot)
        self._subs = [_Subject(b, r, pad_sequences(c, max_len), self.device) for (b, r, c) in subjects]
        n = {s.pair_rows.size for s in self._subs}
        assert len(n) == 1, "Subjects need to have equal data split"     # data_generator_multisub.py:46
        w = {int(s.table.shape[1]) for s in self._subs}
        if len(w) != 1:
            raise ValueError("all subjects must share one input width")
        self.width = w.pop()
        self._plans = [BatchPlan(s.pair_rows.size, batch_size, rank, world, shuffle, rng) for s in self._subs]
        self.keys = [np.asarray(k) for k in keys] if keys is not None else [s.pair_rows for s in self._subs]
        rows = self.batch_size * len(self._subs)
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=self.device)
        self._slots = [dict(betas=z((rows, self.width), torch.float32), cap=z((rows, self.max_len), torch.int32),
                            tgt=z((rows, self.max_len), torch.int32),
                            onehot=z((rows, self.max_len, self.vocab_size), torch.float32) if self.onehot else None)
                       for _ in range(max(int(slots), 2))]
        self._slot = 0
        self._state = z((rows, self.units), torch.float32)               # init_state, data_generator_guse.py:166
        self._bad = z((1,), torch.int32)
        self.on_epoch_end()

    def __len__(self):
        """Nr. of batches per epoch (per rank)."""
        return len(self._plans[0])

    def on_epoch_end(self):
        """Shuffle when an epoch ends (data_generator_guse.py:67-71; each subject on its own,
        data_generator_multisub.py:66-71), then refresh the device-side index arrays (2 x 4 bytes per pair)."""
        self.check()
        for s, p in zip(self._subs, self._plans):
            p.reshuffle()
            s.rows_epoch = torch.from_numpy(s.pair_rows[p.order].astype(np.int32)).to(self.device)
            s.caps_epoch = torch.from_numpy(p.order.astype(np.int32)).to(self.device)

    def check(self):
        """Raises if any batch assembled so far addressed a row outside its table (synchronises)."""
        if int(self._bad.item()) != 0:
            self._bad.zero_()
            raise IndexError("batch assembly addressed a row outside its table, or a token id outside the vocabulary")

    def _chk(self, rc):
        if rc != 0:
            raise _lib.TntError(f"libtnt error {rc}: {self._lib.tnt_data_last_error().decode()}")

    def __getitem__(self, index):
        """One batch: ((betas (B,N), cap_vector (B,T) i32, init_state, init_state), target[, keys])."""
        lib = self._lib
        slot = self._slots[self._slot]
        self._slot = (self._slot + 1) % len(self._slots)
        st = torch.cuda.current_stream(self.device).cuda_stream
        T, V, Bs = self.max_len, self.vocab_size, self.batch_size
        lo, hi = self._plans[0].bounds(index)
        n = hi - lo
        keys = []
        with torch.cuda.device(self.device):
            for k, (s, p) in enumerate(zip(self._subs, self._plans)):
                o = k * n                                                # subject k's rows sit behind subject k-1's
                if n > 0:
                    self._chk(lib.tnt_assemble_betas(s.table.data_ptr(), s.table.shape[0], self.width,
                                                     s.rows_epoch.data_ptr() + 4 * lo, n,
                                                     slot["betas"].data_ptr() + 4 * o * self.width,
                                                     self._bad.data_ptr(), st))
                    self._chk(lib.tnt_assemble_captions(s.caps.data_ptr(), s.caps.shape[0], T,
                                                        s.caps_epoch.data_ptr() + 4 * lo, n,
                                                        slot["cap"].data_ptr() + 4 * o * T, slot["tgt"].data_ptr() + 4 * o * T,
                                                        (slot["onehot"].data_ptr() + 4 * o * T * V) if self.onehot else None,
                                                        V, self._bad.data_ptr(), st))
                if not self.training:
                    keys.append(self.keys[k][p.order[lo:hi]])
        rows = n * len(self._subs)
        target = slot["onehot"][:rows] if self.onehot else slot["tgt"][:rows]
        state = self._state[:rows]
        data = (slot["betas"][:rows], slot["cap"][:rows], state, state)
        if self.training:
            return (data, target)
        return (data, target, np.concatenate(keys) if keys else np.zeros(0, np.int64))

    def launch_count(self):
        return int(self._lib.tnt_data_launch_count())


class DeviceDataGenerator(_DeviceGeneratorBase):
    """data_generator_guse.DataGenerator with the tables in HBM.
    betas (n_keys, N) float32 (NumPy or torch, host or device; uploaded once); pair_rows[i] = beta row of pair i;
    cap_seqs = `tokenizer.texts_to_sequences(captions)` (one id list per pair); keys[i] = what is returned third
    when training=False (the NSD key of pair i, data_generator_guse.py:171)."""

    def __init__(self, betas, pair_rows, cap_seqs, batch_size, units, max_len, vocab_size, keys=None, shuffle=True,
                 training=False, *, onehot=False, device=None, rank=0, world=1, rng=None, slots=3):
        super().__init__([(betas, pair_rows, cap_seqs)], batch_size, units, max_len, vocab_size,
                         [keys] if keys is not None else None, shuffle, training, onehot, device, rank, world, rng, slots)


class MultiSubjectDeviceDataGenerator(_DeviceGeneratorBase):
    """data_generator_multisub.DataGenerator: batch = batch_size rows of subject A followed by batch_size rows of
    subject B (the halves ms2_NIC.call slices, ms2_NIC.py:181-188); every argument is a pair (A, B)."""

    def __init__(self, betas_ab, pair_rows_ab, cap_seqs_ab, batch_size, units, max_len, vocab_size, keys_ab=None,
                 shuffle=True, training=False, *, onehot=False, device=None, rank=0, world=1, rng=None, slots=3):
        super().__init__(list(zip(betas_ab, pair_rows_ab, cap_seqs_ab)), batch_size, units, max_len, vocab_size,
                         keys_ab, shuffle, training, onehot, device, rank, world, rng, slots)
