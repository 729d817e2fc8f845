"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's batch generators.

Follows AttemptFour/DataLoaders/data_generator_guse.py:63-71 (`__len__`, `on_epoch_end`), :123-171
(`__getitem__`, `__data_generation`) and data_generator_multisub.py:62-120 (two subjects: A's rows first, B's
behind them; each subject's pair list shuffled on its own), with the per-sample `np.load(".../subjXX_KID{key}.npy")`
replaced by a row lookup in an in-memory beta matrix (what `pre_load_betas` / `preload_data`, :45-61, set out to do).
Keras helpers restated from their published behaviour (tensorflow/keras is not installable here):
  pad_sequences(seqs, maxlen, truncating='post', padding='post') -> int32 (len(seqs), maxlen): keep the FIRST maxlen
      ids of a longer sequence, zero-fill behind a shorter one;
  to_categorical(y, num_classes) -> float32 one-hot with a new last axis.
Parity unpinned against TF itself (see oracle/keras_semantics.py header); only tests/, smoke() and bench.py's CPU leg
may import this module.
"""
import numpy as np


def pad_sequences(seqs, maxlen):
    """data_generator_guse.py:158: pad_sequences(cap_seqs, maxlen=max_len, truncating='post', padding='post')."""
    out = np.zeros((len(seqs), maxlen), dtype=np.int32)
    for i, s in enumerate(seqs):
        s = list(s)[:maxlen]
        out[i, :len(s)] = s
    return out


def to_categorical(y, num_classes):
    y = np.asarray(y)
    out = np.zeros(y.shape + (num_classes,), dtype=np.float32)
    np.put_along_axis(out, y[..., None].astype(np.int64), 1.0, axis=-1)
    return out


class DataGenerator:
    """data_generator_guse.DataGenerator on in-memory tables.
    betas (n_keys, N) f32; pair_rows (n_pairs,) = row of `betas` each (key, caption) pair reads; cap_seqs = the
    tokenizer's id lists, one per pair; keys (n_pairs,) = what the generator returns third when training=False."""

    def __init__(self, betas, pair_rows, cap_seqs, batch_size, units, max_len, vocab_size, keys=None, shuffle=True,
                 training=False):
        self.betas = np.asarray(betas, dtype=np.float32)
        self.pair_rows = np.asarray(pair_rows, dtype=np.int64)
        self.caps = pad_sequences(cap_seqs, max_len)
        self.keys = np.asarray(keys) if keys is not None else self.pair_rows.copy()
        self.order = np.arange(len(self.pair_rows))
        self.batch_size, self.units, self.max_len, self.vocab_size = batch_size, units, max_len, vocab_size
        self.shuffle, self.training = shuffle, training
        self.on_epoch_end()

    def __len__(self):
        return len(self.order) // self.batch_size                       # :63-65

    def on_epoch_end(self):
        if self.shuffle:
            np.random.shuffle(self.order)                                # :67-71 (same draws as shuffling the pairs)

    def __getitem__(self, index):
        sel = self.order[index * self.batch_size:(index + 1) * self.batch_size]   # :123-127
        return self._data_generation(sel)

    def _data_generation(self, sel):
        betas_batch = self.betas[self.pair_rows[sel]]                    # :138-153
        cap_vector = self.caps[sel]                                      # :156-158
        target = np.zeros_like(cap_vector)                               # :160-163
        target[:, :-1] = cap_vector[:, 1:]
        target = to_categorical(target, self.vocab_size)
        init_state = np.zeros((len(sel), self.units), np.float32)        # :166
        if self.training:
            return ((betas_batch, cap_vector, init_state, init_state), target)
        return ((betas_batch, cap_vector, init_state, init_state), target, self.keys[sel])


class MultiSubjectDataGenerator:
    """data_generator_multisub.DataGenerator: two subjects with equal pair counts, a batch is batch_size rows of A
    followed by batch_size rows of B (2*batch_size rows)."""

    def __init__(self, betas_ab, pair_rows_ab, cap_seqs_ab, batch_size, units, max_len, vocab_size, shuffle=True,
                 training=False):
        self.betas = [np.asarray(b, dtype=np.float32) for b in betas_ab]
        self.pair_rows = [np.asarray(r, dtype=np.int64) for r in pair_rows_ab]
        self.caps = [pad_sequences(c, max_len) for c in cap_seqs_ab]
        assert len(self.pair_rows[0]) == len(self.pair_rows[1]), "Subjects need to have equal data split"   # :46
        self.order = [np.arange(len(r)) for r in self.pair_rows]
        self.batch_size, self.units, self.max_len, self.vocab_size = batch_size, units, max_len, vocab_size
        self.shuffle, self.training = shuffle, training
        self.on_epoch_end()

    def __len__(self):
        return len(self.order[0]) // self.batch_size                    # :62-64

    def on_epoch_end(self):
        if self.shuffle:                                                 # :66-71
            np.random.shuffle(self.order[0])
            np.random.shuffle(self.order[1])

    def __getitem__(self, index):
        s = slice(index * self.batch_size, (index + 1) * self.batch_size)   # :73-78
        sel = [self.order[0][s], self.order[1][s]]
        betas_batch = np.concatenate([self.betas[k][self.pair_rows[k][sel[k]]] for k in (0, 1)], axis=0)   # :93-103
        cap_vector = np.concatenate([self.caps[k][sel[k]] for k in (0, 1)], axis=0)                        # :90,106-107
        target = np.zeros_like(cap_vector)
        target[:, :-1] = cap_vector[:, 1:]
        target = to_categorical(target, self.vocab_size)
        init_state = np.zeros((cap_vector.shape[0], self.units), np.float32)
        keys = np.concatenate([self.pair_rows[k][sel[k]] for k in (0, 1)])
        if self.training:
            return ((betas_batch, cap_vector, init_state, init_state), target)
        return ((betas_batch, cap_vector, init_state, init_state), target, keys)
