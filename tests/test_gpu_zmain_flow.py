"""GPU (-m gpu): the reference's driver flow end to end on the drop-in surface at a small size --
AttemptFour/main.py:113-134 (construct, compile), :147 (build call), :221-281 (callbacks + fit with validation),
:172-190 (weights-only checkpoints), then eval.py:140-193 (load_weights + greedy loop over the test generator)."""
import importlib

import numpy as np
import pytest

from oracle import keras_semantics as K
from tests.util import pkg, syn

ds = importlib.import_module("masters-thesis_b200.dataset")
pytestmark = pytest.mark.gpu


def _new_model(cfg, groups, B):
    m = pkg.NIC((groups, [cfg.G] * cfg.R), cfg.U, 512, cfg.E, cfg.A, cfg.V, cfg.T,
                cfg.dropout_input, cfg.dropout_features, cfg.dropout_text, cfg.dropout_attn, cfg.dropout_lstm,
                cfg.dropout_out, cfg.input_reg, cfg.attn_reg, cfg.lstm_reg, cfg.output_reg,
                inter_units=cfg.H, batch_max=B, seed=42)
    m.compile(pkg.Adam(learning_rate=0.0001, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1), loss=None,
              run_eagerly=True)                                                        # main.py:97,134
    return m


def test_main_py_and_eval_py_flow(tmp_path):
    cfg = K.tiny_config()
    groups = syn.small_groups(cfg.N, cfg.R, seed=1, overlap=True)
    B, rng = 4, np.random.default_rng(0)
    betas = rng.standard_normal((30, cfg.N), dtype=np.float32)

    def pairs(n, seed):
        r = np.random.default_rng(seed)
        return (r.integers(0, 30, n), [[syn.START_ID] + list(map(int, r.integers(4, cfg.V, r.integers(1, 6)))) + [syn.END_ID]
                                       for _ in range(n)])
    (rt, ct), (rv, cv), (re, ce) = pairs(26, 1), pairs(9, 2), pairs(11, 3)
    np.random.seed(42)                                                                # main.py seeds NumPy from config.yaml:24
    train_gen = ds.DeviceDataGenerator(betas, rt, ct, B, cfg.U, cfg.T, cfg.V, shuffle=True, training=True, onehot=True)
    val_gen = ds.DeviceDataGenerator(betas, rv, cv, B, cfg.U, cfg.T, cfg.V, shuffle=False, training=True, onehot=True)
    test_gen = ds.DeviceDataGenerator(betas, re, ce, B, cfg.U, cfg.T, cfg.V, keys=[f"KID{k}" for k in re],
                                      shuffle=False, training=False)

    model = _new_model(cfg, groups, B)
    probs, attn = model(train_gen[0][0])                                              # main.py:147 builds the model
    assert probs.shape == (B, cfg.T, cfg.V) and attn.shape == (cfg.T, B, cfg.R, 1)
    assert np.allclose(probs.sum(-1), 1.0, atol=1e-5)
    lines = []
    model.summary(print_fn=lines.append)                                              # main.py:148
    assert any("Total params" in l for l in lines)

    class LossHistory:                                                                 # Callbacks/EpochLoss.py
        def __init__(self): self.batches, self.epochs = [], []
        def on_train_batch_end(self, batch, logs): self.batches.append(dict(logs))
        def on_epoch_end(self, epoch, logs): self.epochs.append(dict(logs))

    class LRScheduler:                                                                 # main.py:86-92,218-219
        def set_model(self, m): self.model = m
        def on_epoch_begin(self, epoch, logs): self.model.optimizer.lr = 0.0001 if epoch < 1 else 0.00005

    class Checkpoint:                                                                  # main.py:172-190 (save_weights_only)
        def set_model(self, m): self.model = m
        def on_epoch_end(self, epoch, logs): self.model.save_weights(str(tmp_path / f"model-ep{epoch + 1:03d}"))

    hist = LossHistory()
    model.fit(train_gen, epochs=2, steps_per_epoch=len(train_gen), callbacks=[LRScheduler(), hist, Checkpoint()],
              validation_data=val_gen, validation_steps=len(val_gen), initial_epoch=0)
    assert len(hist.batches) == 2 * len(train_gen) and len(hist.epochs) == 2
    assert set(hist.batches[0]) == {"loss", "L2", "accuracy", "attention", "lr"}       # lc_NIC.py:408
    assert {"val_loss", "val_L2", "val_accuracy", "val_attention"} <= set(hist.epochs[0])
    assert hist.batches[0]["lr"] == pytest.approx(1e-4) and hist.batches[-1]["lr"] == pytest.approx(5e-5)
    assert all(np.isfinite(v) for e in hist.epochs for v in e.values())
    train_gen.check(); val_gen.check()

    # eval.py:140-193: a fresh model, weights by name from the checkpoint, greedy loop over the test generator
    ev = _new_model(cfg, groups, B)
    ev.build(cfg.N, B)
    ev.load_weights(str(tmp_path / "model-ep002"), by_name=True, skip_mismatch=True)
    for a, b in zip(ev.get_weights(), model.get_weights()):
        assert np.array_equal(a, b)
    ev.get_layer("lstm").set_weights(model.get_layer("lstm").get_weights())            # main.py:161-162
    n_seen = 0
    for i in range(len(test_gen) + 1):                                                 # eval.py:155 (+1: the remainder)
        (features, _, a0, c0), _, keys = test_gen[i]
        start_seq = np.repeat([syn.START_ID], features.shape[0])                       # eval.py:161
        outputs, outputs_raw, attention_scores, attention_weights = \
            ev.greedy_predict(features, a0, c0, start_seq, cfg.T, cfg.U, None)       # eval.py:163
        n = features.shape[0]
        assert outputs.shape == (n, cfg.T, 1) and outputs_raw.shape == (n, cfg.T, cfg.V)
        assert attention_scores.shape == (cfg.T, n, cfg.R, 1) and attention_weights.shape == (cfg.T, n, cfg.R, cfg.A)
        assert len(keys) == n
        if n:
            assert np.array_equal(outputs[:, :, 0], outputs_raw.argmax(-1))
            assert np.array_equal(outputs, model.greedy_predict(features, a0, c0, start_seq, cfg.T, cfg.U, None)[0])
        n_seen += n
    assert n_seen == 11
    empty = ev.greedy_predict(np.zeros((0, cfg.N), np.float32), np.zeros((0, cfg.U), np.float32),
                              np.zeros((0, cfg.U), np.float32), np.zeros(0, np.int32), cfg.T, cfg.U, None)
    assert empty[0].shape == (0, cfg.T, 1)


def test_main_ms_py_flow_two_subjects(tmp_path):
    """main_ms.py:139-158 + data_generator_multisub: ms2_NIC with the two-subject generator (A's rows first, B's behind)."""
    cfg = K.tiny_config(n_enc=2)
    groups = syn.small_groups(cfg.N, cfg.R, seed=1, overlap=True)
    B, rng = 3, np.random.default_rng(5)                       # rows per subject
    betas = (rng.standard_normal((12, cfg.N), dtype=np.float32), rng.standard_normal((10, cfg.N), dtype=np.float32))
    rows = (rng.integers(0, 12, 14), rng.integers(0, 10, 14))
    caps = tuple([[syn.START_ID] + list(map(int, rng.integers(4, cfg.V, rng.integers(1, 6)))) + [syn.END_ID]
                  for _ in range(14)] for _ in range(2))
    np.random.seed(42)
    gen = ds.MultiSubjectDeviceDataGenerator(betas, rows, caps, B, cfg.U, cfg.T, cfg.V, shuffle=True, training=True, onehot=True)
    m = pkg.NIC((groups, [cfg.G] * cfg.R), cfg.U, 512, cfg.E, cfg.A, cfg.V, cfg.T,
                cfg.dropout_input, cfg.dropout_features, cfg.dropout_text, cfg.dropout_attn, cfg.dropout_lstm,
                cfg.dropout_out, cfg.input_reg, cfg.attn_reg, cfg.lstm_reg, cfg.output_reg,
                n_enc=2, inter_units=cfg.H, batch_max=2 * B, seed=42)
    m.compile(pkg.Adam(learning_rate=0.0001, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1), None)
    out = m(gen[0][0])                                                                 # ms2_NIC.py:205: 4-tuple
    assert len(out) == 4 and out[0].shape == (B, cfg.T, cfg.V) and out[3].shape == (cfg.T, B, cfg.R, 1)
    logs = []

    class Hist:
        def on_train_batch_end(self, batch, l): logs.append(dict(l))
    m.fit(gen, epochs=1, steps_per_epoch=len(gen), callbacks=[Hist()], validation_data=gen, validation_steps=1)
    assert len(logs) == len(gen) == 4
    assert set(logs[0]) == {"loss", "lossA", "lossB", "L2", "accuracyA", "accuracyB", "attentionA", "attentionB", "lr"}   # ms2_NIC.py:366-374
    assert logs[0]["loss"] == pytest.approx((logs[0]["lossA"] + logs[0]["lossB"]) / 2, rel=1e-6)                        # ms2_NIC.py:355
    assert all(np.isfinite(v) for l in logs for v in l.values())
    assert "val_lossA" in m.history
    m.save_weights(str(tmp_path / "ms2"))
    names = set(np.load(str(tmp_path / "ms2.npz")).files)
    assert any(n.startswith("enc0/") for n in names) and any(n.startswith("enc1/") for n in names)
    gen.check()
