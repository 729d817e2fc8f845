"""CPU tests of the Keras checkpoint layer: the HDF5 subset reader/writer and the layer / weight layout
of keras_io.py, on a stand-in for the model (no GPU needed: only names and arrays are involved)."""
import struct
import numpy as np
import pytest

from tests.util import pkg

h5min, keras_io = pkg.h5min if hasattr(pkg, "h5min") else None, pkg.keras_io


def _h5min():
    import importlib
    return importlib.import_module("masters-thesis_b200.h5min")


def test_h5min_roundtrip_nested_groups_attributes_and_large_groups(tmp_path):
    H = _h5min()
    rng = np.random.default_rng(0)
    p = str(tmp_path / "t.h5")
    arrays = {f"nic/locally_dense/dense_{i}/kernel:0": rng.standard_normal((7 + i % 5, 32)).astype(np.float32)
              for i in range(361)}
    with H.File(p, "w") as f:
        f.attrs["layer_names"] = np.array([b"drop_in", b"locally_dense"])
        f.attrs["backend"] = np.bytes_(b"tensorflow")
        g = f.create_group("locally_dense")
        g.attrs["weight_names"] = np.array([k.encode() for k in arrays])
        for k, a in arrays.items():
            g.create_dataset(k, data=a)
        e = f.create_group("drop_in")
        e.attrs["weight_names"] = np.array([], dtype="S1")
        f.create_dataset("scalars/i64", data=np.arange(5, dtype=np.int64))
        f.create_dataset("scalars/f64", data=np.float64(2.5))
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and struct.unpack_from("<Q", raw, 40)[0] == len(raw)     # end-of-file address
    with H.File(p, "r") as f:
        assert [x.decode() for x in f.attrs["layer_names"]] == ["drop_in", "locally_dense"]
        assert bytes(f.attrs["backend"]) == b"tensorflow"
        g = f["locally_dense"]
        names = [x.decode() for x in g.attrs["weight_names"]]
        assert names == list(arrays)
        for k in names:
            assert np.array_equal(np.asarray(g[k]), arrays[k])
        assert len(f["locally_dense/nic/locally_dense"].keys()) == 361
        assert f["drop_in"].attrs["weight_names"].size == 0 and f["drop_in"].keys() == []
        assert np.array_equal(np.asarray(f["scalars/i64"]), np.arange(5)) and float(np.asarray(f["scalars/f64"])) == 2.5
        assert "scalars/i64" in f and "nope" not in f


def test_h5min_rejects_what_it_cannot_read(tmp_path):
    H = _h5min()
    p = tmp_path / "bad.h5"
    p.write_bytes(b"not an hdf5 file at all")
    with pytest.raises(ValueError):
        H.File(str(p), "r")


class _FakeModel:
    """What keras_io needs from NIC: R, n_enc, get_named_weights()."""

    def __init__(self, R=5, n_enc=1, seed=0):
        self.R, self.n_enc = R, n_enc
        rng = np.random.default_rng(seed)
        self.named = {}
        for e in range(n_enc):
            for r in range(R):
                self.named[f"enc{e}/dense_{r}/kernel"] = rng.standard_normal((3 + r, 8)).astype(np.float32)
                self.named[f"enc{e}/dense_{r}/bias"] = rng.standard_normal(8).astype(np.float32)
            for k in ("gamma", "beta", "moving_mean", "moving_variance"):
                self.named[f"enc{e}/input_bn/{k}"] = rng.standard_normal(8).astype(np.float32)
        shapes = {"attention/W1/kernel": (8, 4), "attention/W1/bias": (4,), "attention/W2/kernel": (16, 4),
                  "attention/W2/bias": (4,), "attention/V/kernel": (4, 1), "attention/V/bias": (1,),
                  "emb_text/embeddings": (11, 6), "lstm/kernel": (14, 64), "lstm/recurrent_kernel": (16, 64),
                  "lstm/bias": (64,), "time_distributed_nonlinear/kernel": (16, 12), "time_distributed_nonlinear/bias": (12,),
                  "time_distributed_softmax/kernel": (12, 11), "time_distributed_softmax/bias": (11,)}
        for k, s in shapes.items():
            self.named[k] = rng.standard_normal(s).astype(np.float32)

    def get_named_weights(self):
        return dict(self.named)


def test_keras_layout_names_and_order():
    m = _FakeModel(R=3, n_enc=2)
    layers = keras_io.keras_layers(m)
    assert [l for l, _ in layers] == ["locally_dense", "locally_dense_1", "attention", "emb_text", "lstm",
                                      "time_distributed_nonlinear", "time_distributed_softmax"]
    ld = dict(layers)
    # layer.weights: trainable first (dense kernels / biases in creation order, then gamma, beta), then the moving statistics
    assert [p for _, p in ld["locally_dense"]][-4:] == ["enc0/input_bn/gamma", "enc0/input_bn/beta",
                                                        "enc0/input_bn/moving_mean", "enc0/input_bn/moving_variance"]
    assert ld["locally_dense"][0][0] == "nic/locally_dense/dense/kernel:0"        # first un-named Dense is "dense"
    assert ld["locally_dense_1"][0][0] == "nic/locally_dense_1/dense_3/kernel:0"  # the counter runs on
    assert ld["attention"][0][0] == "nic/attention/dense_6/kernel:0"
    assert [k for k, _ in ld["lstm"]] == ["nic/lstm/lstm_cell/kernel:0", "nic/lstm/lstm_cell/recurrent_kernel:0",
                                          "nic/lstm/lstm_cell/bias:0"]
    every = [p for _, ws in layers for _, p in ws]
    assert sorted(every) == sorted(m.named) and len(set(every)) == len(every)


@pytest.mark.parametrize("by_name", [True, False])
def test_keras_h5_save_and_load_by_name_and_by_order(tmp_path, by_name):
    m = _FakeModel()
    p = str(tmp_path / "model-ep001.h5")
    keras_io.save_h5(m, p)
    got = keras_io.load_h5(_FakeModel(seed=1), p, by_name=by_name)
    assert sorted(got) == sorted(m.named)
    for k in got:
        assert np.array_equal(got[k], m.named[k])


def test_keras_h5_from_a_differently_spelled_file(tmp_path):
    """Reading never depends on our spelling of the weight names: layers are matched by name, weights by position,
    dataset paths come from the file's own `weight_names` (a file written by real Keras spells them its own way),
    full-model files keep the weights under /model_weights, and a layer with another weight count is skipped or refused."""
    H = _h5min()
    m = _FakeModel()
    p = str(tmp_path / "foreign.h5")
    with H.File(p, "w") as f:
        mw = f.create_group("model_weights")
        mw.attrs["layer_names"] = np.array([b"lstm", b"emb_text", b"attention", b"unknown_layer"])
        g = mw.create_group("lstm")
        g.attrs["weight_names"] = np.array([b"lstm_3/whatever/k:0", b"lstm_3/whatever/rk:0", b"lstm_3/whatever/b:0"])
        for n, k in zip(("k", "rk", "b"), ("lstm/kernel", "lstm/recurrent_kernel", "lstm/bias")):
            g.create_dataset(f"lstm_3/whatever/{n}:0", data=m.named[k])
        g = mw.create_group("emb_text")
        g.attrs["weight_names"] = np.array([b"e:0"])
        g.create_dataset("e:0", data=m.named["emb_text/embeddings"])
        g = mw.create_group("attention")                               # wrong number of weights
        g.attrs["weight_names"] = np.array([b"a:0"])
        g.create_dataset("a:0", data=np.zeros(3, np.float32))
        g = mw.create_group("unknown_layer")
        g.attrs["weight_names"] = np.array([], dtype="S1")
    got = keras_io.load_h5(m, p, by_name=True, skip_mismatch=True)
    assert sorted(got) == ["emb_text/embeddings", "lstm/bias", "lstm/kernel", "lstm/recurrent_kernel"]
    assert np.array_equal(got["lstm/recurrent_kernel"], m.named["lstm/recurrent_kernel"])
    with pytest.raises(ValueError):
        keras_io.load_h5(m, p, by_name=True, skip_mismatch=False)
