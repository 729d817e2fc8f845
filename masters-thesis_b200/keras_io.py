"""Keras weights-only HDF5 checkpoints for the drop-in model.

The reference saves `model-ep{epoch:03d}.h5` / `model-latest.h5` with `ModelCheckpoint(save_weights_only=True)`
(AttemptFour/main.py:166-190) and reloads with `model.load_weights(path, by_name=True, skip_mismatch=True)`
(AttemptFour/eval.py:140, main.py:261-267).  Keras' layout (keras/saving/hdf5_format.py, Keras 2.x):

    /                      attrs: layer_names = [b"drop_in", ..., b"locally_dense", ...], backend, keras_version
    /<layer name>/         attrs: weight_names = [b"<weight name>", ...]   (layer.weights: trainable, then non-trainable)
    /<layer name>/<weight name>   one float32 dataset per weight

`by_name=True` matches LAYERS by name and, inside a layer, weights by POSITION (the dataset names are only read from
`weight_names`); `by_name=False` pairs the layers that own weights in order.  Both are restated below.

Layer names of the subclassed model (auto-names of the un-named layers follow Keras' snake_case + per-class counter):
`locally_dense` (`locally_dense_1` for the second encoder of ms2_NIC), `attention`, `emb_text`, `lstm`,
`time_distributed_nonlinear`, `time_distributed_softmax` (lc_NIC.py:84-159; layers.py:6; attention.py:6).  Weight
names are written the way TF 2.x names variables of nested layers (`nic/locally_dense/dense_17/kernel:0`,
`nic/lstm/lstm_cell/kernel:0`, ...): un-named Dense layers count up in creation order (encoder parcels first, then the
attention's W1, W2, V).  TensorFlow cannot run here, so those spellings are an assumption; reading never depends on them.

File access goes through `h5py` when it is importable, else through `h5min.py`, a self-contained reader/writer of the
HDF5 subset Keras uses (contiguous little-endian datasets, fixed-length string attributes, old-style groups).
"""
import numpy as np


def _h5():
    try:
        import h5py
        return h5py
    except ImportError:
        from . import h5min
        return h5min


def keras_layers(model):
    """[(keras layer name, [(keras weight name, product variable name), ...])] in `model.layers` order."""
    R, n_enc = model.R, model.n_enc
    scope = "nic"
    dense = lambda i: "dense" if i == 0 else f"dense_{i}"
    out = []
    for e in range(n_enc):
        lname = "locally_dense" if e == 0 else f"locally_dense_{e}"
        ws = []
        for r in range(R):
            ws.append((f"{scope}/{lname}/{dense(e * R + r)}/kernel:0", f"enc{e}/dense_{r}/kernel"))
            ws.append((f"{scope}/{lname}/{dense(e * R + r)}/bias:0", f"enc{e}/dense_{r}/bias"))
        for k in ("gamma", "beta", "moving_mean", "moving_variance"):       # trainable first, then non-trainable
            ws.append((f"{scope}/{lname}/input_bn/{k}:0", f"enc{e}/input_bn/{k}"))
        out.append((lname, ws))
    d0 = n_enc * R
    out.append(("attention", [(f"{scope}/attention/{dense(d0 + i)}/{kb}:0", f"attention/{nm}/{kb}")
                              for i, nm in enumerate(("W1", "W2", "V")) for kb in ("kernel", "bias")]))
    out.append(("emb_text", [(f"{scope}/emb_text/embeddings:0", "emb_text/embeddings")]))
    out.append(("lstm", [(f"{scope}/lstm/lstm_cell/{k}:0", f"lstm/{k}") for k in ("kernel", "recurrent_kernel", "bias")]))
    for l in ("time_distributed_nonlinear", "time_distributed_softmax"):
        out.append((l, [(f"{scope}/{l}/{k}:0", f"{l}/{k}") for k in ("kernel", "bias")]))
    return out


# layers without weights that `model.layers` lists before the weighted ones (lc_NIC.py:51-55)
_WEIGHTLESS = ["drop_in", "drop_feat", "drop_text", "drop_lstm", "drop_out"]


def save_h5(model, path):
    h5 = _h5()
    named = model.get_named_weights()
    layers = keras_layers(model)
    with h5.File(path, "w") as f:
        f.attrs["layer_names"] = np.array([n.encode() for n in _WEIGHTLESS + [l for l, _ in layers]])
        f.attrs["backend"] = np.bytes_(b"tensorflow")
        f.attrs["keras_version"] = np.bytes_(b"2.8.0")
        for n in _WEIGHTLESS:
            g = f.create_group(n)
            g.attrs["weight_names"] = np.array([], dtype="S1")
        for lname, ws in layers:
            g = f.create_group(lname)
            g.attrs["weight_names"] = np.array([k.encode() for k, _ in ws])
            for kname, pname in ws:
                g.create_dataset(kname, data=np.ascontiguousarray(named[pname], dtype=np.float32))


def _attr_names(obj, key):
    """Keras splits attributes larger than 64 KB into key0, key1, ... (hdf5_format.load_attributes_from_hdf5_group)."""
    if key in obj.attrs:
        vals = list(obj.attrs[key])
    else:
        vals, i = [], 0
        while f"{key}{i}" in obj.attrs:
            vals.extend(obj.attrs[f"{key}{i}"])
            i += 1
    return [v.decode() if isinstance(v, (bytes, np.bytes_)) else str(v) for v in vals]


def load_h5(model, path, by_name=False, skip_mismatch=False):
    """-> {product variable name: array}.  Layers are matched by name (by_name) or, as Keras' topological loading,
    the layers that own weights are paired in order; inside a layer weights are matched by position."""
    h5 = _h5()
    mine = keras_layers(model)
    named = {}
    with h5.File(path, "r") as f:
        root = f["model_weights"] if "model_weights" in f and "layer_names" not in f.attrs else f
        file_layers = []
        for lname in _attr_names(root, "layer_names"):
            g = root[lname]
            wn = _attr_names(g, "weight_names")
            file_layers.append((lname, [np.asarray(g[w]) for w in wn]))
        if by_name:
            index = {l: arrs for l, arrs in file_layers}
            pairs = [(l, ws, index[l]) for l, ws in mine if l in index]
        else:
            weighted = [(l, a) for l, a in file_layers if a]
            if len(weighted) != len(mine):
                raise ValueError(f"{path}: {len(weighted)} layers with weights in the file, the model has {len(mine)}")
            pairs = [(l, ws, a) for (l, ws), (_, a) in zip(mine, weighted)]
        for lname, ws, arrs in pairs:
            if len(arrs) != len(ws):
                if skip_mismatch:
                    continue
                raise ValueError(f"layer {lname}: file holds {len(arrs)} weights, the model expects {len(ws)}")
            for (_, pname), a in zip(ws, arrs):
                named[pname] = a
    return named
