"""Shared helpers for the parity tests: oracle <-> product weight naming, error measures."""
import importlib
import numpy as np

pkg = importlib.import_module("masters-thesis_b200")
syn = pkg.synthetic


def oracle_to_named(cfg, P):
    """oracle parameter dict -> {product variable name: array}."""
    d = {}
    for e in range(cfg.n_enc):
        for r in range(cfg.R):
            d[f"enc{e}/dense_{r}/kernel"] = P[f"enc{e}_W"][r]
            d[f"enc{e}/dense_{r}/bias"] = P[f"enc{e}_b"][r]
        d[f"enc{e}/input_bn/gamma"] = P[f"enc{e}_gamma"]
        d[f"enc{e}/input_bn/beta"] = P[f"enc{e}_beta"]
        d[f"enc{e}/input_bn/moving_mean"] = P[f"enc{e}_mean"]
        d[f"enc{e}/input_bn/moving_variance"] = P[f"enc{e}_var"]
    m = {"att_W1": "attention/W1/kernel", "att_b1": "attention/W1/bias", "att_W2": "attention/W2/kernel",
         "att_b2": "attention/W2/bias", "att_V": "attention/V/kernel", "att_bV": "attention/V/bias",
         "emb": "emb_text/embeddings", "lstm_W": "lstm/kernel", "lstm_U": "lstm/recurrent_kernel",
         "lstm_b": "lstm/bias", "inter_W": "time_distributed_nonlinear/kernel",
         "inter_b": "time_distributed_nonlinear/bias", "out_W": "time_distributed_softmax/kernel",
         "out_b": "time_distributed_softmax/bias"}
    for k, v in m.items():
        d[v] = P[k]
    return d


def make_model(cfg, groups, P=None, batch_max=64, seed=0, clipnorm=0.1, **extra):
    """Product NIC configured like the oracle Config, optionally loaded with the oracle's weights."""
    if getattr(cfg, "attn_loss", ""):
        extra.setdefault("attention_loss", cfg.attn_loss)
    if getattr(cfg, "agc", ()):
        extra.setdefault("agc", tuple(cfg.agc))
    model = pkg.NIC((groups[0], [cfg.G] * cfg.R), cfg.U, 512, cfg.E, cfg.A, cfg.V, cfg.T,
                    cfg.dropout_input, cfg.dropout_features, cfg.dropout_text, cfg.dropout_attn, cfg.dropout_lstm,
                    cfg.dropout_out, cfg.input_reg, cfg.attn_reg, cfg.lstm_reg, cfg.output_reg,
                    n_enc=cfg.n_enc, batch_max=batch_max, seed=seed, inter_units=cfg.H,
                    masked_sparse_ce=cfg.masked_sparse_ce, plain_attention=cfg.plain_attention, **extra)
    model.compile(pkg.Adam(cfg.lr, beta_1=cfg.beta1, beta_2=cfg.beta2, epsilon=cfg.adam_eps, clipnorm=clipnorm), None)
    model.build(cfg.N, batch_max)
    if P is not None:
        model.set_named_weights(oracle_to_named(cfg, P))
    return model


def rel_err(a, b):
    """||a-b|| / max(||b||, tiny)."""
    a = np.asarray(a, dtype=np.float64).reshape(-1)
    b = np.asarray(b, dtype=np.float64).reshape(-1)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def grad_errors(cfg, grads_oracle, grads_dev):
    """{variable name: relative error} between the oracle's tape gradient and the device gradient."""
    named = oracle_to_named(cfg, {**grads_oracle, **{f"enc{e}_{k}": np.zeros(cfg.G) for e in range(cfg.n_enc)
                                                     for k in ("mean", "var")}})
    out = {}
    for name, g in named.items():
        if name in grads_dev:
            out[name] = (rel_err(grads_dev[name], g), float(np.linalg.norm(np.asarray(g, dtype=np.float64))))
    return out


_DEC = {"attention/W1/kernel": "att_W1", "attention/W1/bias": "att_b1", "attention/W2/kernel": "att_W2",
        "attention/W2/bias": "att_b2", "attention/V/kernel": "att_V", "attention/V/bias": "att_bV",
        "emb_text/embeddings": "emb", "lstm/kernel": "lstm_W", "lstm/recurrent_kernel": "lstm_U", "lstm/bias": "lstm_b",
        "time_distributed_nonlinear/kernel": "inter_W", "time_distributed_nonlinear/bias": "inter_b",
        "time_distributed_softmax/kernel": "out_W", "time_distributed_softmax/bias": "out_b"}


def product_to_oracle_name(name):
    """'enc0/dense_17/kernel' -> 'enc0_W' (the oracle keeps one list per encoder), 'lstm/kernel' -> 'lstm_W', ..."""
    import re
    m = re.fullmatch(r"enc(\d)/dense(?:_\d+)?/(kernel|bias)", name)
    if m:
        return f"enc{m.group(1)}_{'W' if m.group(2) == 'kernel' else 'b'}"
    m = re.fullmatch(r"enc(\d)/input_bn/(gamma|beta)", name)
    if m:
        return f"enc{m.group(1)}_{m.group(2)}"
    return _DEC.get(name)
