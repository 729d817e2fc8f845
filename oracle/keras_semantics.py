"""CPU restatement (torch eager, fp32 with an fp64 switch) of the reference hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED: the reference has no
tests/golden vectors for this path and TensorFlow (un-vendored dependency, loosely pinned
tensorflow 2.3.0 / tensorflow-gpu 2.8.0 / keras 2.8.0 in
AttemptFour/requirements_train.txt:46,112-115) is not installable here, so the Keras-2.x
op semantics below are restated from the published Keras behaviour, not checked against TF.

What is restated, op for op (citations relative to /root/reference):
  * encoder        AttemptFour/Model/layers.py:43-53   (LocallyDense: R x gather+Dense, BN, dropout)
                   AttemptFour/Model/img_localDense.py:28-38 (image control: per-location Dense)
  * attention      AttemptFour/Model/attention.py:25-44
  * decoder loop   AttemptFour/Model/lc_NIC.py:223-263 (call_attention)
  * loss/metrics   AttemptFour/Model/lc_NIC.py:328-408 (train_step), :410-459 (test_step),
                   :461-466 (loss_function), :468-486 (accuracy_calculation);
                   loss object AttemptFour/main.py:107-110
  * optimizer      AttemptFour/main.py:97 (Adam 1e-4, b1 .9, b2 .98, eps 1e-8, clipnorm .1)
  * greedy decode  AttemptFour/Model/lc_NIC.py:577-638
  * two subjects   AttemptFour/Model/ms2_NIC.py:177-205,207-291,294-374
  * step variants  attention loss in the differentiated loss (lc_NIC.py:365-367,384), adaptive gradient clipping
                   (AttemptFour/Model/agc.py:20-38, call site lc_NIC.py:388), train_step_sam (lc_NIC.py:713-838)
  * other decoders sample_choice (lc_NIC.py:571-575), top-k / nucleus (AttemptFour/Model/img_NIC.py:509-533),
                   beam search (intent of lc_NIC.py:640-692)

Keras-2.x semantics that are not visible in the reference source:
  1. Dense: act(x @ W + b), W is (in,out).  LeakyReLU(0.2): x if x > 0 else 0.2 x (grad 0.2 at 0).
  2. BatchNormalization(axis=-1) on a 3-D tensor, non-fused path: training => per-channel mean and
     BIASED variance over (B,R); y = x*inv + (beta - mean*inv), inv = gamma*rsqrt(var+1e-3);
     moving <- 0.99*moving + 0.01*batch; init gamma 1, beta 0, moving_mean 0, moving_var 1.
  3. Dropout(p): multiplier 0 or 1/(1-p); identity when not training or p == 0.  Masks come from
     the shared Philox contract (oracle/philox.py) because TF's RNG cannot be replayed.
  4. LSTM v2, implementation 2: z = x@W + h@U + b, gates i,f,c~,o; sigmoid recurrent activation;
     `dropout=` multiplies the step INPUT by one mask per layer call (= per step here);
     kernel_regularizer touches W only; bias zeros with forget slice 1.
  5. Embedding: gather; gradient = IndexedSlices(values (B*T,E), indices = flattened ids).
  6. Softmax: exp(x-max)/sum.  categorical_crossentropy(from_logits=False): p /= sum(p);
     p = clip(p, 1e-7, 1-1e-7); -sum(y log p).  reduction='none' then reduce_mean over batch.
  7. L2(l)(W) = l * sum(W^2).
  8. Adam.apply_gradients with clipnorm (Keras >= 2.4 applies it; switch `apply_clipnorm`):
     per VARIABLE tf.clip_by_norm; for the embedding's IndexedSlices the norm is taken over the
     un-aggregated (B*T,E) values (switch `emb_unaggregated_norm`), duplicates are summed, and
     the sparse Adam path decays/updates ALL rows (== dense Adam on the scatter-added grad).
     ResourceApplyAdam: alpha = lr*sqrt(1-b2^t)/(1-b1^t); m += (g-m)(1-b1); v += (g^2-v)(1-b2);
     w -= alpha*m/(sqrt(v)+eps).
  9. argmax: first maximal index.
"""
from dataclasses import dataclass, field
import math
import numpy as np
import torch

from . import philox

LRELU_ALPHA = 0.2
BN_EPS = 1e-3
BN_MOMENTUM = 0.99
CE_EPS = 1e-7


# ----------------------------------------------------------------------------------------------
# configuration
# ----------------------------------------------------------------------------------------------
@dataclass
class Config:
    """Model hyper-parameters (AttemptFour/config.yaml:24-57; lc_NIC.py:42)."""
    N: int = 327684           # input vertices (config.yaml:46)
    R: int = 345              # parcels after main.py:115's removal
    G: int = 32               # group_size
    U: int = 512              # units
    E: int = 512              # embedding_text
    A: int = 32               # attn_units
    H: int = 256              # dense_inter width, hard-coded lc_NIC.py:141
    V: int = 5001             # top_k + 1
    T: int = 15               # max_length
    dropout_input: float = 0.0
    dropout_features: float = 0.2
    dropout_text: float = 0.2
    dropout_attn: float = 0.2
    dropout_lstm: float = 0.2
    dropout_out: float = 0.2
    input_reg: float = 0.01
    attn_reg: float = 0.001
    lstm_reg: float = 0.00003
    output_reg: float = 0.00001
    lr: float = 1e-4
    beta1: float = 0.9
    beta2: float = 0.98
    adam_eps: float = 10.0e-9
    clipnorm: float = 0.1
    apply_clipnorm: bool = True
    emb_unaggregated_norm: bool = True
    n_enc: int = 1            # 2 => ms2_NIC (first half of the batch -> encoder a, second -> b)
    encoder_kind: str = "parcel"   # "parcel" (layers.py) | "image" (img_localDense.py: x is (B,R,K))
    img_K: int = 512          # per-location feature width of the image control
    # ThinkAndTell/att_model.py variant of the same decoder
    masked_sparse_ce: bool = False   # SparseCategoricalCrossentropy(from_logits=True) x pad mask (att_model.py:306-321)
    plain_attention: bool = False    # attention projections fc1/fc2 without LeakyReLU (att_model.py:11-29)
    # step variants that live in the model file
    attn_loss: str = ""              # "" reported only | "batch": lc_NIC.py:384 enabled | "elementwise": lc_NIC.py:751-752,774
    agc: tuple = ()                  # (clip_factor, eps): lc_NIC.py:388 enabled (agc.py:20-38)

    def no_dropout(self):
        c = Config(**self.__dict__)
        c.dropout_input = c.dropout_features = c.dropout_text = 0.0
        c.dropout_attn = c.dropout_lstm = c.dropout_out = 0.0
        return c


def tiny_config(**kw):
    """SURVEY.md §8c tiny case."""
    base = dict(N=200, R=6, G=8, U=16, E=16, A=8, H=12, V=37, T=5)
    base.update(kw)
    return Config(**base)


# ----------------------------------------------------------------------------------------------
# Keras initialisers (restated; distributions only -- TF's own random streams are not replayable)
# ----------------------------------------------------------------------------------------------
def _trunc_normal(rng, shape, std):
    x = rng.standard_normal(shape)
    bad = np.abs(x) > 2.0
    while bad.any():
        x[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(x) > 2.0
    return (x * std).astype(np.float32)


def he_normal(rng, shape):
    return _trunc_normal(rng, shape, math.sqrt(2.0 / shape[0]) / 0.87962566103423978)


def glorot_normal(rng, shape):
    return _trunc_normal(rng, shape, math.sqrt(2.0 / (shape[0] + shape[1])) / 0.87962566103423978)


def glorot_uniform(rng, shape):
    lim = math.sqrt(6.0 / (shape[0] + shape[1]))
    return rng.uniform(-lim, lim, shape).astype(np.float32)


def orthogonal(rng, shape):
    rows, cols = shape
    a = rng.standard_normal((max(rows, cols), min(rows, cols)))
    q, r = np.linalg.qr(a)
    q = q * np.sign(np.diag(r))
    if rows < cols:
        q = q.T
    return q.reshape(shape).astype(np.float32)


def init_params(cfg, groups, seed=0):
    """Weights as a dict of numpy fp32 arrays.  `groups` = list (per encoder) of lists of index arrays
    (parcel encoder) -- ignored for the image encoder."""
    rng = np.random.default_rng(seed)
    P = {}
    for e in range(cfg.n_enc):
        if cfg.encoder_kind == "parcel":
            sizes = [len(g) for g in groups[e]]
        else:
            sizes = [cfg.img_K] * cfg.R
        P[f"enc{e}_W"] = [he_normal(rng, (n, cfg.G)) for n in sizes]          # layers.py:33, lc_NIC.py:87
        P[f"enc{e}_b"] = [np.zeros(cfg.G, np.float32) for _ in sizes]
        P[f"enc{e}_gamma"] = np.ones(cfg.G, np.float32)                        # layers.py:40
        P[f"enc{e}_beta"] = np.zeros(cfg.G, np.float32)
        P[f"enc{e}_mean"] = np.zeros(cfg.G, np.float32)
        P[f"enc{e}_var"] = np.ones(cfg.G, np.float32)
    P["att_W1"] = he_normal(rng, (cfg.G, cfg.A)); P["att_b1"] = np.zeros(cfg.A, np.float32)   # attention.py:21
    P["att_W2"] = he_normal(rng, (cfg.U, cfg.A)); P["att_b2"] = np.zeros(cfg.A, np.float32)   # attention.py:22
    P["att_V"] = glorot_uniform(rng, (cfg.A, 1)); P["att_bV"] = np.zeros(1, np.float32)        # attention.py:23
    P["emb"] = rng.uniform(-0.08, 0.08, (cfg.V, cfg.E)).astype(np.float32)                    # lc_NIC.py:105-112
    P["lstm_W"] = glorot_uniform(rng, (cfg.G + cfg.E, 4 * cfg.U))                             # lc_NIC.py:118-124
    P["lstm_U"] = orthogonal(rng, (cfg.U, 4 * cfg.U))
    b = np.zeros(4 * cfg.U, np.float32); b[cfg.U:2 * cfg.U] = 1.0
    P["lstm_b"] = b
    P["inter_W"] = glorot_normal(rng, (cfg.U, cfg.H)); P["inter_b"] = np.zeros(cfg.H, np.float32)  # lc_NIC.py:139-148
    P["out_W"] = glorot_normal(rng, (cfg.H, cfg.V)); P["out_b"] = np.zeros(cfg.V, np.float32)      # lc_NIC.py:150-159
    return P


def trainable_names(cfg):
    """Variables in the order the product's flat layout uses (encoder(s), attention, embedding, lstm, head)."""
    names = []
    for e in range(cfg.n_enc):
        names += [f"enc{e}_W", f"enc{e}_b", f"enc{e}_gamma", f"enc{e}_beta"]
    names += ["att_W1", "att_b1", "att_W2", "att_b2", "att_V", "att_bV", "emb",
              "lstm_W", "lstm_U", "lstm_b", "inter_W", "inter_b", "out_W", "out_b"]
    return names


def l2_lambda(cfg, name):
    """Which kernels carry an L2 regulariser (lc_NIC.py:84-159)."""
    if name.startswith("enc") and name.endswith("_W"):
        return cfg.input_reg
    if name in ("att_W1", "att_W2"):
        return cfg.attn_reg
    if name == "lstm_W":
        return cfg.lstm_reg
    if name in ("inter_W", "out_W"):
        return cfg.output_reg
    return 0.0


def to_torch(P, dtype=torch.float32, requires_grad=False):
    out = {}
    for k, v in P.items():
        if isinstance(v, list):
            out[k] = [torch.tensor(a, dtype=dtype, requires_grad=requires_grad) for a in v]
        else:
            out[k] = torch.tensor(v, dtype=dtype, requires_grad=requires_grad)
    return out


def to_numpy(P):
    out = {}
    for k, v in P.items():
        if isinstance(v, list):
            out[k] = [a.detach().cpu().numpy().astype(np.float32) for a in v]
        else:
            out[k] = v.detach().cpu().numpy().astype(np.float32)
    return out


# ----------------------------------------------------------------------------------------------
# forward
# ----------------------------------------------------------------------------------------------
def lrelu(x):
    return torch.where(x > 0, x, LRELU_ALPHA * x)


def _mask(shape, rate, site, step, seed, training, dtype):
    if not training or rate <= 0:
        return None
    return torch.from_numpy(philox.keep_mask(shape, rate, site, step, seed, np.float64)).to(dtype)


def _drop(x, rate, site, step, seed, training):
    m = _mask(tuple(x.shape), rate, site, step, seed, training, x.dtype)
    return x if m is None else x * m


# Hook for the data-parallel tests: a differentiable SUM-all-reduce (torch.distributed.nn.functional.all_reduce)
# applied to the BatchNorm batch sums.  None = single process (the reference's semantics).
BN_SUM_REDUCE = None


def encoder(cfg, P, e, groups, x, training, seed, stats_out=None, second_dropout=False):
    """LocallyDense (layers.py:43-53) / image LocallyDense (img_localDense.py:28-38)."""
    W, bias = P[f"enc{e}_W"], P[f"enc{e}_b"]
    if cfg.encoder_kind == "parcel":
        out = [lrelu(x[:, torch.as_tensor(idx, dtype=torch.long)] @ W[r] + bias[r])
               for r, idx in enumerate(groups[e])]                        # layers.py:45
    else:
        out = [lrelu(x[:, r, :] @ W[r] + bias[r]) for r in range(cfg.R)]  # img_localDense.py:31
    out = torch.stack(out).permute(1, 0, 2)                               # (B,R,G) layers.py:48-49
    gamma, beta = P[f"enc{e}_gamma"], P[f"enc{e}_beta"]
    if training and BN_SUM_REDUCE is not None:
        # data-parallel restatement: per-channel (sum x, sum x^2, count) are summed over the ranks
        n = BN_SUM_REDUCE(torch.tensor(float(out.shape[0] * out.shape[1]), dtype=out.dtype))
        mean = BN_SUM_REDUCE(out.sum(dim=(0, 1))) / n
        var = BN_SUM_REDUCE((out * out).sum(dim=(0, 1))) / n - mean * mean
        if stats_out is not None:
            stats_out[f"enc{e}_mean"] = mean.detach()
            stats_out[f"enc{e}_var"] = var.detach()
    elif training:
        mean = out.mean(dim=(0, 1))
        var = ((out - mean) ** 2).mean(dim=(0, 1))                        # biased
        if stats_out is not None:
            stats_out[f"enc{e}_mean"] = mean.detach()
            stats_out[f"enc{e}_var"] = var.detach()
    else:
        mean, var = P[f"enc{e}_mean"], P[f"enc{e}_var"]
    inv = gamma * torch.rsqrt(var + BN_EPS)
    y = out * inv + (beta - mean * inv)                                   # layers.py:50
    y = _drop(y, cfg.dropout_features, philox.SITE_FEAT, e, seed, training)   # layers.py:51
    if second_dropout:                                                    # ms2_NIC.py:214,257
        y = _drop(y, cfg.dropout_features, philox.SITE_FEAT2, e, seed, training)
    return y, out


def attention(cfg, P, hidden, features, step, training, seed, b_offset=0, B_total=None):
    """attention.py:25-44.  Returns (context, weights (B,R,1), hidden layer (B,R,A))."""
    act = (lambda v: v) if cfg.plain_attention else lrelu                 # att_model.py:15-17: bare Dense layers
    f1 = act(features @ P["att_W1"] + P["att_b1"])                        # W1 carries LeakyReLU (lc_NIC.py:95-101)
    h2 = act(hidden.unsqueeze(1) @ P["att_W2"] + P["att_b2"])
    s = torch.tanh(f1 + h2)                                               # attention.py:30-33
    if training and cfg.dropout_attn > 0:
        Bt = B_total if B_total is not None else s.shape[0]
        m = _mask((Bt,) + tuple(s.shape[1:]), cfg.dropout_attn, philox.SITE_ATTN, step, seed, True, s.dtype)
        s = s * m[b_offset:b_offset + s.shape[0]]                         # attention.py:36
    score = s @ P["att_V"] + P["att_bV"]                                  # attention.py:38
    w = torch.softmax(score, dim=1)                                       # attention.py:40
    ctx = (w * features).sum(dim=1)                                       # attention.py:42
    return ctx, w, s


def lstm_step(cfg, P, x, h, c, step, training, seed, b_offset=0, B_total=None):
    """One Keras LSTM(v2, implementation 2) step with the layer's input dropout (lc_NIC.py:255)."""
    if training and cfg.dropout_lstm > 0:
        Bt = B_total if B_total is not None else x.shape[0]
        m = _mask((Bt, x.shape[1]), cfg.dropout_lstm, philox.SITE_LSTM_IN, step, seed, True, x.dtype)
        x = x * m[b_offset:b_offset + x.shape[0]]
    z = x @ P["lstm_W"] + h @ P["lstm_U"] + P["lstm_b"]
    U = cfg.U
    i = torch.sigmoid(z[:, 0:U]); f = torch.sigmoid(z[:, U:2 * U])
    g = torch.tanh(z[:, 2 * U:3 * U]); o = torch.sigmoid(z[:, 3 * U:4 * U])
    c2 = f * c + i * g
    h2 = o * torch.tanh(c2)
    return h2, c2


def decoder(cfg, P, features, cap, a0, c0, training, seed, b_offset=0, B_total=None):
    """lc_NIC.call_attention after the encoder (lc_NIC.py:233-263).  Returns probs (B,T,V), attn (T,B,R,1)
    and the gathered embedding tensor (whose .grad is the IndexedSlices `values`)."""
    B, T = cap.shape
    Bt = B_total if B_total is not None else B
    emb_values = P["emb"][cap.long()]                                     # (B,T,E) lc_NIC.py:233
    if emb_values.requires_grad:
        emb_values.retain_grad()
    text = emb_values
    if training and cfg.dropout_text > 0:
        m = _mask((Bt, T, cfg.E), cfg.dropout_text, philox.SITE_TEXT, 0, seed, True, text.dtype)
        text = text * m[b_offset:b_offset + B]
    a, c = a0, c0
    attn, outs = [], []
    for t in range(T):                                                    # lc_NIC.py:244
        ctx, w, _ = attention(cfg, P, a, features, t, training, seed, b_offset, Bt)
        attn.append(w)
        sample = torch.cat([ctx, text[:, t, :]], dim=-1)                  # lc_NIC.py:253
        a, c = lstm_step(cfg, P, sample, a, c, t, training, seed, b_offset, Bt)
        o = a
        if training and cfg.dropout_lstm > 0:                             # lc_NIC.py:256 (dropout_lstm layer)
            m = _mask((Bt, cfg.U), cfg.dropout_lstm, philox.SITE_LSTM_OUT, t, seed, True, a.dtype)
            o = a * m[b_offset:b_offset + B]
        outs.append(o)
    out = torch.stack(outs, dim=1)                                        # (B,T,U) lc_NIC.py:258
    inter = lrelu(out @ P["inter_W"] + P["inter_b"])                      # lc_NIC.py:261
    if training and cfg.dropout_out > 0:
        m = _mask((Bt, T, cfg.H), cfg.dropout_out, philox.SITE_OUT, 0, seed, True, inter.dtype)
        inter = inter * m[b_offset:b_offset + B]
    logits = inter @ P["out_W"] + P["out_b"]
    probs = torch.softmax(logits, dim=-1)
    return probs, torch.stack(attn), emb_values, logits


def forward(cfg, P, groups, data, training, seed=0, stats_out=None):
    """`model(data, training)`: lc_NIC.py:163,223 (n_enc==1) or ms2_NIC.py:177-205 (n_enc==2).
    data = (betas, cap (B,T) int, a0, c0).  Returns dict(probs, attn, emb_values list, logits)."""
    x, cap, a0, c0 = data
    B = x.shape[0]
    if cfg.dropout_input > 0 and training:                                # lc_NIC.py:227
        x = x * _mask(tuple(x.shape), cfg.dropout_input, philox.SITE_INPUT, 0, seed, True, x.dtype)
    if cfg.n_enc == 1:
        feats, _ = encoder(cfg, P, 0, groups, x, training, seed, stats_out)
        probs, attn, ev, logits = decoder(cfg, P, feats, cap, a0, c0, training, seed)
        return dict(probs=probs, attn=attn, emb_values=[ev], logits=logits, feats=feats)
    # ms2: halves, both forced training=True by the reference (ms2_NIC.py:195,202,419,426);
    # callers pass training=True for ms2 in both train_step and test_step.
    h = B // 2
    probs, attn, evs, logits, fts = [], [], [], [], []
    for e, sl in enumerate((slice(0, h), slice(h, B))):
        feats, _ = encoder(cfg, P, e, groups, x[sl], training, seed, stats_out, second_dropout=True)
        p, a, ev, lg = decoder(cfg, P, feats, cap[sl], a0[sl], c0[sl], training, seed,
                               b_offset=sl.start, B_total=B)
        probs.append(p); attn.append(a); evs.append(ev); logits.append(lg); fts.append(feats)
    return dict(probs=torch.cat(probs), attn=torch.cat(attn, dim=1), emb_values=evs,
                logits=torch.cat(logits), feats=torch.cat(fts))


# ----------------------------------------------------------------------------------------------
# loss / metrics
# ----------------------------------------------------------------------------------------------
def categorical_crossentropy(target_onehot, probs):
    """Keras backend.categorical_crossentropy(from_logits=False), reduction none (main.py:107-110)."""
    p = probs / probs.sum(dim=-1, keepdim=True)
    p = torch.clamp(p, CE_EPS, 1.0 - CE_EPS)
    return -(target_onehot * torch.log(p)).sum(dim=-1)


def ce_and_accuracy(target_ids, probs, V):
    """lc_NIC.py:370-375 : per step mean over batch, then mean over steps."""
    B, T = target_ids.shape
    onehot = torch.nn.functional.one_hot(target_ids.long(), V).to(probs.dtype)
    ce = 0.0
    acc = 0.0
    for t in range(T):
        ce = ce + categorical_crossentropy(onehot[:, t], probs[:, t]).mean()       # lc_NIC.py:461-466
        acc = acc + (onehot[:, t].argmax(dim=1) == probs[:, t].argmax(dim=1)).to(probs.dtype).sum() / B
    return ce / T, acc / T


def masked_sparse_ce_and_accuracy(target_ids, logits):
    """ThinkAndTell/att_model.py:256-262,306-321 with train.py:262-263's loss object: per step
    reduce_mean( sparse_softmax_cross_entropy_with_logits(real, pred) * (real != 0) ) -- the mean runs over the
    whole batch, masked zeros included -- summed over the steps and divided by target.shape[1] = T + 1 (the
    target of that model carries T + 1 tokens for T loss terms).  Accuracy is reported as in lc_NIC."""
    B, T = target_ids.shape
    ce = 0.0
    acc = 0.0
    for t in range(T):
        y = target_ids[:, t].long()
        lse = torch.logsumexp(logits[:, t], dim=-1)
        loss_ = (lse - logits[:, t].gather(1, y[:, None])[:, 0]) * (y != 0).to(logits.dtype)
        ce = ce + loss_.mean()
        acc = acc + (y == logits[:, t].argmax(dim=1)).to(logits.dtype).sum() / B
    return ce / (T + 1), acc / T


def attention_mse(attn):
    """lc_NIC.py:365-367: reduce_sum over axis=1 of the squeezed (T,B,R) tensor (the BATCH axis), MSE vs 1."""
    s = attn.squeeze(-1).sum(dim=1)
    return ((1.0 - s) ** 2).mean()


def attention_mse_elementwise(attn):
    """train_step_sam (lc_NIC.py:751-752): MSE(ones(attention_scores.shape), attention_scores)."""
    return ((1.0 - attn) ** 2).mean()


def l2_loss(cfg, P):
    total = 0.0
    for name in trainable_names(cfg):
        lam = l2_lambda(cfg, name)
        if lam > 0:
            ws = P[name] if isinstance(P[name], list) else [P[name]]
            for w in ws:
                total = total + lam * (w ** 2).sum()
    return total


def make_target(cap):
    """data_generator_guse.py:160-162: shift left by one, trailing 0 (ids; one-hot is applied lazily)."""
    tgt = np.zeros_like(cap)
    tgt[:, :-1] = cap[:, 1:]
    return tgt


def eval_step(cfg, P_np, groups, batch, training=False, seed=0, dtype=torch.float32):
    """test_step (lc_NIC.py:410-459 / ms2_NIC.py:376-465)."""
    P = to_torch(P_np, dtype)
    x, cap, a0, c0, tgt = _batch_to_torch(batch, dtype)
    with torch.no_grad():
        out = forward(cfg, P, groups, (x, cap, a0, c0), training or cfg.n_enc == 2, seed)
        return _metrics(cfg, P, out, tgt)


def _metrics(cfg, P, out, tgt):
    if cfg.n_enc == 1:
        if cfg.masked_sparse_ce:
            ce, acc = masked_sparse_ce_and_accuracy(tgt, out["logits"])
        else:
            ce, acc = ce_and_accuracy(tgt, out["probs"], cfg.V)
        return dict(loss=ce, L2=l2_loss(cfg, P), accuracy=acc, attention=attention_mse(out["attn"]))
    h = tgt.shape[0] // 2
    ceA, accA = ce_and_accuracy(tgt[:h], out["probs"][:h], cfg.V)
    ceB, accB = ce_and_accuracy(tgt[h:], out["probs"][h:], cfg.V)
    return dict(loss=(ceA + ceB) / 2.0, lossA=ceA, lossB=ceB, L2=l2_loss(cfg, P), accuracyA=accA, accuracyB=accB,
                attentionA=attention_mse(out["attn"][:, :h]), attentionB=attention_mse(out["attn"][:, h:]))


def _batch_to_torch(batch, dtype):
    x, cap, a0, c0, tgt = batch
    return (torch.as_tensor(np.asarray(x), dtype=dtype), torch.as_tensor(np.asarray(cap)).long(),
            torch.as_tensor(np.asarray(a0), dtype=dtype), torch.as_tensor(np.asarray(c0), dtype=dtype),
            torch.as_tensor(np.asarray(tgt)).long())


# ----------------------------------------------------------------------------------------------
# train step (GradientTape + Adam.apply_gradients)
# ----------------------------------------------------------------------------------------------
def new_opt_state(P_np):
    z = lambda v: [np.zeros_like(a) for a in v] if isinstance(v, list) else np.zeros_like(v)
    return dict(t=0, m={k: z(v) for k, v in P_np.items()}, v={k: z(v) for k, v in P_np.items()})


def loss_and_grads(cfg, P_np, groups, batch, seed=0, dtype=torch.float32, training=True):
    """Forward + tape.gradient(total_loss, trainable_variables) (lc_NIC.py:354-387).
    Returns (metrics dict of floats, grads dict of numpy arrays [dense; emb is the scatter-added grad],
    emb_values_sqnorm (float: sum of squares of the un-aggregated IndexedSlices values), bn batch stats)."""
    P = to_torch(P_np, dtype)
    names = trainable_names(cfg)
    leaves = []
    for n in names:
        vs = P[n] if isinstance(P[n], list) else [P[n]]
        for v in vs:
            v.requires_grad_(True)
            leaves.append(v)
    x, cap, a0, c0, tgt = _batch_to_torch(batch, dtype)
    stats = {}
    out = forward(cfg, P, groups, (x, cap, a0, c0), training, seed, stats)
    met = _metrics(cfg, P, out, tgt)
    total = met["loss"] + met["L2"]
    if cfg.attn_loss == "batch":                                          # lc_NIC.py:384
        total = total + met["attention"]
    elif cfg.attn_loss == "elementwise":                                  # lc_NIC.py:774
        met["attention_elementwise"] = attention_mse_elementwise(out["attn"])
        total = total + met["attention_elementwise"]
    total.backward()
    grads = {}
    for n in names:
        if isinstance(P[n], list):
            grads[n] = [(v.grad if v.grad is not None else torch.zeros_like(v)).detach().numpy().astype(np.float64
                        if dtype == torch.float64 else np.float32) for v in P[n]]
        else:
            g = P[n].grad if P[n].grad is not None else torch.zeros_like(P[n])
            grads[n] = g.detach().numpy().astype(np.float64 if dtype == torch.float64 else np.float32)
    emb_sq = float(sum((ev.grad.double() ** 2).sum() for ev in out["emb_values"]))
    met = {k: float(v.detach()) if torch.is_tensor(v) else float(v) for k, v in met.items()}
    stats = {k: v.numpy() for k, v in stats.items()}
    return met, grads, emb_sq, stats, out


def clip_by_norm_scale(sqnorm, clipnorm):
    """tf.clip_by_norm: t * clip / max(||t||, clip)."""
    nrm = math.sqrt(sqnorm) if sqnorm > 0 else 0.0
    return clipnorm / max(nrm, clipnorm)


def unitwise_norm(x):
    """agc.py:6-18: vectors -> one norm; 2-D (in, out) -> per output column, keepdims."""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim <= 1:
        return np.sqrt((x ** 2).sum())
    return np.sqrt((x ** 2).sum(axis=0, keepdims=True))


def adaptive_clip_grad(cfg, P_np, grads, emb_values):
    """agc.adaptive_clip_grad (agc.py:20-38) over (trainable_variables, gradients).  `emb_values` = the embedding's
    IndexedSlices values (B*T, E): its unit-wise gradient norm is taken over them (agc.py:24-29) and the clipped
    slices are what the optimizer later sums / takes its clip norm over.  Returns (new grads, new emb_sq)."""
    clip, eps = cfg.agc
    out = {}
    emb_sq = None
    for n in trainable_names(cfg):
        is_list = isinstance(P_np[n], list)
        ws = P_np[n] if is_list else [P_np[n]]
        gs = grads[n] if is_list else [grads[n]]
        res = []
        for w, g in zip(ws, gs):
            p_norm = unitwise_norm(w)
            max_norm = np.maximum(p_norm, eps) * clip
            if n == "emb":
                vals = np.asarray(emb_values, dtype=np.float64)
                g_norm = unitwise_norm(vals)
                fac = np.where(g_norm < max_norm, 1.0, max_norm / np.maximum(g_norm, 1e-6))
                emb_sq = float(((vals * fac) ** 2).sum())
            else:
                g_norm = unitwise_norm(g)
                fac = np.where(g_norm < max_norm, 1.0, max_norm / np.maximum(g_norm, 1e-6))
            res.append((np.asarray(g, dtype=np.float64) * fac).astype(np.float32))
        out[n] = res if is_list else res[0]
    return out, emb_sq


def adam_apply(cfg, P_np, opt, grads, emb_sq):
    """optimizer.apply_gradients (lc_NIC.py:389) with per-variable clipnorm (main.py:97).  In place."""
    opt["t"] += 1
    t = opt["t"]
    alpha = cfg.lr * math.sqrt(1.0 - cfg.beta2 ** t) / (1.0 - cfg.beta1 ** t)
    for n in trainable_names(cfg):
        is_list = isinstance(P_np[n], list)
        ws = P_np[n] if is_list else [P_np[n]]
        gs = grads[n] if is_list else [grads[n]]
        ms = opt["m"][n] if is_list else [opt["m"][n]]
        vs = opt["v"][n] if is_list else [opt["v"][n]]
        for w, g, m, v in zip(ws, gs, ms, vs):
            g = g.astype(np.float32)
            if cfg.apply_clipnorm and cfg.clipnorm is not None:
                if n == "emb" and cfg.emb_unaggregated_norm:
                    sq = emb_sq
                else:
                    sq = float((g.astype(np.float64) ** 2).sum())
                g = g * np.float32(cfg.clipnorm) / np.float32(max(math.sqrt(sq), cfg.clipnorm))
            m += (g - m) * np.float32(1.0 - cfg.beta1)
            v += (g * g - v) * np.float32(1.0 - cfg.beta2)
            w -= np.float32(alpha) * m / (np.sqrt(v) + np.float32(cfg.adam_eps))


def update_moving_stats(cfg, P_np, stats):
    for e in range(cfg.n_enc):
        for k in ("mean", "var"):
            key = f"enc{e}_{k}"
            if key in stats:
                P_np[key] = (P_np[key] * np.float32(BN_MOMENTUM)
                             + stats[key].astype(np.float32) * np.float32(1.0 - BN_MOMENTUM))


def train_step(cfg, P_np, opt, groups, batch, seed=0, dtype=torch.float32):
    """lc_NIC.train_step (lc_NIC.py:328-408) / ms2_NIC.train_step (ms2_NIC.py:294-374).  Mutates P_np, opt."""
    met, grads, emb_sq, stats, out = loss_and_grads(cfg, P_np, groups, batch, seed, dtype)
    if cfg.agc:                                                           # lc_NIC.py:388
        vals = np.concatenate([ev.grad.reshape(-1, cfg.E).numpy() for ev in out["emb_values"]])
        grads_a, emb_sq = adaptive_clip_grad(cfg, P_np, grads, vals)
        adam_apply(cfg, P_np, opt, grads_a, emb_sq)
    else:
        adam_apply(cfg, P_np, opt, grads, emb_sq)
    update_moving_stats(cfg, P_np, stats)
    met["lr"] = cfg.lr
    return met, grads


def train_step_sam(cfg, P_np, opt, groups, batch, seed=0, rho=0.05, eps=1e-12, dtype=torch.float32):
    """train_step_sam (lc_NIC.py:713-838).  First pass: loss + L2 + element-wise attention MSE; e_w = g * rho /
    (global_norm(g) + eps) with the embedding entering through its IndexedSlices values (tf.linalg.global_norm) and
    applied as their scatter (lc_NIC.py:788-789); second pass at w + e_w without the attention loss, different
    dropout draws (seed + 1: two separate model calls); w -= e_w; apply_gradients(second gradients).  The moving
    statistics move twice (two training calls).  Mutates P_np, opt."""
    c1 = Config(**{**cfg.__dict__, "attn_loss": "elementwise"})
    _, g1, emb_sq1, stats1, _ = loss_and_grads(c1, P_np, groups, batch, seed, dtype)
    update_moving_stats(cfg, P_np, stats1)
    sq = 0.0
    for n in trainable_names(cfg):
        if n == "emb":
            sq += emb_sq1
        else:
            gs = g1[n] if isinstance(g1[n], list) else [g1[n]]
            sq += float(sum((np.asarray(g, dtype=np.float64) ** 2).sum() for g in gs))
    scale = np.float32(rho / (np.float32(math.sqrt(sq)) + np.float32(eps)))
    e_ws = {}
    for n in trainable_names(cfg):
        if isinstance(P_np[n], list):
            e_ws[n] = [(g.astype(np.float32) * scale) for g in g1[n]]
            for w, e in zip(P_np[n], e_ws[n]):
                w += e
        else:
            e_ws[n] = g1[n].astype(np.float32) * scale
            P_np[n] += e_ws[n]
    c2 = Config(**{**cfg.__dict__, "attn_loss": ""})
    met, g2, emb_sq2, stats2, _ = loss_and_grads(c2, P_np, groups, batch, seed + 1, dtype)
    for n in trainable_names(cfg):
        if isinstance(P_np[n], list):
            for w, e in zip(P_np[n], e_ws[n]):
                w -= e
        else:
            P_np[n] -= e_ws[n]
    adam_apply(cfg, P_np, opt, g2, emb_sq2)
    update_moving_stats(cfg, P_np, stats2)
    met["attention"] = 0.0                                                # lc_NIC.py:800: reset, not recomputed
    met["lr"] = cfg.lr
    return met, g2


# ----------------------------------------------------------------------------------------------
# greedy decode
# ----------------------------------------------------------------------------------------------
def choose_words(probs, step, top_k=0, top_p=0.0, seed=0):
    """The random word choice of sample_choice (tf.random.categorical(log p), lc_NIC.py:571-575), top-k and nucleus
    (img_NIC.py:509-533; built as intended: the k MOST likely words / the shortest most-likely prefix whose mass
    exceeds p -- the committed helpers sort ascending).  TF's RNG cannot be replayed, so the draw follows the shared
    contract of include/tnt.h (tnt_sample): sort descending (ties: lower id first), keep n words, inverse CDF with
    the row's Philox uniform over fp64 running sums.  probs (B,V) fp32 -> ids (B,), margin (B,) = distance of the
    threshold to the nearest running-sum boundary relative to the kept mass (a tiny margin marks a near-tie)."""
    p = np.asarray(probs, dtype=np.float32)
    B, V = p.shape
    u = philox.sample_uniform(B, step, seed)
    ids = np.zeros(B, np.int64)
    margin = np.zeros(B)
    for b in range(B):
        order = np.lexsort((np.arange(V), -p[b].astype(np.float64)))      # descending p, ties: lower id first
        cs = np.cumsum(p[b][order].astype(np.float64))
        n = V
        if top_k and top_k > 0:
            n = min(n, int(top_k))
        if 0.0 < top_p < 1.0:
            over = np.nonzero(cs > np.float32(top_p))[0]
            if over.size:
                n = min(n, int(over[0]) + 1)
        thr = u[b] * cs[n - 1]
        hit = np.nonzero(cs[:n] > thr)[0]
        j = int(hit[0]) if hit.size else n - 1
        ids[b] = order[j]
        margin[b] = min(abs(cs[j] - thr), abs(thr - (cs[j - 1] if j else 0.0))) / cs[n - 1]
    return ids, margin


def greedy_predict(cfg, P_np, groups, betas, a0, c0, start_seq, max_len, dtype=torch.float32, enc=0, sample=None):
    """greedy_predict_attention (lc_NIC.py:577-638); inference mode (BN moving stats, no dropout).
    Returns outputs (B,max_len,1) int64, outputs_raw (B,max_len,V), attention_scores (max_len,B,R,1),
    attention_weights (max_len,B,R,A).  sample = (top_k, top_p, seed): random word choice (choose_words) instead of
    arg-max; a fifth return value then lists the margins (max_len, B)."""
    P = to_torch(P_np, dtype)
    with torch.no_grad():
        x = torch.as_tensor(np.asarray(betas), dtype=dtype)
        features, _ = encoder(cfg, P, enc, groups, x, False, 0)            # lc_NIC.py:593-594
        word = torch.as_tensor(np.asarray(start_seq)).long()
        text = P["emb"][word]                                              # lc_NIC.py:596
        a = torch.as_tensor(np.asarray(a0), dtype=dtype)
        c = torch.as_tensor(np.asarray(c0), dtype=dtype)
        outs, raws, scores, hid, margins = [], [], [], [], []
        for i in range(max_len):                                           # lc_NIC.py:607
            ctx, w, s = attention(cfg, P, a, features, i, False, 0)
            scores.append(w.numpy()); hid.append(s.numpy())
            xin = torch.cat([ctx, text], dim=-1)
            a, c = lstm_step(cfg, P, xin, a, c, i, False, 0)
            inter = lrelu(a @ P["inter_W"] + P["inter_b"])
            probs = torch.softmax(inter @ P["out_W"] + P["out_b"], dim=-1)  # (B,V)
            raws.append(probs.numpy())
            if sample is None:
                word = torch.as_tensor(np.argmax(probs.numpy(), axis=-1))   # lc_NIC.py:627 first max index
            else:
                w_, m_ = choose_words(probs.numpy(), i, sample[0], sample[1], sample[2])
                word = torch.as_tensor(w_); margins.append(m_)
            outs.append(word.numpy())
            text = P["emb"][word]                                           # lc_NIC.py:632
    outputs = np.stack(outs, axis=1)[:, :, None].astype(np.int64)
    outputs_raw = np.stack(raws, axis=1)
    if sample is not None:
        return outputs, outputs_raw, np.array(scores), np.array(hid), np.array(margins)
    return outputs, outputs_raw, np.array(scores), np.array(hid)


def beam_search(cfg, P_np, groups, betas, start_seq, max_len, beam=3, end_id=-1, a0=None, c0=None,
                dtype=torch.float32, enc=0):
    """Beam search over the decode step of greedy_predict_attention -- the intent of lc_NIC.beam_search
    (lc_NIC.py:640-692; the committed function calls `self.attention` with the wrong arity and never runs), specified
    as include/tnt.h states it: zero initial state (`:651-652`) unless given, scores = sums of log-probabilities
    (log softmax of the dense_out pre-activation), no length normalisation, every beam of a sample starts from the
    same prefix (only beam 0 is live at step 0), a beam that produced end_id is frozen (continues with id 0 at no
    cost), candidates ordered by (score desc, parent beam asc, word asc).
    Returns sequences (B, beam, max_len) int64 best first, scores (B, beam) float32."""
    P = to_torch(P_np, dtype)
    NEG = np.float32(-1.0e30)
    with torch.no_grad():
        x = torch.as_tensor(np.asarray(betas), dtype=dtype)
        B = x.shape[0]
        feats, _ = encoder(cfg, P, enc, groups, x, False, 0)
        W = beam
        feats = feats.repeat_interleave(W, dim=0)                          # row b*W + w
        z = torch.zeros((B, cfg.U), dtype=dtype)
        a = (z if a0 is None else torch.as_tensor(np.asarray(a0), dtype=dtype)).repeat_interleave(W, dim=0)
        c = (z if c0 is None else torch.as_tensor(np.asarray(c0), dtype=dtype)).repeat_interleave(W, dim=0)
        word = torch.as_tensor(np.asarray(start_seq)).long().reshape(-1).repeat_interleave(W)
        score = np.full((B, W), NEG, np.float32); score[:, 0] = 0.0
        fin = np.zeros((B, W), bool)
        ids = np.zeros((B, W, max_len), np.int64)
        for i in range(max_len):
            text = P["emb"][word]
            ctx, _, _ = attention(cfg, P, a, feats, i, False, 0)
            a, c = lstm_step(cfg, P, torch.cat([ctx, text], dim=-1), a, c, i, False, 0)
            inter = lrelu(a @ P["inter_W"] + P["inter_b"])
            logits = (inter @ P["out_W"] + P["out_b"]).numpy().astype(np.float32)
            mx = logits.max(axis=-1, keepdims=True)
            lse = mx[:, 0] + np.log(np.exp(logits - mx).sum(axis=-1, dtype=np.float32))
            logp = logits - lse[:, None]
            new_score = np.zeros_like(score); new_fin = np.zeros_like(fin)
            new_ids = np.zeros_like(ids); parent = np.zeros((B, W), np.int64); new_word = np.zeros((B, W), np.int64)
            for b in range(B):
                cand = []                                                   # (score, parent, rank, word)
                for w in range(W):
                    row = b * W + w
                    if fin[b, w]:
                        cand.append((score[b, w], w, 0, 0))
                        cand += [(NEG, w, k, 0) for k in range(1, W)]
                    elif score[b, w] <= NEG:
                        cand += [(NEG, w, k, 0) for k in range(W)]
                    else:
                        order = np.lexsort((np.arange(cfg.V), -logits[row].astype(np.float64)))[:W]
                        cand += [(np.float32(score[b, w] + logp[row, v]), w, k, int(v)) for k, v in enumerate(order)]
                cand.sort(key=lambda t: (-float(t[0]), t[1], t[2]))
                for k in range(W):
                    sc, pw, _, v = cand[k]
                    new_score[b, k] = sc; parent[b, k] = pw; new_word[b, k] = v
                    new_fin[b, k] = fin[b, pw] or v == end_id
                    new_ids[b, k] = ids[b, pw]; new_ids[b, k, i] = v
            flat_parent = torch.as_tensor((np.arange(B)[:, None] * W + parent).reshape(-1))
            a, c = a[flat_parent], c[flat_parent]
            word = torch.as_tensor(new_word.reshape(-1))
            score, fin, ids = new_score, new_fin, new_ids
    return ids, score
