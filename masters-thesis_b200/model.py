"""Python mirror of the reference's model surface for the hot path.

`NIC` keeps the constructor, `compile`, `__call__`, `train_step`, `test_step`, `greedy_predict`,
`fit`, `get_weights/set_weights`, `get_layer`, `save_weights/load_weights` and `summary` of
`class NIC(tf.keras.Model)` (AttemptFour/Model/lc_NIC.py:36-638; ms2_NIC.py for n_enc=2) so that
main.py:113-134,147,326,342 / eval.py:140,163 style drivers run unchanged.  All arithmetic happens in
libtnt.so (hand-written sm_100a CUDA) through the C ABI of include/tnt.h; torch is used only for
device memory, streams and torch.distributed.  There is no CPU path.
"""
import ctypes as C
import os
import numpy as np
import torch

from . import _lib
from .init import init_variable


class Adam:
    """Duck type of tf.keras.optimizers.Adam(learning_rate, beta_1, beta_2, epsilon, clipnorm) (main.py:97)."""

    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, clipnorm=None, **_):
        self.learning_rate = float(learning_rate)
        self.beta_1, self.beta_2, self.epsilon = float(beta_1), float(beta_2), float(epsilon)
        self.clipnorm = clipnorm
        self.iterations = 0

    @property
    def lr(self):
        return self.learning_rate

    @lr.setter
    def lr(self, v):
        self.learning_rate = float(v)


def _opt_field(opt, names, default):
    for n in names:
        if isinstance(opt, dict) and n in opt:
            return opt[n]
        if hasattr(opt, n):
            v = getattr(opt, n)
            try:
                return float(v.numpy()) if hasattr(v, "numpy") else (None if v is None else float(v))
            except TypeError:
                continue
    return default


PRECISIONS = {"fp32": 0, "bf16x3": 1, "bf16": 2}          # enum tnt_precision
ATTN_LOSS = {None: 0, False: 0, "batch": 1, "elementwise": 2}   # enum tnt_attn_loss


class _CudaView:
    """__cuda_array_interface__ holder so torch can view a library-owned device buffer (DP exchange buffers)."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


class _Layer:
    """What `model.get_layer(name)` returns: get_weights / set_weights over the layer's variables
    (main.py:161-162)."""

    def __init__(self, model, prefix):
        self._m, self._prefix = model, prefix
        self.name = prefix

    def _vars(self):
        vs = [v for v in self._m._vars if v["name"].startswith(self._prefix + "/") or
              ("/" + self._prefix + "/") in v["name"]]
        if not vs:
            raise ValueError(f"No such layer: {self._prefix}")
        return vs

    def get_weights(self):
        return [self._m._read_var(v) for v in self._vars()]

    def set_weights(self, weights):
        vs = self._vars()
        if len(vs) != len(weights):
            raise ValueError(f"layer {self._prefix} expects {len(vs)} arrays, got {len(weights)}")
        for v, a in zip(vs, weights):
            self._m._write_var(v, a)


class NIC:
    """Drop-in for lc_NIC.NIC (n_enc=1) and ms2_NIC.NIC (n_enc=2)."""

    def __init__(self, groups, units, embedding_features, embedding_text, attn_units, vocab_size, max_length,
                 dropout_input, dropout_features, dropout_text, dropout_attn, dropout_lstm, dropout_out,
                 input_reg, attn_reg, lstm_reg, output_reg, *, n_enc=1, batch_max=64, seed=42, device=None,
                 inter_units=256, process_group=None, masked_sparse_ce=False, plain_attention=False,
                 precision="fp32", attention_loss=None, agc=None):
        """Keyword-only extras (none of them exists as a constructor argument upstream; each stands for an edit of
        the reference source): `precision` = "fp32" | "bf16x3" | "bf16" (GEMM operand precision, include/tnt.h);
        `attention_loss` = None | "batch" (un-comment lc_NIC.py:384) | "elementwise" (train_step_sam's form);
        `agc` = None | (clip_factor, eps) (un-comment lc_NIC.py:388: agc.adaptive_clip_grad)."""
        self._libh = _lib.load()                       # raises if libtnt.so is missing: no fallback
        in_groups, out_groups = groups
        assert len(in_groups) == len(out_groups), "Input groups don't match ouput groups"   # layers.py:30
        if len(set(int(o) for o in out_groups)) != 1:
            raise ValueError("all parcels must share one output width (group_size)")
        self.in_groups = [np.ascontiguousarray(np.asarray(g, dtype=np.int32)) for g in in_groups]
        self.R, self.G = len(in_groups), int(out_groups[0])
        self.units, self.E, self.A = int(units), int(embedding_text), int(attn_units)
        self.embedding_features = embedding_features      # accepted, unused (as in lc_NIC)
        self.V, self.T, self.H = int(vocab_size), int(max_length), int(inter_units)
        self.n_enc = int(n_enc)
        # ThinkAndTell/att_model.py variant: masked sparse CE on logits; bare attention projections
        self.masked_sparse_ce, self.plain_attention = bool(masked_sparse_ce), bool(plain_attention)
        if self.masked_sparse_ce and self.n_enc != 1:
            raise ValueError("the masked sparse-CE variant is defined for the single-encoder model only")
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        self.precision = precision
        if attention_loss not in ATTN_LOSS:
            raise ValueError(f"attention_loss must be one of {list(ATTN_LOSS)}")
        if attention_loss and self.n_enc != 1:
            raise ValueError("the attention-loss switch is defined for the single-encoder model (lc_NIC.py:384)")
        self.attention_loss = attention_loss
        self.agc = None if agc is None else (float(agc[0]), float(agc[1]))
        self.dropouts = (dropout_input, dropout_features, dropout_text, dropout_attn, dropout_lstm, dropout_out)
        self.regs = (input_reg, attn_reg, lstm_reg, output_reg)
        self.device = torch.device(device if device is not None else "cuda")
        self.batch_max = int(batch_max)
        self._seed = int(seed)
        self._step = 0
        self._ctx = None
        self._N = None
        self.optimizer = Adam(1e-3)
        self.loss = None
        self._pg = process_group
        self._copy_stream = None
        self._slot_done = [None, None]
        self._slot = 0
        self._stage = {}
        self._stage_slots = [{}, {}]
        self._pf = []
        self._bucket_cache = {}
        self._buf_cache = {}
        self._dp_ready = False
        self._deferred = False
        self.stop_training = False

    # ---- Keras-like plumbing -------------------------------------------------------------------
    def compile(self, optimizer=None, loss=None, run_eagerly=True, **_):
        """main.py:134.  Only Adam(+clipnorm) with the probabilities-CE loss of main.py:107-110 is on the
        hot path; anything else is refused loudly."""
        if optimizer is not None:
            self.optimizer = optimizer
        self.loss = loss
        # the loss object only selects the CE variant: SparseCategoricalCrossentropy(from_logits=True)
        # (ThinkAndTell/train.py:262-263) or the string "masked_sparse_ce" -> att_model.py:306-321
        if loss == "masked_sparse_ce" or getattr(loss, "from_logits", False):
            self.masked_sparse_ce = True
        if self._ctx is not None:
            self._rebuild(self.batch_max)

    def _opt(self):
        o = self.optimizer
        lr = _opt_field(o, ("learning_rate", "lr"), 1e-3)
        b1 = _opt_field(o, ("beta_1",), 0.9)
        b2 = _opt_field(o, ("beta_2",), 0.999)
        eps = _opt_field(o, ("epsilon",), 1e-7)
        cn = _opt_field(o, ("clipnorm",), None)
        return lr, b1, b2, eps, (cn if cn is not None else -1.0)

    @property
    def world(self):
        import torch.distributed as dist
        return dist.get_world_size(self._pg) if (dist.is_available() and dist.is_initialized()) else 1

    def _rebuild(self, batch_max, N=None):
        lib = self._libh
        N = int(N if N is not None else self._N)
        old = None
        if self._ctx is not None:
            old = (self._w, self._m, self._v, self._state, lib.tnt_get_adam_step(self._ctx))
            lib.tnt_destroy(self._ctx)                   # synchronises the side stream
            self._ctx = None
            self._deferred = False
        self._buf_cache = {}
        self._dp_ready = False
        lr, b1, b2, eps, cn = self._opt()
        cfg = _lib.TntConfig(N=N, R=self.R, G=self.G, U=self.units, E=self.E, A=self.A, H=self.H, V=self.V, T=self.T,
                             B_max=int(batch_max), n_enc=self.n_enc, precision=PRECISIONS[self.precision],
                             dropout_input=self.dropouts[0], dropout_features=self.dropouts[1],
                             dropout_text=self.dropouts[2], dropout_attn=self.dropouts[3],
                             dropout_lstm=self.dropouts[4], dropout_out=self.dropouts[5],
                             input_reg=self.regs[0], attn_reg=self.regs[1], lstm_reg=self.regs[2],
                             output_reg=self.regs[3], lr=lr, beta1=b1, beta2=b2, adam_eps=eps, clipnorm=cn,
                             emb_unaggregated_norm=1, second_feature_dropout=1 if self.n_enc == 2 else 0,
                             variant=(1 if self.masked_sparse_ce else 0) | (2 if self.plain_attention else 0),
                             attn_loss=ATTN_LOSS[self.attention_loss],
                             agc_clip=self.agc[0] if self.agc else 0.0, agc_eps=self.agc[1] if self.agc else 0.0)
        idx = np.concatenate(self.in_groups * self.n_enc).astype(np.int32)
        sizes = np.array([len(g) for g in self.in_groups] * self.n_enc, dtype=np.int64)
        off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        ctx = C.c_void_p()
        with torch.cuda.device(self.device):
            rc = lib.tnt_create(C.byref(cfg), idx.ctypes.data, off.ctypes.data, C.byref(ctx))
        _lib.check(lib, None, rc)
        self._ctx, self._N, self.batch_max = ctx, N, int(batch_max)
        n, ns = lib.tnt_param_floats(ctx), lib.tnt_state_floats(ctx)
        self._vars = []
        info = _lib.TntVarInfo()
        for i in range(lib.tnt_num_vars(ctx)):
            _lib.check(lib, ctx, lib.tnt_var_info_get(ctx, i, C.byref(info)))
            self._vars.append(dict(name=info.name.decode(), offset=info.offset, rows=info.rows, cols=info.cols,
                                   l2=info.l2, trainable=bool(info.trainable)))
        if old is None:
            z = lambda k: torch.zeros(k, dtype=torch.float32, device=self.device)
            self._w, self._m, self._v, self._g, self._state = z(n), z(n), z(n), z(n), z(max(ns, 4))
            rng = np.random.default_rng(self._seed)
            w_host = np.zeros(n, np.float32)
            s_host = np.zeros(max(ns, 4), np.float32)
            for v in self._vars:
                a = init_variable(rng, v["name"], v["rows"], v["cols"], self.units)
                tgt = w_host if v["trainable"] else s_host
                tgt[v["offset"]:v["offset"] + a.size] = a.reshape(-1)
            self._w.copy_(torch.from_numpy(w_host))
            self._state.copy_(torch.from_numpy(s_host))
        else:
            self._w, self._m, self._v, self._state, t = old
            self._g = torch.zeros_like(self._w)
            self._bucket_cache = {}
            lib.tnt_set_adam_step(ctx, t)
        _lib.check(lib, ctx, lib.tnt_bind_params(ctx, self._w.data_ptr(), self._m.data_ptr(), self._v.data_ptr(),
                                                  self._g.data_ptr(), self._state.data_ptr()))
        self._stage = {}
        self._stage_slots = [{}, {}]
        self._pf = []
        self.built = True

    def _ensure(self, B, N):
        if self._ctx is None:
            self._rebuild(max(self.batch_max, B), N)
        elif N != self._N:
            raise ValueError(f"input width {N} != {self._N} the model was built with")
        elif B > self.batch_max:
            self._rebuild(B)

    def __del__(self):
        try:
            if self._ctx is not None:
                self._libh.tnt_destroy(self._ctx)
        except Exception:
            pass

    # ---- host <-> device staging ----------------------------------------------------------------
    def _dev(self, name, x, dtype):
        """Device tensor holding `x`: CUDA tensors pass through zero-copy; numpy / CPU tensors are copied
        (asynchronously when the source is pinned) into a persistent staging buffer on the current stream."""
        if isinstance(x, torch.Tensor) and x.is_cuda:
            return x.to(dtype).contiguous()
        if not isinstance(x, torch.Tensor):
            a = np.asarray(x)
            if hasattr(x, "numpy") and not isinstance(x, np.ndarray):
                a = np.asarray(x.numpy())
            x = torch.from_numpy(np.ascontiguousarray(a))
        if x.dtype != dtype:
            x = x.to(dtype)
        buf = self._stage.get(name)
        if buf is None or buf.shape != x.shape or buf.dtype != dtype:
            buf = torch.empty(x.shape, dtype=dtype, device=self.device)
            self._stage[name] = buf
        buf.copy_(x, non_blocking=True)
        return buf

    def prefetch(self, data):
        """Start the host->device copies of a LATER batch on a side stream while the current step computes --
        the job Keras' fit() gives its enqueuer workers (`main.py:269-281`).  `data` is what train_step /
        test_step take; the step that is later called with the SAME object picks the device copies up
        instead of copying again.  Two staging slots alternate, so one batch can be in flight per step."""
        inputs, target = data[0], data[1]
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        slot = self._slot
        self._slot ^= 1
        B, N = int(inputs[0].shape[0]), int(np.prod(inputs[0].shape[1:]))
        self._ensure(B, N)
        cs = self._copy_stream
        if any(isinstance(t, torch.Tensor) and t.is_cuda for t in (*inputs, target)):
            # device-resident batch (dataset.DeviceDataGenerator): its assembly kernels were enqueued on the
            # current stream; the side stream must not touch the tensors before they are written
            produced = torch.cuda.Event()
            produced.record(torch.cuda.current_stream(self.device))
            cs.wait_event(produced)
        for ev in self._slot_done[slot] or ():
            cs.wait_event(ev)                           # the step that last read this slot has finished
        with torch.cuda.stream(cs):
            saved = self._stage
            self._stage = self._stage_slots[slot]
            try:
                B, xb, xc, xa, xs = self._inputs(inputs)
                tgt = self._target_ids(target, B)
            finally:
                self._stage = saved
            ev = torch.cuda.Event()
            ev.record(cs)
        self._pf = [e for e in self._pf if e[1] != slot] + [(data, slot, ev, (B, xb, xc, xa, xs, tgt))]

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _target_ids(self, target, B):
        """The reference hands train_step a one-hot (B,T,V) target (data_generator_guse.py:163); ids (B,T)
        are accepted too.  One-hot input is reduced to ids on the device (first max index)."""
        shape = tuple(target.shape)
        if len(shape) == 3:
            oh = self._dev("target_onehot", target, torch.float32)
            ids = self._stage.get("target_ids")
            if ids is None or ids.shape != (B, self.T):
                ids = torch.empty((B, self.T), dtype=torch.int32, device=self.device)
                self._stage["target_ids"] = ids
            _lib.check(self._libh, self._ctx, self._libh.tnt_onehot_to_ids(self._ctx, oh.data_ptr(), B * self.T,
                                                                            ids.data_ptr(), self._stream()))
            return ids
        return self._dev("target_ids_in", target, torch.int32)

    def _inputs(self, data):
        betas, cap, a0, c0 = data
        B, N = int(betas.shape[0]), int(np.prod(betas.shape[1:]))
        self._ensure(B, N)
        xb = self._dev("betas", betas, torch.float32)
        xc = self._dev("cap", cap, torch.int32)
        xa = self._dev("a0", a0, torch.float32)
        xs = self._dev("c0", c0, torch.float32)
        if tuple(xc.shape) != (B, self.T):
            raise ValueError(f"caption tensor must be (B,{self.T}), got {tuple(xc.shape)}")
        return B, xb, xc, xa, xs

    def _buffer(self, name, typestr, itemsize):
        """Torch view of a library-owned device buffer; cached per context (the data-parallel step asks for four of
        them every step and the ctypes + __cuda_array_interface__ round trip costs ~50 us of host time each)."""
        key = (name, typestr)
        hit = self._buf_cache.get(key)
        if hit is not None:
            return hit
        t = self._buffer_uncached(name, typestr, itemsize)
        self._buf_cache[key] = t
        return t

    def _buffer_uncached(self, name, typestr, itemsize):
        p, nb = C.c_void_p(), C.c_int64()
        _lib.check(self._libh, self._ctx, self._libh.tnt_buffer(self._ctx, name.encode(), C.byref(p), C.byref(nb)))
        return torch.as_tensor(_CudaView(p.value, nb.value // itemsize, typestr), device=self.device)

    def debug_buffer(self, name, dtype="f4"):
        """Internal activation by name (tests)."""
        t = self._buffer_uncached(name, "<" + dtype, int(dtype[1:]))
        return t.cpu().numpy()

    # ---- the hot path ---------------------------------------------------------------------------
    def __call__(self, data, training=False):
        """model(data, training) -> (probs (B,T,V), attention_scores (T,B,R,1))  (lc_NIC.py:163,223-263).
        n_enc == 2 returns (predA, attnA, predB, attnB) and always runs training=True (ms2_NIC.py:177-205)."""
        B, xb, xc, xa, xs = self._inputs(data)
        if self.n_enc == 2:
            training = True
        probs = torch.empty((B, self.T, self.V), dtype=torch.float32, device=self.device)
        attn = torch.empty((self.T, B, self.R, 1), dtype=torch.float32, device=self.device)
        seed = self._next_seed()
        _lib.check(self._libh, self._ctx, self._libh.tnt_forward(
            self._ctx, xb.data_ptr(), xc.data_ptr(), xa.data_ptr(), xs.data_ptr(), B, int(bool(training)), seed,
            probs.data_ptr(), attn.data_ptr(), self._stream()))
        p, a = probs.cpu().numpy(), attn.cpu().numpy()
        _lib.check(self._libh, self._ctx, self._libh.tnt_check(self._ctx, self._stream()))
        if self.n_enc == 2:
            h = B // 2
            return p[:h], a[:, :h], p[h:], a[:, h:]
        return p, a

    call = __call__

    def _next_seed(self):
        self._step += 1
        return (self._seed * 0x9E3779B97F4A7C15 + self._step) & 0xFFFFFFFFFFFFFFFF

    def _bucket(self, which):
        """Slice of the flat gradient: 0 encoder, 1 decoder, 2 head, 3 decoder without head (tnt_grad_bucket)."""
        if which not in self._bucket_cache:
            o, n = C.c_int64(), C.c_int64()
            _lib.check(self._libh, self._ctx, self._libh.tnt_grad_bucket(self._ctx, which, C.byref(o), C.byref(n)))
            self._bucket_cache[which] = self._g[o.value:o.value + n.value]
        return self._bucket_cache[which]

    def _dp_setup(self):
        """Data parallelism lives in the library (tnt_dp_init): rank 0 draws the NCCL ids, torch.distributed ships
        them, every rank joins.  Collective: all ranks reach it in their first step after (re)building."""
        import torch.distributed as dist
        world = self.world
        if world == 1 or self._dp_ready:
            return
        rank = dist.get_rank(self._pg)
        buf = np.zeros(256, np.uint8)
        if rank == 0:
            _lib.check(self._libh, self._ctx, self._libh.tnt_dp_unique_id(buf.ctypes.data))
        dev = self.device if dist.get_backend(self._pg) == "nccl" else torch.device("cpu")
        t = torch.from_numpy(buf).to(dev)
        ranks = dist.get_process_group_ranks(self._pg if self._pg is not None else dist.group.WORLD)
        dist.broadcast(t, src=ranks[0], group=self._pg)
        buf = np.ascontiguousarray(t.cpu().numpy())
        torch.cuda.synchronize(self.device)
        _lib.check(self._libh, self._ctx, self._libh.tnt_dp_init(self._ctx, rank, world, buf.ctypes.data, self._stream()))
        self._dp_ready = True

    def _run_step(self, data, train, training, seed=None, apply=True, read_metrics=True, before_apply=None):
        """One tnt_step call: forward, loss, tape gradient, data-parallel exchanges, clip + Adam (include/tnt.h)."""
        lib = self._libh
        (inputs, target) = data[0], data[1]
        pf_slot = None
        hit = [e for e in self._pf if e[0] is data]
        if hit:
            _, pf_slot, ev, (B, xb, xc, xa, xs, tgt) = hit[0]
            self._pf = [e for e in self._pf if e[0] is not data]
            torch.cuda.current_stream(self.device).wait_event(ev)
        else:
            B, xb, xc, xa, xs = self._inputs(inputs)
            tgt = self._target_ids(target, B)
        ctx = self._ctx
        world = self.world
        seed = self._next_seed() if seed is None else int(seed)
        if world > 1:
            import torch.distributed as dist
            self._dp_setup()
            seed = (seed + 0x632BE59BD9B4E019 * dist.get_rank(self._pg)) & 0xFFFFFFFFFFFFFFFF   # rank-local dropout masks
        st = self._stream()
        do_apply = bool(train and apply)
        if do_apply:
            lib.tnt_set_lr(ctx, self._opt()[0])
        late = do_apply and before_apply is not None          # something happens between the gradient and the update
        _lib.check(lib, ctx, lib.tnt_step(ctx, xb.data_ptr(), xc.data_ptr(), tgt.data_ptr(), xa.data_ptr(), xs.data_ptr(),
                                           B, int(training), int(train), int(do_apply and not late), seed, st))
        if late:
            before_apply()
            _lib.check(lib, ctx, lib.tnt_phase_optimizer(ctx, st))
        if do_apply:
            self._deferred = True                # the decoder variables are updated on the library's side stream
            if hasattr(self.optimizer, "iterations") and isinstance(self.optimizer.iterations, int):
                self.optimizer.iterations += 1
        if pf_slot is not None:
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(self.device))
            self._slot_done[pf_slot] = [done]
        if not read_metrics:
            return None
        m = _lib.TntMetrics()
        _lib.check(lib, ctx, lib.tnt_read_metrics(ctx, C.byref(m), st))
        return self._metrics_dict(m, world, None, train)

    def _metrics_dict(self, m, world, attn_override, train):
        """Keys of lc_NIC.py:408 / ms2_NIC.py:366-374 (train) and lc_NIC.py:459 (test)."""
        rows = [m.rows[0], m.rows[1]]              # global under data parallelism (the library reduces the sums)
        att = [m.attn_mse[0], m.attn_mse[1]]
        if self.n_enc == 1:
            # att_model.py:262: the variant divides its T step losses by T + 1
            lrows = rows[0] * (self.T + 1) / self.T if self.masked_sparse_ce else rows[0]
            out = {"loss": m.ce_sum[0] / lrows, "L2": m.l2, "accuracy": m.correct[0] / rows[0], "attention": att[0]}
        else:
            la, lb = m.ce_sum[0] / rows[0], m.ce_sum[1] / rows[1]
            out = {"loss": (la + lb) / 2.0, "lossA": la, "lossB": lb, "L2": m.l2,
                   "accuracyA": m.correct[0] / rows[0], "accuracyB": m.correct[1] / rows[1],
                   "attentionA": att[0], "attentionB": att[1]}
        if train:
            out["lr"] = self._opt()[0]
        out_elem = getattr(m, "attn_elem_mse", 0.0)
        if self.attention_loss == "elementwise":
            out["attention_elementwise"] = out_elem
        return out

    def trace(self, enable=True):
        """In-situ kernel timing: CUDA events on the launching stream around named kernel groups (tnt_trace)."""
        self._libh.tnt_trace(self._ctx, int(bool(enable)))

    def trace_read(self):
        """-> list of (name, ms) in launch order for everything recorded since trace(True)."""
        names = C.create_string_buffer(1 << 16)
        ms = (C.c_float * 4096)()
        n = self._libh.tnt_trace_read(self._ctx, names, len(names), ms, 4096)
        ns = names.value.decode().split("\n")
        return [(ns[i], float(ms[i])) for i in range(max(n, 0))]

    def train_step_async(self, xb, xc, tgt_ids, xa, xs, seed=None):
        """tnt_train_step on DEVICE tensors, nothing read back: the device-resident step bench.py times (one C call per
        step on any number of GPUs; metrics stay in the library's `metrics_raw` buffer)."""
        B = int(xb.shape[0])
        self._ensure(B, int(xb.shape[1]))
        seed = self._next_seed() if seed is None else int(seed)
        if self.world > 1:
            import torch.distributed as dist
            self._dp_setup()
            seed = (seed + 0x632BE59BD9B4E019 * dist.get_rank(self._pg)) & 0xFFFFFFFFFFFFFFFF
        self._libh.tnt_set_lr(self._ctx, self._opt()[0])
        _lib.check(self._libh, self._ctx, self._libh.tnt_train_step(
            self._ctx, xb.data_ptr(), xc.data_ptr(), tgt_ids.data_ptr(), xa.data_ptr(), xs.data_ptr(), B, seed,
            self._stream()))
        self._deferred = True                    # weights must be read through _drain() (tnt_side_join)

    def train_step(self, data, seed=None):
        """lc_NIC.train_step (lc_NIC.py:328-408): data = ((betas, cap, a0, c0), target)."""
        return self._run_step(data, train=True, training=True, seed=seed)

    def test_step(self, data, seed=None):
        """lc_NIC.test_step (lc_NIC.py:410-459); ms2 keeps training=True (ms2_NIC.py:415-427)."""
        return self._run_step(data, train=False, training=(self.n_enc == 2), seed=seed)

    def greedy_predict(self, img_input, a0, c0, start_seq, max_len, units=None, tokenizer=None, training=False,
                       enc=0, return_raw=True, return_attention=True):
        """greedy_predict_attention (lc_NIC.py:577-638) -> (outputs (B,max_len,1) int64, outputs_raw (B,max_len,V),
        attention_scores (max_len,B,R,1), attention_weights (max_len,B,R,A)).  The whole loop runs on the
        device with no host round trip per word."""
        return self._decode(img_input, a0, c0, start_seq, max_len, training, enc, return_raw, return_attention, None)

    def sample_predict(self, img_input, a0, c0, start_seq, max_len, units=None, tokenizer=None, training=False,
                       enc=0, return_raw=True, return_attention=True, top_k=0, top_p=0.0, seed=None):
        """The same decode loop with a random word choice instead of arg-max: `sample_choice`
        (tf.random.categorical(log p), lc_NIC.py:571-575) by default, the k most likely words with `top_k`
        (img_NIC.py:509-513) or the nucleus with `top_p` (img_NIC.py:515-533).  The draw follows the shared
        Philox contract (include/tnt.h: tnt_sample), so a seed reproduces the captions.  Same return tuple."""
        seed = self._next_seed() if seed is None else int(seed)
        return self._decode(img_input, a0, c0, start_seq, max_len, training, enc, return_raw, return_attention,
                            (int(top_k), float(top_p), seed))

    def _decode(self, img_input, a0, c0, start_seq, max_len, training, enc, return_raw, return_attention, sample):
        assert training == False, "training is set to True"   # lc_NIC.py:591
        B, N = int(img_input.shape[0]), int(np.prod(img_input.shape[1:]))
        if B == 0:                                                # ragged remainder of an evenly divided split (eval.py:155)
            return (np.zeros((0, max_len, 1), np.int64),
                    np.zeros((0, max_len, self.V), np.float32) if return_raw else None,
                    np.zeros((max_len, 0, self.R, 1), np.float32) if return_attention else None,
                    np.zeros((max_len, 0, self.R, self.A), np.float32) if return_attention else None)
        self._ensure(min(B, max(self.batch_max, 1)), N)
        outs, raws, atts, hids = [], [], [], []
        lib, ctx = self._libh, self._ctx
        for s in range(0, B, self.batch_max):
            e = min(B, s + self.batch_max)
            n = e - s
            xb = self._dev("g_betas", img_input[s:e], torch.float32)
            xa = self._dev("g_a0", a0[s:e], torch.float32)
            xs = self._dev("g_c0", c0[s:e], torch.float32)
            ss = start_seq[s:e]
            st_ = self._dev("g_start", ss.reshape(-1) if isinstance(ss, torch.Tensor) else np.asarray(ss).reshape(-1),
                            torch.int32)
            ids = torch.empty((n, max_len), dtype=torch.int64, device=self.device)
            raw = torch.empty((n, max_len, self.V), dtype=torch.float32, device=self.device) if return_raw else None
            att = torch.empty((max_len, n, self.R, 1), dtype=torch.float32, device=self.device) if return_attention else None
            hid = torch.empty((max_len, n, self.R, self.A), dtype=torch.float32, device=self.device) if return_attention else None
            ptr = lambda t: t.data_ptr() if t is not None else None
            if sample is None:
                rc = lib.tnt_greedy(ctx, xb.data_ptr(), xa.data_ptr(), xs.data_ptr(), st_.data_ptr(), n, int(max_len),
                                    int(enc), ids.data_ptr(), ptr(raw), ptr(att), ptr(hid), self._stream())
            else:
                # rows of a later chunk draw from their own counters: the row index of the contract is global
                rc = lib.tnt_sample(ctx, xb.data_ptr(), xa.data_ptr(), xs.data_ptr(), st_.data_ptr(), n, int(max_len),
                                    int(enc), sample[0], sample[1], (sample[2] + 0x9E3779B97F4A7C15 * (s // self.batch_max)) & 0xFFFFFFFFFFFFFFFF,
                                    ids.data_ptr(), ptr(raw), ptr(att), ptr(hid), self._stream())
            _lib.check(lib, ctx, rc)
            outs.append(ids.cpu().numpy()[:, :, None])
            if return_raw:
                raws.append(raw.cpu().numpy())
            if return_attention:
                atts.append(att.cpu().numpy()); hids.append(hid.cpu().numpy())
        _lib.check(lib, ctx, lib.tnt_check(ctx, self._stream()))
        outputs = np.concatenate(outs, axis=0)
        assert outputs.shape == (B, max_len, 1)                 # lc_NIC.py:637
        return (outputs, np.concatenate(raws, axis=0) if return_raw else None,
                np.concatenate(atts, axis=1) if return_attention else None,
                np.concatenate(hids, axis=1) if return_attention else None)

    def beam_search(self, features, start_seq, max_len, units=None, beam=3, end_id=None, a0=None, c0=None, enc=0):
        """Beam search over the decode step (what lc_NIC.beam_search, lc_NIC.py:640-692, set out to do; the
        committed function raises on its first attention call).  Zero initial state as there (`:651-652`) unless
        a0/c0 are given.  Scores are sums of log-probabilities; a beam that produced `end_id` is frozen.
        -> (sequences (B, beam, max_len) int64 best first, scores (B, beam) float32)."""
        B, N = int(features.shape[0]), int(np.prod(features.shape[1:]))
        if B * beam > max(self.batch_max, 1) and self._ctx is not None:
            self._rebuild(B * beam)
        self._ensure(B * beam, N)
        lib, ctx = self._libh, self._ctx
        z = np.zeros((B, self.units), np.float32)
        xb = self._dev("g_betas", features, torch.float32)
        xa = self._dev("g_a0", z if a0 is None else a0, torch.float32)
        xs = self._dev("g_c0", z if c0 is None else c0, torch.float32)
        st_ = self._dev("g_start", start_seq.reshape(-1) if isinstance(start_seq, torch.Tensor)
                        else np.asarray(start_seq).reshape(-1), torch.int32)
        ids = torch.empty((B, beam, max_len), dtype=torch.int64, device=self.device)
        sc = torch.empty((B, beam), dtype=torch.float32, device=self.device)
        _lib.check(lib, ctx, lib.tnt_beam_search(ctx, xb.data_ptr(), xa.data_ptr(), xs.data_ptr(), st_.data_ptr(), B,
                                                  int(max_len), int(enc), int(beam), -1 if end_id is None else int(end_id),
                                                  ids.data_ptr(), sc.data_ptr(), self._stream()))
        out = ids.cpu().numpy(), sc.cpu().numpy()
        _lib.check(lib, ctx, lib.tnt_check(ctx, self._stream()))
        return out

    def kernel_variants(self):
        """Which kernels the last calls ran (tnt_kernel_variants): a function of the shapes only."""
        return self._libh.tnt_kernel_variants(self._ctx).decode() if self._ctx is not None else ""

    def train_step_sam(self, data, rho=0.05, eps=1e-12, seed=None):
        """train_step_sam (lc_NIC.py:713-838): gradient at w (loss includes the element-wise attention loss, :774),
        climb to w + rho g/||g||, gradient there (no attention loss), step back, Adam with the second gradient.
        The two passes draw different dropout masks (two model calls): the second uses seed + 1."""
        lib, ctx = self._libh, self._ctx
        seed = self._next_seed() if seed is None else int(seed)
        if self.n_enc != 1:
            raise ValueError("train_step_sam is defined for the single-encoder model")
        self._inputs(data[0])                                        # builds the context on the first call
        lib, ctx = self._libh, self._ctx
        saved = self.attention_loss
        try:
            self.attention_loss = "elementwise"
            _lib.check(lib, ctx, lib.tnt_set_attn_loss(ctx, 2))
            self._run_step(data, train=True, training=True, seed=seed, apply=False)
            _lib.check(lib, ctx, lib.tnt_sam_perturb(ctx, float(rho), float(eps), self._stream()))
            self.attention_loss = None
            _lib.check(lib, ctx, lib.tnt_set_attn_loss(ctx, 0))
            met = self._run_step(data, train=True, training=True, seed=(seed + 1) & 0xFFFFFFFFFFFFFFFF, apply=True,
                                 before_apply=lambda: _lib.check(lib, ctx, lib.tnt_sam_restore(ctx, self._stream())))
        finally:
            self.attention_loss = saved
            lib.tnt_set_attn_loss(ctx, ATTN_LOSS[saved])
        met["attention"] = 0.0                                       # lc_NIC.py:800: reset before the second pass
        return met

    # ---- weights (main.py:155-162,172-190; eval.py:140) -------------------------------------------
    def _drain(self):
        """A deferred decoder-bucket update (data parallel) may still be running on the library's side stream:
        make the current stream wait for it before weights are touched from the host side."""
        if self._deferred and self._ctx is not None:
            _lib.check(self._libh, self._ctx, self._libh.tnt_side_join(self._ctx, self._stream()))
            self._deferred = False

    def _read_var(self, v):
        self._drain()
        src = self._w if v["trainable"] else self._state
        n = v["rows"] * max(v["cols"], 1)
        a = src[v["offset"]:v["offset"] + n].cpu().numpy()
        return a.reshape(v["rows"], v["cols"]) if v["cols"] > 0 else a

    def _write_var(self, v, a):
        self._drain()
        a = np.asarray(a, dtype=np.float32)
        n = v["rows"] * max(v["cols"], 1)
        if a.size != n:
            raise ValueError(f"shape mismatch for {v['name']}: expected {n} values, got {a.shape}")
        dst = self._w if v["trainable"] else self._state
        dst[v["offset"]:v["offset"] + n].copy_(torch.from_numpy(np.ascontiguousarray(a.reshape(-1))))

    def _require_built(self):
        if self._ctx is None:
            raise RuntimeError("model is not built yet: call it once on a batch (main.py:147) or use build(N)")

    def build(self, input_width, batch=None):
        self._rebuild(max(self.batch_max, batch or 1), input_width)

    @property
    def weights(self):
        self._require_built()
        return [v["name"] for v in self._vars]

    def get_weights(self):
        """All variables in Keras order (per layer: trainable then non-trainable)."""
        self._require_built()
        return [self._read_var(v) for v in self._vars]

    def set_weights(self, weights):
        self._require_built()
        if len(weights) != len(self._vars):
            raise ValueError(f"expected {len(self._vars)} arrays, got {len(weights)}")
        for v, a in zip(self._vars, weights):
            self._write_var(v, a)

    def get_named_weights(self):
        self._require_built()
        return {v["name"]: self._read_var(v) for v in self._vars}

    def set_named_weights(self, d, skip_mismatch=False):
        self._require_built()
        for v in self._vars:
            if v["name"] in d:
                try:
                    self._write_var(v, d[v["name"]])
                except ValueError:
                    if not skip_mismatch:
                        raise

    def get_layer(self, name):
        self._require_built()
        return _Layer(self, name)

    def save_weights(self, path, **_):
        """Weights-only checkpoint (main.py:166-190: ModelCheckpoint(..., save_weights_only=True) -> `.h5`).
        `*.h5` / `*.hdf5`: Keras' HDF5 layout (layer_names / weight_names attributes, one group per layer:
        keras_io.py); anything else: `.npz` keyed by variable name."""
        from . import keras_io
        self._require_built()
        if str(path).endswith((".h5", ".hdf5", ".keras.h5")):
            keras_io.save_h5(self, str(path))
        else:
            np.savez(path if str(path).endswith(".npz") else str(path) + ".npz", **self.get_named_weights())

    def load_weights(self, path, by_name=False, skip_mismatch=False, **_):
        """Keras semantics (eval.py:140 passes by_name=True, skip_mismatch=True): a shape mismatch or a variable
        that is missing from the file raises unless skip_mismatch.  Returns {"loaded", "skipped", "missing",
        "unexpected"} name lists so that a partial load is never silent."""
        from . import keras_io
        self._require_built()
        p = str(path)
        if p.endswith((".h5", ".hdf5")):
            named = keras_io.load_h5(self, p, by_name=by_name, skip_mismatch=skip_mismatch)
        else:
            p = p if p.endswith(".npz") else p + ".npz"
            with np.load(p) as z:
                named = {k: z[k] for k in z.files}
        report = {"loaded": [], "skipped": [], "missing": [], "unexpected": [k for k in named if k not in
                                                                             {v["name"] for v in self._vars}]}
        for v in self._vars:
            if v["name"] not in named:
                report["missing"].append(v["name"])
                continue
            try:
                self._write_var(v, named[v["name"]])
                report["loaded"].append(v["name"])
            except ValueError:
                if not skip_mismatch:
                    raise
                report["skipped"].append(v["name"])
        if (report["missing"] or report["unexpected"]) and not (skip_mismatch or by_name):
            raise ValueError(f"checkpoint does not match the model: missing {report['missing'][:4]}..., "
                             f"unexpected {report['unexpected'][:4]}... (pass by_name=True / skip_mismatch=True to load partially)")
        self.last_load_report = report
        return report

    def loss_and_gradients(self, data, seed=None):
        """Forward + tape gradient without the optimizer update (tests / diagnostics): returns
        (metrics, {variable name: d(total_loss)/d(var)}) where total_loss = CE + L2 as in lc_NIC.py:381-387.
        Weights, Adam state and BatchNorm moving statistics are left untouched."""
        if self._ctx is not None:
            state = self._state.clone()
        met = self._run_step(data, train=True, training=True, seed=seed, apply=False)
        g, w = self._g.cpu().numpy(), self._w.cpu().numpy()
        out = {}
        for v in self._vars:
            if v["trainable"]:
                n = v["rows"] * max(v["cols"], 1)
                a = g[v["offset"]:v["offset"] + n] + 2.0 * v["l2"] * w[v["offset"]:v["offset"] + n]
                out[v["name"]] = a.reshape(v["rows"], v["cols"]) if v["cols"] > 0 else a
        try:
            self._state.copy_(state)
        except NameError:
            pass
        return met, out

    def count_params(self):
        self._require_built()
        return int(sum(v["rows"] * max(v["cols"], 1) for v in self._vars))

    def summary(self, print_fn=print):
        self._require_built()
        layers = {}
        for v in self._vars:
            key = v["name"].rsplit("/", 1)[0]
            key = key.split("/dense_")[0] + ("/dense_*" if "/dense_" in key else "")
            layers[key] = layers.get(key, 0) + v["rows"] * max(v["cols"], 1)
        print_fn('Model: "nic" (libtnt / sm_100a)')
        for k, n in layers.items():
            print_fn(f"  {k:<44s}{n:>12,d}")
        print_fn(f"Total params: {self.count_params():,d}")

    def launch_count(self):
        return int(self._libh.tnt_launch_count(self._ctx)) if self._ctx is not None else 0

    # ---- Keras fit() (main.py:269-281) --------------------------------------------------------------
    def fit(self, x, epochs=1, steps_per_epoch=None, callbacks=None, validation_data=None, validation_steps=None,
            initial_epoch=0, verbose=0, **_):
        """Iterates a keras.utils.Sequence-like generator (len, [], on_epoch_end), calling train_step /
        test_step per batch and the callback hooks with the logs Keras passes.  The reference's custom train_step /
        test_step return plain tensors, not Keras metric objects, so Keras reports the LAST batch's values as the
        epoch logs and as val_* (what ModelCheckpoint(monitor='val_loss') at main.py:166-190 and LossHistory see);
        `history` keeps the same."""
        cbs = list(callbacks or [])
        for cb in cbs:
            if hasattr(cb, "set_model"):
                cb.set_model(self)
        call = lambda name, *a: [getattr(cb, name)(*a) for cb in cbs if hasattr(cb, name)]
        history = {}
        call("on_train_begin", {})
        for epoch in range(initial_epoch, epochs):
            call("on_epoch_begin", epoch, {})
            n = steps_per_epoch if steps_per_epoch is not None else len(x)
            logs = {}
            nxt = x[0] if n > 0 else None
            if nxt is not None:
                self.prefetch(nxt)
            for i in range(n):
                call("on_train_batch_begin", i, {})
                cur = nxt
                if i + 1 < n:
                    nxt = x[i + 1]
                    self.prefetch(nxt)                 # uploads batch i+1 while batch i computes
                logs = dict(self.train_step(cur))
                call("on_train_batch_end", i, dict(logs))
            if validation_data is not None:
                nv = validation_steps if validation_steps is not None else len(validation_data)
                call("on_test_begin", {})
                vl = {}
                for i in range(nv):
                    call("on_test_batch_begin", i, {})
                    vl = self.test_step(validation_data[i][:2])
                    call("on_test_batch_end", i, dict(vl))
                call("on_test_end", dict(vl))
                logs.update({"val_" + k: v for k, v in vl.items()})
                if hasattr(validation_data, "on_epoch_end"):
                    validation_data.on_epoch_end()
            for k, v in logs.items():
                history.setdefault(k, []).append(v)
            call("on_epoch_end", epoch, dict(logs))
            if hasattr(x, "on_epoch_end"):
                x.on_epoch_end()
            if verbose:
                print(f"epoch {epoch + 1}/{epochs} " + " ".join(f"{k}={v:.4f}" for k, v in logs.items()))
            if self.stop_training:
                break
        call("on_train_end", {})
        self.history = history
        return self


def image_groups(n_locations, feature_width):
    """Group table of the image control (img_localDense.py:31): location r reads x[:, r, :] of a
    (B, n_locations, feature_width) input, i.e. columns [r*K, (r+1)*K) of the flattened row."""
    return [np.arange(r * feature_width, (r + 1) * feature_width, dtype=np.int32) for r in range(n_locations)]
