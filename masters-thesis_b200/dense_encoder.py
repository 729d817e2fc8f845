"""Dense voxel encoder of AttemptFour/Model/NIC.py (SURVEY 8a row A1''), B200 path.

Mirrors the encoder part of `class NIC(tf.keras.Model)` (NIC.py:57-69 layers, :122-128 call):

    img_input = self.dropout_input(img_input); features = self.dense_in(img_input)
    features = self.dropout(features); features = self.expand(features); features = self.batch_norm(features)

`DenseEncoder(input_size, embedding_dim, dropout_input, dropout, input_reg)` takes the corresponding constructor
arguments of NIC.py:22; `__call__(x, training)` returns the `(B, 1, embedding_dim)` features as a CUDA tensor;
`backward(dfeatures)` + `apply_gradients()` are what `tape.gradient` / `optimizer.apply_gradients` (NIC.py train
step) do to its four variables.  Weights in Keras order: dense_img/kernel, dense_img/bias, batch_norm/gamma,
batch_norm/beta, batch_norm/moving_mean, batch_norm/moving_variance.  No CPU path: libtnt.so is required."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .init import _trunc_normal, _TRUNC
from .model import Adam, _opt_field


class DenseEncoder:
    def __init__(self, input_size, embedding_dim, dropout_input=0.0, dropout=0.0, input_reg=0.0, *, batch_max=64,
                 seed=42, device=None):
        self._libh = _lib.load()
        self.K, self.E = int(input_size), int(embedding_dim)
        self.dropout_input, self.dropout, self.input_reg = float(dropout_input), float(dropout), float(input_reg)
        self.batch_max = int(batch_max)
        self.device = torch.device(device if device is not None else "cuda")
        self.optimizer = Adam(1e-3)
        self._seed, self._step = int(seed), 0
        self._ctx = None
        self._build()

    def compile(self, optimizer=None, loss=None, **_):
        if optimizer is not None:
            self.optimizer = optimizer
        self._build(keep=True)

    def _opt(self):
        o = self.optimizer
        return (_opt_field(o, ("learning_rate", "lr"), 1e-3), _opt_field(o, ("beta_1",), 0.9),
                _opt_field(o, ("beta_2",), 0.999), _opt_field(o, ("epsilon",), 1e-7), _opt_field(o, ("clipnorm",), None))

    def _build(self, keep=False):
        lib = self._libh
        lr, b1, b2, eps, clip = self._opt()
        cfg = _lib.TntDenseConfig(K=self.K, E=self.E, B_max=self.batch_max, dropout_input=self.dropout_input,
                                  dropout=self.dropout, input_reg=self.input_reg, lr=lr, beta1=b1, beta2=b2,
                                  adam_eps=eps, clipnorm=float(clip) if clip else 0.0)
        old = (self._w, self._m, self._v, self._state, self._adam_t) if keep and self._ctx is not None else None
        self.close()
        ctx = C.c_void_p()
        with torch.cuda.device(self.device):
            rc = lib.tnt_dense_create(C.byref(cfg), C.byref(ctx))
        if rc != 0:
            raise _lib.TntError(f"tnt_dense_create failed ({rc}): {lib.tnt_dense_last_error(None).decode()}")
        self._ctx = ctx
        n = lib.tnt_dense_param_floats(ctx)
        if old is None:
            self._w = torch.zeros(n, device=self.device)
            self._m, self._v = torch.zeros_like(self._w), torch.zeros_like(self._w)
            self._state = torch.zeros(2 * self.E, device=self.device)
            self._adam_t = 0
            rng = np.random.default_rng(self._seed)
            glorot = _trunc_normal(rng, (self.K, self.E), (2.0 / (self.K + self.E)) ** 0.5 / _TRUNC)   # NIC.py:66
            self.set_weights([glorot, np.zeros(self.E, np.float32),
                              np.ones(self.E, np.float32), np.zeros(self.E, np.float32), np.zeros(self.E, np.float32),
                              np.ones(self.E, np.float32)], _bound=False)
        else:
            self._w, self._m, self._v, self._state, self._adam_t = old
        self._g = torch.zeros_like(self._w)
        self._check(lib.tnt_dense_bind_params(ctx, self._w.data_ptr(), self._m.data_ptr(), self._v.data_ptr(),
                                              self._g.data_ptr(), self._state.data_ptr()))
        lib.tnt_dense_set_adam_step(ctx, self._adam_t)

    def _check(self, rc):
        if rc != 0:
            raise _lib.TntError(f"libtnt dense error {rc}: {self._libh.tnt_dense_last_error(self._ctx).decode()}")

    def close(self):
        if getattr(self, "_ctx", None) is not None:
            self._libh.tnt_dense_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    # ---- NIC.py:122-128 ------------------------------------------------------------------------------
    def __call__(self, img_input, training=False, seed=None):
        x = img_input if torch.is_tensor(img_input) else torch.from_numpy(np.ascontiguousarray(img_input, np.float32))
        x = x.to(self.device, torch.float32).contiguous().view(x.shape[0], -1)
        if x.shape[1] != self.K:
            raise ValueError(f"expected input width {self.K}, got {x.shape[1]}")
        B = int(x.shape[0])
        if seed is None:
            seed = (self._seed * 0x9E3779B97F4A7C15 + self._step) & 0xFFFFFFFFFFFFFFFF
            self._step += 1
        feats = torch.empty(B, self.E, device=self.device)
        self._check(self._libh.tnt_dense_forward(self._ctx, x.data_ptr(), B, int(bool(training)), int(seed),
                                                 feats.data_ptr(), self._stream()))
        self._x_keepalive = x
        return feats.view(B, 1, self.E)

    def forward_noalloc(self, x, out):
        """Inference forward through tnt_dense_forward_noalloc: no allocation anywhere on the call (the library's scratch
        is a static device array, include/tnt.h), so `x` (B,K) and `out` (B,E) must already be fp32, contiguous and on
        this encoder's device.  K <= 65536, E <= 512; calls that overlap on the GPU race.  No reference counterpart."""
        if not (torch.is_tensor(x) and torch.is_tensor(out) and x.is_cuda and out.is_cuda and x.dtype == torch.float32 and
                out.dtype == torch.float32 and x.is_contiguous() and out.is_contiguous()):
            raise ValueError("forward_noalloc takes contiguous fp32 CUDA tensors (it converts and allocates nothing)")
        B = int(x.shape[0])
        if x.numel() != B * self.K or out.numel() != B * self.E:
            raise ValueError(f"expected x ({B}, {self.K}) and out ({B}, {self.E})")
        self._check(self._libh.tnt_dense_forward_noalloc(self._ctx, x.data_ptr(), B, out.data_ptr(), self._stream()))
        return out

    def backward(self, dfeatures):
        """d(loss)/d(features) (B,1,E) or (B,E) of the last training call -> gradients of the four variables."""
        d = dfeatures if torch.is_tensor(dfeatures) else torch.from_numpy(np.ascontiguousarray(dfeatures, np.float32))
        d = d.to(self.device, torch.float32).contiguous().view(-1, self.E)
        self._check(self._libh.tnt_dense_backward(self._ctx, d.data_ptr(), self._stream()))
        self._d_keepalive = d

    def apply_gradients(self):
        lib = self._libh
        lib.tnt_dense_set_lr(self._ctx, self._opt()[0])
        self._check(lib.tnt_dense_apply_gradients(self._ctx, self._stream()))
        self._adam_t += 1

    # ---- weights ---------------------------------------------------------------------------------------
    def _slices(self):
        K, E = self.K, self.E
        o = K * E
        return [("dense_img/kernel", self._w, 0, (K, E)), ("dense_img/bias", self._w, o, (E,)),
                ("batch_norm/gamma", self._w, o + E, (E,)), ("batch_norm/beta", self._w, o + 2 * E, (E,)),
                ("batch_norm/moving_mean", self._state, 0, (E,)), ("batch_norm/moving_variance", self._state, E, (E,))]

    def get_weights(self):
        return [buf[o:o + int(np.prod(shp))].cpu().numpy().reshape(shp) for _, buf, o, shp in self._slices()]

    def set_weights(self, weights, _bound=True):
        sl = self._slices()
        if len(weights) != len(sl):
            raise ValueError(f"expected {len(sl)} arrays, got {len(weights)}")
        for (name, buf, o, shp), a in zip(sl, weights):
            a = np.asarray(a, np.float32)
            if a.shape != tuple(shp):
                raise ValueError(f"shape mismatch for {name}: expected {shp}, got {a.shape}")
            buf[o:o + a.size].copy_(torch.from_numpy(np.ascontiguousarray(a.reshape(-1))))

    def gradients(self):
        return {name: self._g[o:o + int(np.prod(shp))].cpu().numpy().reshape(shp)
                for name, buf, o, shp in self._slices()[:4]}

    def launch_count(self):
        return int(self._libh.tnt_dense_launch_count(self._ctx))
