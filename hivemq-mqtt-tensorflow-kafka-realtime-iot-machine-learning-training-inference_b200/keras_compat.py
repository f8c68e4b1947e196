"""The slice of the ``tf.keras`` surface the reference scripts use, running on the B200 kernels.

Mirrors (same names, argument meaning and error behaviour where it matters):

* model building  — ``Input``, ``Dense(units, activation, activity_regularizer=l1(..))``, ``Model(inputs, outputs)``
  (python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:161-167), ``Sequential().add(LSTM/RepeatVector/
  TimeDistributed(Dense))`` (python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.py:169-176)
* ``compile(optimizer='adam', loss='mean_squared_error', metrics=['accuracy'])`` (cardata-v1.py:169-171)
* ``fit(...).history`` (cardata-v1.py:192-194, nb2.0:871-877), ``predict(x, callbacks=[...])`` (cardata-v1.py:231),
  ``save(path)`` / ``load_model(path)`` (cardata-v1.py:199-202) on Keras-format ``.h5`` files
* new conveniences with the notebook's math: ``reconstruction_error(x)`` (nb2.0:1015) and ``score(x, threshold=5)``
  (nb2.0:1129,1163)

All arithmetic happens in libb2sensor.so; there is no NumPy/CPU compute path here.
"""
from __future__ import annotations

import ctypes as C
import json
import math
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _ffi, h5lite

_ctx_cache: Dict[int, int] = {}


def get_ctx(device: Optional[int] = None) -> int:
    """Process-wide b2_ctx per CUDA device (created lazily; raises without a B200)."""
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    if device not in _ctx_cache:
        l = _ffi.lib()
        h = C.c_void_p()
        _ffi.check(l.b2_ctx_create(device, C.byref(h)), "b2_ctx_create")
        _ctx_cache[device] = h.value
    return _ctx_cache[device]


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def set_option(name: str, value: int, device: Optional[int] = None) -> None:
    """Tuning / debugging switch of the device's context (``b2_ctx_set_option``): "fast_tanh", "score_cfg", "no_bulk",
    "force_generic", "avro_generic", "lstm_tc", "score_tc", "tc_ctas", "decode_r2", "decode_tc", "decode_mma", "zero_skip".  The environment (B2_FAST_TANH, ...) only sets the defaults,
    once, when the context is created."""
    _ffi.check(_ffi.lib().b2_ctx_set_option(get_ctx(device), name.encode(), int(value)), "b2_ctx_set_option")


def get_option(name: str, device: Optional[int] = None) -> int:
    v = C.c_int()
    _ffi.check(_ffi.lib().b2_ctx_get_option(get_ctx(device), name.encode(), C.byref(v)), "b2_ctx_get_option")
    return int(v.value)


_hint_cache: Dict[int, int] = {}


def set_record_bytes_hint(device: int, avg_record_bytes: int) -> None:
    """The ONE place that forwards the average record size of the batches to come to ``b2_ctx_set_record_bytes_hint`` (decode
    calls and the partition driver both come through here, so the cached value can never go stale)."""
    avg = int(avg_record_bytes)
    if _hint_cache.get(device) != avg:
        _ffi.check(_ffi.lib().b2_ctx_set_record_bytes_hint(get_ctx(device), avg), "b2_ctx_set_record_bytes_hint")
        _hint_cache[device] = avg


def normalize_fn(x, kind, lo, rng, device: Optional[int] = None):
    """``normalize_fn`` (cardata-v1.py:40-131) as a stand-alone operator on a row-major ``[N, D]`` float32 matrix:
    ``(v - lo) / (hi - lo) * 2.0 - 1.0`` per scaled feature, 0.0 for the four forced-zero ones, in binary32 with the
    reference's operation order.  Returns a device tensor (torch in) or a NumPy array (NumPy in)."""
    is_np = not isinstance(x, torch.Tensor)
    xd = torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float32))) if is_np else x
    dev = xd.device if xd.is_cuda else torch.device("cuda", torch.cuda.current_device() if device is None else device)
    xd = xd.to(dev, torch.float32).contiguous()
    if xd.dim() != 2 or xd.shape[1] != len(kind):
        raise ValueError(f"expected rows of shape (N, {len(kind)}), got {tuple(xd.shape)}")
    out = torch.empty_like(xd)
    nm = _ffi.make_norm(kind, lo, rng)
    with torch.cuda.device(dev):
        _ffi.check(_ffi.lib().b2_normalize(get_ctx(dev.index), _ffi.ptr(xd), int(xd.shape[0]), int(xd.shape[1]), C.byref(nm),
                                           _ffi.ptr(out), _stream_ptr()), "b2_normalize")
    return out.cpu().numpy() if is_np else out


# ----------------------------------------------------------------------------------------------
# layers / regularisers / callbacks (declarative only)
# ----------------------------------------------------------------------------------------------
class L1:
    def __init__(self, l1: float = 0.01):
        self.l1 = float(np.float32(l1))

    def get_config(self):
        return {"class_name": "L1L2", "config": {"l1": self.l1, "l2": 0.0}}


class regularizers:
    @staticmethod
    def l1(l=0.01):
        return L1(l)


class _Tensor:
    """Symbolic tensor of the functional API: remembers the chain of layers that produced it."""

    def __init__(self, shape, chain):
        self.shape, self.chain = shape, chain


def Input(shape):
    return _Tensor(tuple(shape), [InputLayer(shape)])


class Layer:
    _counts: Dict[str, int] = {}

    def __init__(self, name=None):
        base = type(self).__name__.lower()
        if name is None:
            k = Layer._counts.get(base, 0)
            Layer._counts[base] = k + 1
            name = base if k == 0 else f"{base}_{k}"
        self.name = name

    def __call__(self, t: _Tensor) -> _Tensor:
        return _Tensor(self.output_shape(t.shape), t.chain + [self])

    def output_shape(self, s):
        return s


class InputLayer(Layer):
    def __init__(self, shape, name=None):
        super().__init__(name or None)
        self.shape = tuple(shape)


class Dense(Layer):
    def __init__(self, units, activation=None, activity_regularizer=None, use_bias=True, name=None, **_):
        super().__init__(name)
        if not use_bias:
            raise NotImplementedError("Dense(use_bias=False) is not used by the reference and not supported")
        act = "linear" if activation is None else activation
        if act not in _ffi.ACT_CODES:
            raise ValueError(f"unsupported activation {activation!r} (supported: {sorted(_ffi.ACT_CODES)})")
        self.units, self.activation, self.activity_regularizer = int(units), act, activity_regularizer

    def output_shape(self, s):
        return s[:-1] + (self.units,)


class LSTM(Layer):
    def __init__(self, units, activation="tanh", return_sequences=False, input_shape=None, recurrent_activation="sigmoid",
                 name=None, **_):
        super().__init__(name)
        self.units, self.activation, self.return_sequences, self.input_shape = int(units), activation, return_sequences, input_shape
        self.recurrent_activation = recurrent_activation


class RepeatVector(Layer):
    def __init__(self, n, name=None):
        super().__init__(name)
        self.n = int(n)


class TimeDistributed(Layer):
    def __init__(self, layer, name=None):
        super().__init__(name)
        self.layer = layer


class Callback:
    """Same hooks the reference's OutputCallback overrides (cardata-v1.py:212-226)."""

    def on_predict_batch_end(self, batch, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass


class History:
    def __init__(self):
        self.history: Dict[str, List[float]] = {}



class Windows:
    """``dataset.window(look_back, shift=1, drop_remainder=True).flat_map(lambda w: w.batch(look_back))``
    (LSTM-…/cardata-v1.py:184-185) as a *view*: it keeps the row stream ``rows[N, features]`` and the window geometry;
    ``Model.predict(Windows(...))`` hands both to ``b2_lstm_forward_windows`` so the kernel reads every window in place.
    ``targets()`` is the script's ``dataset.skip(look_back)`` (:189), truncated like ``Dataset.zip``; ``numpy()`` materialises
    the windows (for tests and small inputs)."""

    def __init__(self, rows, look_back: int, shift: int = 1):
        if look_back < 1 or shift < 1:
            raise ValueError("look_back and shift must be >= 1")
        self.rows, self.look_back, self.shift = rows, int(look_back), int(shift)

    def __len__(self):
        n = int(self.rows.shape[0])
        return 0 if n < self.look_back else (n - self.look_back) // self.shift + 1

    def numpy(self):
        r = self.rows.detach().cpu().numpy() if isinstance(self.rows, torch.Tensor) else np.asarray(self.rows)
        idx = np.arange(len(self))[:, None] * self.shift + np.arange(self.look_back)[None, :]
        return r[idx]

    def targets(self):
        r = self.rows.detach().cpu().numpy() if isinstance(self.rows, torch.Tensor) else np.asarray(self.rows)
        y = r[self.look_back:]
        return y[:min(len(y), len(self))][:, None, :]

class ModelCheckpoint(Callback):
    """``ModelCheckpoint(filepath, save_best_only=True)`` on val_loss (nb2.0:862-864)."""

    def __init__(self, filepath, monitor="val_loss", save_best_only=False, verbose=0):
        self.filepath, self.monitor, self.save_best_only, self.best, self.model = filepath, monitor, save_best_only, math.inf, None

    def on_epoch_end(self, epoch, logs=None):
        cur = (logs or {}).get(self.monitor)
        if not self.save_best_only or (cur is not None and cur < self.best):
            if cur is not None:
                self.best = cur
            self.model.save(self.filepath.format(epoch=epoch + 1))


# ----------------------------------------------------------------------------------------------
# initialisers (model_config: GlorotUniform kernels, zero biases; LSTM: orthogonal recurrent, unit_forget_bias)
# ----------------------------------------------------------------------------------------------
def _glorot(g: np.random.Generator, shape):
    lim = math.sqrt(6.0 / (shape[0] + shape[1]))
    return g.uniform(-lim, lim, size=shape).astype(np.float32)


def init_dense_weights(widths: Sequence[int], seed: int = 314) -> np.ndarray:
    g = np.random.Generator(np.random.PCG64(seed))
    parts = []
    for a, b in zip(widths[:-1], widths[1:]):
        parts += [_glorot(g, (a, b)).ravel(), np.zeros(b, dtype=np.float32)]
    return np.concatenate(parts)


def init_lstm_weights(features: int = 18, units=(32, 16, 16, 32), seed: int = 314) -> np.ndarray:
    g = np.random.Generator(np.random.PCG64(seed))
    parts = []
    inp = features
    for u in units:
        parts.append(_glorot(g, (inp, 4 * u)).ravel())
        q, _ = np.linalg.qr(g.standard_normal((4 * u, u)))
        parts.append(q.T.astype(np.float32).ravel())
        b = np.zeros(4 * u, dtype=np.float32)
        b[u:2 * u] = 1.0
        parts.append(b)
        inp = u
    parts += [_glorot(g, (units[-1], features)).ravel(), np.zeros(features, dtype=np.float32)]
    return np.concatenate(parts)


# ----------------------------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------------------------
def _to_device_f32(x, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        t = x
        if t.dtype != torch.float32:
            t = t.to(torch.float32)  # Keras casts float64 inputs to float32 on entry
        if not t.is_cuda:
            t = t.pin_memory().to(device, non_blocking=True)
        return t.contiguous()
    a = np.ascontiguousarray(x, dtype=np.float32)
    return torch.from_numpy(a).pin_memory().to(device, non_blocking=True)


class Model:
    """Dense autoencoder (any Dense stack ≤ 8 layers × 64 units; the reference's D-14-7-7-D stack gets the fused
    FFMA2 kernels) or the reference's LSTM stack."""

    def __init__(self, inputs=None, outputs=None, name="model", _layers=None, seed: int = 314, device: Optional[int] = None):
        self.name = name
        self.layers: List[Layer] = list(_layers) if _layers is not None else list(outputs.chain)
        self._device = device
        self._ae = None
        self._lstm = None
        self._compiled = False
        self._adam = dict(lr=float(np.float32(1e-3)), beta1=float(np.float32(0.9)), beta2=float(np.float32(0.999)),
                          eps=float(np.float32(1e-7)), iter=0)
        self._norm = None
        self._seed = seed
        self._weights: Optional[np.ndarray] = None
        self._h5_layer_names: Optional[List[str]] = None
        self.history = None
        if self.layers:
            self._analyse()

    # -- structure ---------------------------------------------------------------------------------
    def _analyse(self):
        self._built = False
        self._analyse_impl()
        self._built = True

    def _analyse_impl(self):
        ls = [l for l in self.layers if not isinstance(l, InputLayer)]
        self.kind = "lstm" if any(isinstance(l, LSTM) for l in ls) else "dense"
        if self.kind == "dense":
            if not all(isinstance(l, Dense) for l in ls):
                raise NotImplementedError("only stacks of Dense layers (autoencoder) are supported")
            inp = [l for l in self.layers if isinstance(l, InputLayer)]
            if not inp:
                raise ValueError("model has no Input layer")
            self.widths = [int(inp[0].shape[-1])] + [l.units for l in ls]
            self.acts = [_ffi.ACT_CODES[l.activation] for l in ls]
            reg = ls[0].activity_regularizer
            if any(l.activity_regularizer is not None for l in ls[1:]):
                raise NotImplementedError("activity regulariser is supported on the first Dense layer only (as in the reference)")
            self.l1 = float(reg.l1) if reg is not None else 0.0
            self.n_params = sum(a * b + b for a, b in zip(self.widths[:-1], self.widths[1:]))
        else:
            lstm = [l for l in ls if isinstance(l, LSTM)]
            td = [l for l in ls if isinstance(l, TimeDistributed)]
            rv = [l for l in ls if isinstance(l, RepeatVector)]
            ok = (len(lstm) == 4 and len(td) == 1 and len(rv) == 1 and isinstance(td[0].layer, Dense)
                  and all(l.activation == "relu" and l.recurrent_activation == "sigmoid" for l in lstm)
                  and [l.return_sequences for l in lstm] == [True, False, True, True]
                  and td[0].layer.activation == "linear")
            if not ok:
                raise NotImplementedError("only the reference LSTM stack (LSTM seq → LSTM → RepeatVector → LSTM seq → LSTM seq "
                                          "→ TimeDistributed(Dense), relu/sigmoid) is supported")
            self.units = [l.units for l in lstm]
            self.look_back = rv[0].n
            shp = lstm[0].input_shape
            self.features = int(shp[-1]) if shp else td[0].layer.units
            if td[0].layer.units != self.features:
                raise NotImplementedError("TimeDistributed(Dense(n)) must map back to the input feature count")
            inp = self.features
            n = 0
            for u in self.units:
                n += inp * 4 * u + u * 4 * u + 4 * u
                inp = u
            self.n_params = n + self.units[-1] * self.features + self.features

    def _ensure_built(self):
        if not getattr(self, "_built", False):
            self._analyse()  # raises NotImplementedError / ValueError for unsupported or incomplete stacks

    @property
    def input_dim(self):
        return self.widths[0] if self.kind == "dense" else self.features

    def count_params(self):
        return self.n_params

    def summary(self, print_fn=print):
        print_fn(f'Model: "{self.name}"')
        for l in self.layers:
            print_fn(f"  {l.name:<24}{type(l).__name__}")
        print_fn(f"Total params: {self.n_params:,}")

    # -- weights -------------------------------------------------------------------------------------
    def get_flat_weights(self) -> np.ndarray:
        if self._ae is not None:
            out = np.empty(self.n_params, dtype=np.float32)
            _ffi.check(_ffi.lib().b2_ae_get_weights(self._ae, out.ctypes.data), "b2_ae_get_weights")
            self._weights = out
        if self._weights is None:
            self._weights = (init_dense_weights(self.widths, self._seed) if self.kind == "dense"
                             else init_lstm_weights(self.features, self.units, self._seed))
        return self._weights.copy()

    def set_flat_weights(self, w: np.ndarray):
        w = np.ascontiguousarray(w, dtype=np.float32).ravel()
        if w.size != self.n_params:
            raise ValueError(f"expected {self.n_params} weights, got {w.size}")
        self._weights = w.copy()
        if self._ae is not None:
            _ffi.check(_ffi.lib().b2_ae_set_weights(self._ae, self._weights.ctypes.data), "b2_ae_set_weights")
        if self._lstm is not None:
            _ffi.lib().b2_lstm_destroy(self._lstm)
            self._lstm = None

    def get_weights(self) -> List[np.ndarray]:
        flat, out, p = self.get_flat_weights(), [], 0
        for shp in self._weight_shapes():
            n = int(np.prod(shp))
            out.append(flat[p:p + n].reshape(shp))
            p += n
        return out

    def set_weights(self, ws: Sequence[np.ndarray]):
        self.set_flat_weights(np.concatenate([np.asarray(w, dtype=np.float32).ravel() for w in ws]))

    def _weight_shapes(self):
        if self.kind == "dense":
            s = []
            for a, b in zip(self.widths[:-1], self.widths[1:]):
                s += [(a, b), (b,)]
            return s
        s, inp = [], self.features
        for u in self.units:
            s += [(inp, 4 * u), (u, 4 * u), (4 * u,)]
            inp = u
        return s + [(self.units[-1], self.features), (self.features,)]

    def set_normalization(self, kind, lo, rng, order: str = "f32"):
        """Fuse ``normalize_fn`` (cardata-v1.py:40-131) into score/predict/fit: raw rows in, normalised inside the kernel.
        ``order="f64"`` is cardata-v3.py's variant (:53-168): Avro values decoded as float64, scaled in float64, rounded to
        float32 last — it needs float64 values, so only the fused Avro decode + score accepts it."""
        if order not in ("f32", "f64"):
            raise ValueError("order must be 'f32' or 'f64'")
        self._norm = None if kind is None else (list(kind), list(lo), list(rng), 1 if order == "f64" else 0)
        self._norm_struct_key = None  # the cached C struct belongs to the previous table
        return self

    # -- native handles ------------------------------------------------------------------------------
    def _device_index(self):
        return self._device if self._device is not None else torch.cuda.current_device()

    def _handle(self):
        l = _ffi.lib()
        if self.kind == "dense":
            if self._ae is None:
                w = self.get_flat_weights()
                d = _ffi.AeDesc()
                d.n_layers = len(self.acts)
                for i, v in enumerate(self.widths):
                    d.widths[i] = v
                for i, v in enumerate(self.acts):
                    d.act[i] = v
                d.l1_activity = self.l1
                h = C.c_void_p()
                _ffi.check(l.b2_ae_create(get_ctx(self._device_index()), C.byref(d), w.ctypes.data, C.byref(h)), "b2_ae_create")
                self._ae = h.value
            return self._ae
        if self._lstm is None:
            w = self.get_flat_weights()
            d = _ffi.LstmDesc()
            d.features = self.features
            for i, u in enumerate(self.units):
                d.units[i] = u
            h = C.c_void_p()
            _ffi.check(l.b2_lstm_create(get_ctx(self._device_index()), C.byref(d), w.ctypes.data, C.byref(h)), "b2_lstm_create")
            self._lstm = h.value
        return self._lstm

    def __del__(self):
        try:
            l = _ffi.lib()
            if self._ae is not None:
                l.b2_ae_destroy(self._ae)
            if self._lstm is not None:
                l.b2_lstm_destroy(self._lstm)
        except Exception:
            pass

    def _norm_struct(self):
        if getattr(self, "_norm_struct_key", None) is None:  # rebuilt only after set_normalization installed a new table
            self._norm_struct_cached = _ffi.make_norm(*self._norm) if self._norm is not None else _ffi.make_norm()
            self._norm_struct_key = True
        return self._norm_struct_cached

    # -- inference -------------------------------------------------------------------------------------
    def score_device(self, x: torch.Tensor, threshold: float = 5.0, mse: Optional[torch.Tensor] = None,
                     flag: Optional[torch.Tensor] = None, recon: Optional[torch.Tensor] = None,
                     counters: Optional[torch.Tensor] = None):
        """Enqueue the fused kernel on device buffers (no host sync). Returns (mse, flag)."""
        self._ensure_built()
        if self.kind != "dense":
            raise TypeError("score is defined for the Dense autoencoder")
        n = x.shape[0]
        if x.dim() != 2 or x.shape[1] != self.widths[0]:
            raise ValueError(f"expected input of shape (N, {self.widths[0]}), got {tuple(x.shape)}")
        if mse is None:
            mse = torch.empty(n, dtype=torch.float32, device=x.device)
        if flag is None:
            flag = torch.empty(n, dtype=torch.uint8, device=x.device)
        nm = self._norm_struct()
        _ffi.check(_ffi.lib().b2_ae_score(self._handle(), _ffi.ptr(x), n, C.byref(nm), float(threshold), _ffi.ptr(mse),
                                          _ffi.ptr(flag), _ffi.ptr(recon), _ffi.ptr(counters), _stream_ptr()), "b2_ae_score")
        return mse, flag

    def _score_host_overlapped(self, x: torch.Tensor, threshold: float, dev, chunks: int = 4):
        """Pinned float32 rows on the host → (mse, flag) in a pinned result buffer kept per size: the rows cross PCIe in chunks of
        falling size on a copy stream, chunk c is scored while chunk c + 1 arrives, and its 5 B/row results go back on a third
        stream (the same three-stream overlap as decode.decode_score_host)."""
        n, d = x.shape
        ws = getattr(self, "_host_ws", None)
        if ws is None or ws["dev"] != dev:
            ws = self._host_ws = {"dev": dev, "side": torch.cuda.Stream(device=dev), "back": torch.cuda.Stream(device=dev), "cap": 0,
                                  "ev": [torch.cuda.Event() for _ in range(16)], "done": [torch.cuda.Event() for _ in range(16)], "out": {}}
        if n > ws["cap"]:
            ws["cap"] = n
            ws["x"] = torch.empty((n, d), dtype=torch.float32, device=dev)
            ws["res"] = torch.empty(n * 5 + 16, dtype=torch.uint8, device=dev)
        out = ws["out"].get(n)
        if out is None:
            out = ws["out"][n] = torch.empty(n * 5, dtype=torch.uint8).pin_memory()
        xd, res = ws["x"][:n], ws["res"]
        mse, flag = res[: n * 4].view(torch.float32), res[n * 4: n * 5]
        main, side, back = torch.cuda.current_stream(), ws["side"], ws["back"]
        chunks = max(1, min(chunks, 16))
        tot = chunks * (chunks + 1) // 2
        bounds = sorted(set(n * (tot - (chunks - c) * (chunks - c + 1) // 2) // tot for c in range(chunks + 1)))
        side.wait_stream(main)
        with torch.cuda.stream(side):
            for k, (c0, c1) in enumerate(zip(bounds[:-1], bounds[1:])):
                xd[c0:c1].copy_(x[c0:c1], non_blocking=True)
                ws["ev"][k].record(side)
        for k, (c0, c1) in enumerate(zip(bounds[:-1], bounds[1:])):
            main.wait_event(ws["ev"][k])
            self.score_device(xd[c0:c1], threshold, mse=mse[c0:c1], flag=flag[c0:c1])
            ws["done"][k].record(main)
            with torch.cuda.stream(back):
                back.wait_event(ws["done"][k])
                out[c0 * 4: c1 * 4].copy_(res[c0 * 4: c1 * 4], non_blocking=True)
                out[n * 4 + c0: n * 4 + c1].copy_(res[n * 4 + c0: n * 4 + c1], non_blocking=True)
        main.wait_stream(back)
        main.synchronize()
        return out[: n * 4].view(torch.float32).numpy(), out[n * 4:].numpy()

    def score(self, x, threshold: float = 5.0, reuse_output: bool = False):
        """``mse = mean((x - predict(x))**2, axis=1)`` (nb2.0:1015) and ``flag = mse > threshold`` (nb2.0:1163).
        Returns (mse f32[N], flag u8[N]) as NumPy arrays.  With ``reuse_output`` and a pinned float32 CPU tensor the copies and the
        kernel overlap chunk by chunk and the returned arrays are views of a pinned buffer that the next call of the same size
        overwrites."""
        dev = torch.device("cuda", self._device_index())
        if (reuse_output and isinstance(x, torch.Tensor) and not x.is_cuda and x.dtype == torch.float32 and x.dim() == 2
                and x.shape[0] >= 4096 and x.is_contiguous() and x.is_pinned() and self.kind == "dense"):
            self._ensure_built()
            with torch.cuda.device(dev):
                return self._score_host_overlapped(x, threshold, dev)
        xd = _to_device_f32(x, dev)
        mse, flag = self.score_device(xd, threshold)
        return mse.cpu().numpy(), flag.cpu().numpy()

    def reconstruction_error(self, x):
        return self.score(x, threshold=math.inf)[0]

    def predict(self, x, batch_size: Optional[int] = None, callbacks: Optional[Sequence[Callback]] = None, verbose=0):
        """Forward pass over all rows; one kernel launch regardless of ``batch_size``.  ``callbacks`` get
        ``on_predict_batch_end(batch, {'outputs': [rows_of_that_batch]})`` per ``batch_size`` chunk (default 32) exactly as
        TF 2.0.0's predict loop feeds the reference's OutputCallback (cardata-v1.py:218-224)."""
        self._ensure_built()
        dev = torch.device("cuda", self._device_index())
        if isinstance(x, Windows):
            if self.kind == "dense":
                raise ValueError("Windows input is for the LSTM model")
            if x.look_back != self.look_back:
                raise ValueError(f"model was built with look_back={self.look_back}, got windows of {x.look_back}")
            rows = _to_device_f32(x.rows, dev)
            if rows.dim() != 2 or rows.shape[1] != self.features:
                raise ValueError(f"expected a row stream of shape (N, {self.features}), got {tuple(rows.shape)}")
            y = torch.empty((len(x), x.look_back, self.features), dtype=torch.float32, device=dev)
            if len(x) == 0:
                return y.cpu().numpy()
            _ffi.check(_ffi.lib().b2_lstm_forward_windows(self._handle(), _ffi.ptr(rows), rows.shape[0], x.look_back, x.shift,
                                                          _ffi.ptr(y), _stream_ptr()), "b2_lstm_forward_windows")
            out = y.cpu().numpy()
            return self._predict_callbacks(out, batch_size, callbacks)
        xd = _to_device_f32(x, dev)
        if self.kind == "dense":
            if xd.dim() != 2 or xd.shape[1] != self.widths[0]:
                raise ValueError(f"expected input of shape (N, {self.widths[0]}), got {tuple(xd.shape)}")
            recon = torch.empty_like(xd)
            self.score_device(xd, math.inf, recon=recon)
            out = recon.cpu().numpy()
            if callbacks and any(hasattr(cb, "consume_rows") for cb in callbacks):
                # result sinks that format on the GPU (sink.OutputCallback) take every row in one launch, from device memory
                for cb in callbacks:
                    if hasattr(cb, "consume_rows"):
                        cb.consume_rows(recon)
                callbacks = [cb for cb in callbacks if not hasattr(cb, "consume_rows")]
        else:
            if xd.dim() != 3 or xd.shape[2] != self.features:
                raise ValueError(f"expected input of shape (B, T, {self.features}), got {tuple(xd.shape)}")
            if xd.shape[1] != self.look_back:
                raise ValueError(f"model was built with look_back={self.look_back}, got T={xd.shape[1]}")
            y = torch.empty_like(xd)
            _ffi.check(_ffi.lib().b2_lstm_forward(self._handle(), _ffi.ptr(xd), xd.shape[0], xd.shape[1], _ffi.ptr(y),
                                                  _stream_ptr()), "b2_lstm_forward")
            out = y.cpu().numpy()
        return self._predict_callbacks(out, batch_size, callbacks)

    @staticmethod
    def _predict_callbacks(out, batch_size, callbacks):
        if callbacks:
            bs = batch_size or 32
            for b, s in enumerate(range(0, out.shape[0], bs)):
                for cb in callbacks:
                    cb.on_predict_batch_end(b, {"outputs": [out[s:s + bs]]})
        return out

    # -- training ----------------------------------------------------------------------------------------
    def compile(self, optimizer="adam", loss="mean_squared_error", metrics=None, **_):
        if (optimizer if isinstance(optimizer, str) else getattr(optimizer, "name", "")).lower() != "adam":
            raise NotImplementedError("only optimizer='adam' is supported (the reference's training_config)")
        if loss not in ("mean_squared_error", "mse"):
            raise NotImplementedError("only loss='mean_squared_error' is supported")
        self._metrics = list(metrics or [])
        self._compiled = True

    def fit(self, x, y=None, epochs=1, batch_size=32, shuffle=True, validation_data=None, verbose=0, callbacks=None,
            seed: Optional[int] = None, data_parallel: bool = False):
        """Mini-batch Adam on MSE(x, AE(x)) + l1·Σ|h1|/B, all steps of an epoch inside ONE kernel launch.
        ``y`` must be None or ``x`` (autoencoder; the notebooks call ``fit(train_x, train_x)``, nb2.0:871).

        ``batch_size`` ≤ 256 (the reference uses 32/100/128) runs the single-CTA persistent kernel.  Larger batches, or
        ``data_parallel=True`` under an initialised ``torch.distributed`` NCCL group (``x`` = this rank's shard,
        ``batch_size`` = per-rank batch), run the grid-wide gradient kernel, ONE all-reduce of the P+3 floats and the Adam
        kernel per step."""
        self._ensure_built()
        if not self._compiled:
            raise RuntimeError("You must compile your model before training/testing. Use `model.compile(optimizer, loss)`.")
        if self.kind != "dense":
            return self._fit_lstm(x, y, epochs, batch_size, shuffle, verbose, callbacks, seed)
        if y is not None and y is not x:
            if np.shape(y) != np.shape(x) or not np.array_equal(np.asarray(y), np.asarray(x)):
                raise NotImplementedError("autoencoder training only: y must be x")
        dev = torch.device("cuda", self._device_index())
        xd = _to_device_f32(x, dev)
        n = xd.shape[0]
        l = _ffi.lib()
        h = self._handle()
        self._restore_adam_slots()
        if batch_size > 256 or data_parallel:
            return self._fit_large(xd, epochs, batch_size, shuffle, validation_data, verbose, callbacks, seed, data_parallel)
        steps = (n + batch_size - 1) // batch_size
        hist = History()
        hist.history = {"loss": [], "accuracy": []} if "accuracy" in getattr(self, "_metrics", []) else {"loss": []}
        g = np.random.Generator(np.random.PCG64(self._seed if seed is None else seed))
        adam = _ffi.Adam(self._adam["lr"], self._adam["beta1"], self._adam["beta2"], self._adam["eps"], self._adam["iter"])
        nm = self._norm_struct()
        loss_d = torch.empty(steps, dtype=torch.float32, device=dev)
        acc_d = torch.empty(steps, dtype=torch.float32, device=dev)
        sizes = np.full(steps, batch_size, dtype=np.float64)
        sizes[-1] = n - batch_size * (steps - 1)
        for cb in callbacks or []:
            if isinstance(cb, ModelCheckpoint):
                cb.model = self
        for ep in range(epochs):
            order = None
            if shuffle:
                order = torch.from_numpy(g.permutation(n).astype(np.int32)).to(dev)
            _ffi.check(l.b2_ae_fit_steps(h, _ffi.ptr(xd), _ffi.ptr(order), n, batch_size, C.byref(nm), C.byref(adam),
                                         _ffi.ptr(loss_d), _ffi.ptr(acc_d), _stream_ptr()), "b2_ae_fit_steps")
            # Keras reports the sample-weighted mean of the per-batch values
            lo = loss_d.cpu().numpy().astype(np.float64)
            ac = acc_d.cpu().numpy().astype(np.float64)
            logs = {"loss": float((lo * sizes).sum() / n)}
            if "accuracy" in hist.history:
                logs["accuracy"] = float((ac * sizes).sum() / n)
            if validation_data is not None:
                vx = validation_data[0] if isinstance(validation_data, (tuple, list)) else validation_data
                vm = self.reconstruction_error(vx)
                logs["val_loss"] = float(np.mean(vm.astype(np.float64)))
            for k, v in logs.items():
                hist.history.setdefault(k, []).append(v)
            if verbose:
                print(f"Epoch {ep + 1}/{epochs} - " + " - ".join(f"{k}: {v:.4f}" for k, v in logs.items()))
            for cb in callbacks or []:
                cb.on_epoch_end(ep, logs)
        self._adam["iter"] = int(adam.iter)
        self.history = hist
        return hist

    def _fit_lstm(self, x, y, epochs, batch_size, shuffle, verbose, callbacks, seed):
        """``model.fit(zip((dataset_x, dataset_y)).batch(batch_size), epochs=…)`` of the LSTM script (LSTM-…/cardata-v1.py:191-194):
        back-propagation through time + Adam, every step of an epoch inside one launch (``b2_lstm_fit_steps``).

        ``x``: ``[N, T, F]`` array / tensor with targets ``y`` of the same shape (``[N, F]`` is accepted for ``T == 1``), or a
        :class:`Windows` view with ``look_back == 1`` and ``y=None`` — then the targets are the script's ``dataset.skip(look_back)``
        and both are read in place from the row stream.  Batches are consecutive (a ``tf.data`` pipeline is not shuffled by
        ``fit``); ``shuffle=True`` permutes the sequences per epoch."""
        dev = torch.device("cuda", self._device_index())
        F, T = self.features, None
        if isinstance(x, Windows):
            if y is not None:
                raise ValueError("Windows input brings its own targets (dataset.skip(look_back)); pass y=None")
            if x.look_back != 1:
                raise NotImplementedError("training on sliding windows is implemented for look_back = 1 (the reference's setting); "
                                          "pass materialised [N, T, F] inputs and targets for longer windows")
            rows = _to_device_f32(x.rows, dev)
            n = min(len(x), max(int(rows.shape[0]) - 1, 0))  # Dataset.zip stops at the shorter side (Windows.targets())
            T = 1
            xd, yd = rows, rows[1:]
            xs, ys = x.shift, 1  # window i starts at row i * shift; target i is row 1 + i (dataset.skip(look_back))
        else:
            xd = _to_device_f32(x, dev)
            if xd.dim() != 3 or xd.shape[2] != F:
                raise ValueError(f"expected input of shape (N, T, {F}), got {tuple(xd.shape)}")
            n, T = int(xd.shape[0]), int(xd.shape[1])
            if y is None:
                raise ValueError("LSTM training needs targets y")
            yd = _to_device_f32(y, dev)
            if yd.dim() == 2 and T == 1:
                yd = yd[:, None, :]
            if tuple(yd.shape) != tuple(xd.shape):
                raise ValueError(f"targets must have the model's output shape {tuple(xd.shape)}, got {tuple(yd.shape)}")
            xs = ys = T
        if n < 1:
            raise ValueError("no training sequences")
        l = _ffi.lib()
        d = _ffi.LstmDesc()
        d.features = F
        for i, u in enumerate(self.units):
            d.units[i] = u
        P = self.n_params
        w = torch.from_numpy(self.get_flat_weights().copy()).to(dev)
        st = getattr(self, "_lstm_adam", None)
        if st is None or st[0].numel() != P or st[0].device != dev:
            st = self._lstm_adam = (torch.zeros(P, dtype=torch.float32, device=dev), torch.zeros(P, dtype=torch.float32, device=dev))
        m, v = st
        batch_size = int(batch_size)
        steps = (n + batch_size - 1) // batch_size
        sb = int(l.b2_lstm_fit_scratch_bytes(C.byref(d), T, min(batch_size, n)))
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
        loss_d = torch.empty(steps, dtype=torch.float32, device=dev)
        acc_d = torch.empty(steps, dtype=torch.float32, device=dev)
        adam = _ffi.Adam(self._adam["lr"], self._adam["beta1"], self._adam["beta2"], self._adam["eps"], self._adam["iter"])
        hist = History()
        hist.history = {"loss": [], "accuracy": []} if "accuracy" in getattr(self, "_metrics", []) else {"loss": []}
        g = np.random.Generator(np.random.PCG64(self._seed if seed is None else seed))
        sizes = np.full(steps, batch_size, dtype=np.float64)
        sizes[-1] = n - batch_size * (steps - 1)
        for ep in range(epochs):
            xe, ye, xse, yse = xd, yd, xs, ys
            if shuffle:
                perm = torch.from_numpy(g.permutation(n)).to(dev)
                if isinstance(x, Windows):
                    xe, ye, xse, yse = xd[perm * xs].contiguous(), yd[perm * ys].contiguous(), 1, 1
                else:
                    xe, ye = xd[perm].contiguous(), yd[perm].contiguous()
            _ffi.check(l.b2_lstm_fit_steps(get_ctx(dev.index), C.byref(d), _ffi.ptr(w), _ffi.ptr(m), _ffi.ptr(v), _ffi.ptr(xe), xse,
                                           _ffi.ptr(ye), yse, n, T, min(batch_size, n), C.byref(adam), _ffi.ptr(loss_d), _ffi.ptr(acc_d),
                                           _ffi.ptr(scratch), sb, _stream_ptr()), "b2_lstm_fit_steps")
            lo = loss_d.cpu().numpy().astype(np.float64)
            logs = {"loss": float((lo * sizes).sum() / n)}
            if "accuracy" in hist.history:
                logs["accuracy"] = float((acc_d.cpu().numpy().astype(np.float64) * sizes).sum() / n)
            for k, val in logs.items():
                hist.history.setdefault(k, []).append(val)
            if verbose:
                print(f"Epoch {ep + 1}/{epochs} - " + " - ".join(f"{k}: {val:.4f}" for k, val in logs.items()))
            for cb in callbacks or []:
                cb.on_epoch_end(ep, logs)
        self._adam["iter"] = int(adam.iter)
        self.set_flat_weights(w.cpu().numpy())  # the forward kernels repack the weights on their next use
        self.history = hist
        return hist

    def _restore_adam_slots(self):
        """Push Adam m/v loaded from an .h5 (load_model) into the native handle once, so that training resumes."""
        opt = getattr(self, "_h5_optimizer", None)
        if not opt or self._ae is None:
            return
        dense = [l for l in self.layers if isinstance(l, Dense)]
        names = getattr(self, "_h5_inner_names", None) or [l.name for l in dense]
        m_parts, v_parts = [], []
        try:
            for nm in names:
                for kind in ("kernel", "bias"):
                    m_parts.append([val for k, val in opt.items() if k.endswith(f"Adam/{nm}/{kind}/m:0")][0].ravel())
                    v_parts.append([val for k, val in opt.items() if k.endswith(f"Adam/{nm}/{kind}/v:0")][0].ravel())
        except IndexError:
            self._h5_optimizer = None
            return
        m = np.ascontiguousarray(np.concatenate(m_parts), dtype=np.float32)
        v = np.ascontiguousarray(np.concatenate(v_parts), dtype=np.float32)
        if m.size == self.n_params:
            _ffi.check(_ffi.lib().b2_ae_set_adam(self._ae, m.ctypes.data, v.ctypes.data), "b2_ae_set_adam")
        self._h5_optimizer = None

    def _exchange_buffers(self, dev):
        """Symmetric-memory exchange buffers for the fused data-parallel step (one per model and process group, kept): every
        rank's buffer mapped into every process, so the training kernel stores its gradient vector straight into its peers'
        memory over NVLink.  Returns None when symmetric memory is unavailable on ANY rank (then NCCL all-reduce is used)."""
        import torch.distributed as dist

        ex = getattr(self, "_exch", None)
        if ex is not None:
            return ex or None
        world, ok = dist.get_world_size(), 1
        try:
            import torch.distributed._symmetric_memory as symm

            if os.environ.get("B2_DP_NCCL") == "1" or world > 8:
                raise RuntimeError("fused exchange disabled")
            nfl = int(_ffi.lib().b2_ae_exchange_bytes()) // 4
            buf = symm.empty(nfl, dtype=torch.float32, device=dev)
            buf.zero_()
            hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
            ptrs = (C.c_uint64 * world)(*[int(p) for p in hdl.buffer_ptrs])
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001 — any failure means "no symmetric memory here"
            ok, buf, hdl, ptrs = 0, None, None, None
            self._exch_error = repr(e)
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            self._exch = False
            return None
        dist.barrier()
        self._exch = {"buf": buf, "hdl": hdl, "ptrs": ptrs, "seq": 0, "rank": dist.get_rank(), "world": world}
        return self._exch

    def _fit_large(self, xd, epochs, batch_size, shuffle, validation_data, verbose, callbacks, seed, data_parallel):
        """Large-batch / data-parallel training: ONE kernel per step (``b2_ae_step_fused``) — per-CTA gradient sums, fold, the
        cross-GPU sum over symmetric memory (peer stores + flags inside the kernel) and Adam.  Falls back to gradient kernel →
        NCCL all-reduce → Adam kernel when symmetric memory cannot be set up (``self.dp_path`` says which ran)."""
        import torch.distributed as dist

        dp = data_parallel and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        l, h, dev, n = _ffi.lib(), self._handle(), xd.device, xd.shape[0]
        P = self.n_params
        nm = self._norm_struct()
        adam = _ffi.Adam(self._adam["lr"], self._adam["beta1"], self._adam["beta2"], self._adam["eps"], self._adam["iter"])
        ex = self._exchange_buffers(dev) if dp else None
        fused = (not dp) or ex is not None
        self.dp_path = ("fused step kernel, peer stores over symmetric memory" if ex is not None else
                        "fused step kernel (single GPU)" if fused else "b2_ae_grad + NCCL all-reduce + b2_ae_apply")
        world = dist.get_world_size() if dp else 1
        if fused:
            scratch_bytes = int(l.b2_ae_step_scratch_bytes(h, min(batch_size, n)))
            scratch = torch.zeros(scratch_bytes, dtype=torch.uint8, device=dev)
        else:
            scratch_bytes = int(l.b2_ae_grad_scratch_bytes(h, min(batch_size, n)))
            scratch = torch.empty(scratch_bytes, dtype=torch.uint8, device=dev)
            grads = torch.empty(P + 3, dtype=torch.float32, device=dev)
        steps = (n + batch_size - 1) // batch_size
        stats = torch.empty((steps, 2), dtype=torch.float32, device=dev)
        hist = History()
        hist.history = {"loss": [], "accuracy": []} if "accuracy" in getattr(self, "_metrics", []) else {"loss": []}
        g = np.random.Generator(np.random.PCG64(self._seed if seed is None else seed))
        for cb in callbacks or []:
            if isinstance(cb, ModelCheckpoint):
                cb.model = self
        for ep in range(epochs):
            xs = xd[torch.from_numpy(g.permutation(n)).to(dev)].contiguous() if shuffle else xd
            for s in range(steps):
                xb = xs[s * batch_size:(s + 1) * batch_size]
                if fused:
                    if ex is not None:
                        ex["seq"] += 1
                    _ffi.check(l.b2_ae_step_fused(h, _ffi.ptr(xb), xb.shape[0], C.byref(nm), C.byref(adam), _ffi.ptr(scratch), scratch_bytes,
                                                  int(xb.shape[0]) * world, None if ex is None else ex["ptrs"], world,
                                                  0 if ex is None else ex["rank"], 0 if ex is None else ex["seq"], _ffi.ptr(stats[s]),
                                                  _stream_ptr()), "b2_ae_step_fused")
                    continue
                _ffi.check(l.b2_ae_grad(h, _ffi.ptr(xb), xb.shape[0], C.byref(nm), _ffi.ptr(grads), _ffi.ptr(scratch), scratch_bytes,
                                        _stream_ptr()), "b2_ae_grad")
                if dp:
                    dist.all_reduce(grads)  # NCCL over NVLink: P+3 floats (gradient sums, loss sum, correct, rows)
                _ffi.check(l.b2_ae_apply(h, _ffi.ptr(grads), C.byref(adam), _ffi.ptr(stats[s]), _stream_ptr()), "b2_ae_apply")
            st = stats.cpu().numpy().astype(np.float64)
            if not np.isfinite(st).all():
                raise RuntimeError("data-parallel step: a peer GPU did not arrive at the exchange (step skipped); see b2_ae_step_fused")
            sizes = np.full(steps, batch_size, dtype=np.float64)
            sizes[-1] = n - batch_size * (steps - 1)
            logs = {"loss": float((st[:, 0] * sizes).sum() / n)}
            if "accuracy" in hist.history:
                logs["accuracy"] = float((st[:, 1] * sizes).sum() / n)
            if validation_data is not None:
                vx = validation_data[0] if isinstance(validation_data, (tuple, list)) else validation_data
                logs["val_loss"] = float(np.mean(self.reconstruction_error(vx).astype(np.float64)))
            for k, v in logs.items():
                hist.history.setdefault(k, []).append(v)
            if verbose:
                print(f"Epoch {ep + 1}/{epochs} - " + " - ".join(f"{k}: {v:.4f}" for k, v in logs.items()))
            for cb in callbacks or []:
                cb.on_epoch_end(ep, logs)
        self._adam["iter"] = int(adam.iter)
        self.history = hist
        return hist

    # -- persistence -----------------------------------------------------------------------------------------
    def _model_config(self) -> dict:
        layers = []
        for l in self.layers:
            if isinstance(l, InputLayer):
                layers.append({"class_name": "InputLayer", "name": l.name, "config": {
                    "batch_input_shape": [None, *l.shape], "dtype": "float32", "sparse": False, "name": l.name}, "inbound_nodes": []})
            elif isinstance(l, Dense):
                layers.append({"class_name": "Dense", "name": l.name, "config": {
                    "name": l.name, "trainable": True, "dtype": "float32", "units": l.units, "activation": l.activation,
                    "use_bias": True,
                    "kernel_initializer": {"class_name": "GlorotUniform", "config": {"seed": None}},
                    "bias_initializer": {"class_name": "Zeros", "config": {}},
                    "kernel_regularizer": None, "bias_regularizer": None,
                    "activity_regularizer": l.activity_regularizer.get_config() if l.activity_regularizer else None,
                    "kernel_constraint": None, "bias_constraint": None}})
            else:
                raise NotImplementedError("save() supports Dense models")
        for prev, cur in zip(layers[:-1], layers[1:]):
            cur["inbound_nodes"] = [[[prev["name"], 0, 0, {}]]]
        return {"class_name": "Model", "config": {"name": self.name, "layers": layers,
                                                  "input_layers": [[layers[0]["name"], 0, 0]],
                                                  "output_layers": [[layers[-1]["name"], 0, 0]]}}

    def save(self, path: str):
        """Write a Keras-2.2.4-tf style ``.h5`` (model_config, training_config, model_weights, optimizer_weights)."""
        if self.kind != "dense":
            raise NotImplementedError("save() supports Dense models")
        ws = self.get_weights()
        dense = [l for l in self.layers if isinstance(l, Dense)]
        names = [l.name for l in self.layers]
        mw: Dict[str, object] = {"@attrs": {"layer_names": np.array(names), "backend": "tensorflow", "keras_version": "2.2.4-tf"}}
        for l in self.layers:
            if isinstance(l, InputLayer):
                mw[l.name] = {"@attrs": {"weight_names": np.array([], dtype="S1")}}
        for i, l in enumerate(dense):
            mw[l.name] = {l.name: {"kernel:0": ws[2 * i], "bias:0": ws[2 * i + 1]},
                          "@attrs": {"weight_names": np.array([f"{l.name}/kernel:0", f"{l.name}/bias:0"])}}
        tree: Dict[str, object] = {"model_weights": mw}
        if self._ae is not None and self._adam["iter"] > 0:
            m = np.empty(self.n_params, dtype=np.float32)
            v = np.empty(self.n_params, dtype=np.float32)
            _ffi.check(_ffi.lib().b2_ae_get_adam(self._ae, m.ctypes.data, v.ctypes.data), "b2_ae_get_adam")
            adam: Dict[str, object] = {"iter:0": np.array(self._adam["iter"], dtype=np.int64)}
            p = 0
            # Keras' load_model_from_hdf5 reads optimizer_weights.attrs['weight_names'] unconditionally; its order is the
            # optimizer's variable order: iter, every slot m (kernel, bias per layer), then every slot v
            # (models/autoencoder_sensor_anomaly_detection.h5 carries these 17 names for its 4 Dense layers).
            wn = ["training/Adam/iter:0"]
            for slot in ("m", "v"):
                for l in dense:
                    wn += [f"training/Adam/{l.name}/kernel/{slot}:0", f"training/Adam/{l.name}/bias/{slot}:0"]
            for i, l in enumerate(dense):
                kshape, bshape = ws[2 * i].shape, ws[2 * i + 1].shape
                kn, bn = int(np.prod(kshape)), int(np.prod(bshape))
                adam[l.name] = {"kernel": {"m:0": m[p:p + kn].reshape(kshape), "v:0": v[p:p + kn].reshape(kshape)},
                                "bias": {"m:0": m[p + kn:p + kn + bn], "v:0": v[p + kn:p + kn + bn]}}
                p += kn + bn
            tree["optimizer_weights"] = {"training": {"Adam": adam}, "@attrs": {"weight_names": np.array(wn)}}
        training_config = {"optimizer_config": {"class_name": "Adam", "config": {
            "name": "Adam", "learning_rate": self._adam["lr"], "decay": 0.0, "beta_1": self._adam["beta1"],
            "beta_2": self._adam["beta2"], "epsilon": self._adam["eps"], "amsgrad": False}},
            "loss": "mean_squared_error", "metrics": ["accuracy"], "weighted_metrics": None, "sample_weight_mode": None,
            "loss_weights": None}
        h5lite.write_h5(path, tree, {"keras_version": "2.2.4-tf", "backend": "tensorflow",
                                     "model_config": json.dumps(self._model_config()),
                                     "training_config": json.dumps(training_config)})


class Sequential(Model):
    def __init__(self, layers=None, name="sequential", **kw):
        super().__init__(name=name, _layers=[], **kw)
        for l in layers or []:
            self.add(l)

    def add(self, layer: Layer):
        self.layers.append(layer)
        try:
            self._analyse()
        except (NotImplementedError, ValueError):
            pass  # stack not complete yet; validated again on first use

    def _handle(self):
        self._analyse()
        return super()._handle()


def load_model(path: str, device: Optional[int] = None) -> Model:
    """``tf.keras.models.load_model`` for Keras ``.h5`` files (cardata-v1.py:202, nb2.0:888): rebuilds the Dense stack from
    ``model_config`` and loads ``model_weights`` (+ Adam ``iter``/``m``/``v`` when present so training can resume)."""
    f = h5lite.H5File(path)
    if "model_config" not in f.attrs:
        raise ValueError(f"{path}: no model_config attribute (not a Keras full-model file)")
    cfg = json.loads(f.attrs["model_config"])
    lcfgs = cfg["config"]["layers"] if isinstance(cfg["config"], dict) else cfg["config"]
    layers: List[Layer] = []
    for lc in lcfgs:
        cn, c = lc["class_name"], lc["config"]
        if cn == "InputLayer":
            layers.append(InputLayer(tuple(c["batch_input_shape"][1:]), name=c["name"]))
        elif cn == "Dense":
            reg = c.get("activity_regularizer")
            r = L1(reg["config"]["l1"]) if reg and reg["config"].get("l1") else None
            if reg and reg["config"].get("l2"):
                raise NotImplementedError("l2 activity regulariser not supported")
            if "batch_input_shape" in c and not layers:
                layers.append(InputLayer(tuple(c["batch_input_shape"][1:]), name=c["name"] + "_input"))
            layers.append(Dense(c["units"], activation=c["activation"], activity_regularizer=r, use_bias=c.get("use_bias", True),
                                name=c["name"]))
        else:
            raise NotImplementedError(f"{path}: layer type {cn} is not supported by load_model")
    m = Model(_layers=layers, name=cfg["config"].get("name", "model") if isinstance(cfg["config"], dict) else "sequential",
              device=device)
    mw = f["model_weights"]
    parts = []
    inner_names = []
    for ln in [str(s) for s in np.atleast_1d(mw.attrs["layer_names"])]:
        g = mw[ln]
        wn = [str(s) for s in np.atleast_1d(g.attrs.get("weight_names", []))]
        if not wn:
            continue
        inner_names.append(wn[0].split("/")[0])
        by_kind = {("kernel" if "kernel" in w else "bias"): g[w].read() for w in wn}
        parts += [by_kind["kernel"].astype(np.float32).ravel(), by_kind["bias"].astype(np.float32).ravel()]
    m.set_flat_weights(np.concatenate(parts))
    if "training_config" in f.attrs:
        tc = json.loads(f.attrs["training_config"])
        oc = tc.get("optimizer_config", {})
        if oc.get("class_name") == "Adam":
            c = oc["config"]
            m._adam.update(lr=float(c["learning_rate"]), beta1=float(c["beta_1"]), beta2=float(c["beta_2"]), eps=float(c["epsilon"]))
            m.compile("adam", tc.get("loss", "mean_squared_error"), tc.get("metrics"))
    m._h5_optimizer = None
    m._h5_inner_names = inner_names
    if "optimizer_weights" in f:
        ow = f["optimizer_weights"]
        found: Dict[str, np.ndarray] = {}
        ow.visit(lambda p, o: found.__setitem__(p, o.read()) if not o.is_group else None)
        m._h5_optimizer = found
        for p, v in found.items():
            if p.endswith("iter:0"):
                m._adam["iter"] = int(np.asarray(v).ravel()[0])
    return m


class _NS:
    pass


# tf.keras-like namespaces so that scripts can do `keras.layers.Dense`, `keras.models.load_model`, ...
layers = _NS()
layers.Input, layers.Dense, layers.LSTM, layers.RepeatVector, layers.TimeDistributed = Input, Dense, LSTM, RepeatVector, TimeDistributed
models = _NS()
models.Model, models.Sequential, models.load_model = Model, Sequential, load_model
callbacks = _NS()
callbacks.Callback, callbacks.ModelCheckpoint = Callback, ModelCheckpoint

# module-level helpers reachable as bs.keras.<name> (bs.keras is this module): set_option / get_option / normalize_fn
