"""CPU ORACLE — test infrastructure only, never the product path.

A NumPy restatement of the reference's algorithm for the scoring / training / LSTM hot path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import this.

PARITY UNPINNED: the reference ships no tests, golden vectors or expected outputs for this path
and TensorFlow / tensorflow-io cannot be installed here, so this restatement could not be checked against the
reference's own Keras run.  It is cross-checked instead against independent implementations available offline:
glibc ``strtof`` and ``pyarrow.csv`` for the decimal→binary32 parse, ``torch`` CPU autograd for the gradients,
a float64 shadow for every float32 routine, and the derived known-answer vectors of SURVEY.md App. G.

Each function cites the reference lines it follows (paths relative to the reference root).
"""
from __future__ import annotations

import ctypes
import ctypes.util
import json
import struct
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

# ----------------------------------------------------------------------------------------------
# CSV decode  — tf.io.decode_csv(entry, record_defaults) as used at
# python-scripts/autoencoder-anomaly-detection/Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-34
# ----------------------------------------------------------------------------------------------
_libc = ctypes.CDLL(ctypes.util.find_library("c") or "libc.so.6")
_libc.strtof.restype = ctypes.c_float
_libc.strtof.argtypes = [ctypes.c_char_p, ctypes.POINTER(ctypes.c_char_p)]

REC_OK, REC_FIELD_COUNT, REC_BAD_NUMBER, REC_EMPTY_FIELD, REC_BAD_QUOTE, REC_TRUNCATED, REC_BAD_BRANCH, REC_TOO_LONG = range(8)


def strtof(tok: bytes) -> Tuple[float, bool]:
    """Correctly rounded decimal→binary32 via glibc; returns (value, fully_consumed).

    [upstream] TF parses CSV float fields with ``strings::safe_strtof`` (double-conversion StringToFloat,
    leading/trailing spaces allowed); both are correctly rounded, so they agree bit-for-bit on valid input."""
    t = tok.strip(b" \t")
    if not t or b"\0" in t:
        return 0.0, False
    low = t.lower().lstrip(b"+-")
    if low.startswith(b"0x") or low.startswith(b"nan(") or (low[:1].isalpha() and low not in (b"inf", b"infinity", b"nan")):
        return float("nan"), False  # hex floats / junk are rejected by the kernel as well
    end = ctypes.c_char_p()
    buf = ctypes.create_string_buffer(t)
    v = _libc.strtof(buf, ctypes.byref(end))
    consumed = ctypes.cast(end, ctypes.c_void_p).value - ctypes.addressof(buf)
    return float(v), consumed == len(t)


def split_csv_record(rec: bytes, delim: bytes = b",", use_quote_delim: bool = True) -> Tuple[List[bytes], int]:
    """RFC-4180 split of ONE record the way decode_csv does it ([upstream] tensorflow/core/kernels/decode_csv_op.cc):
    a field starting with '"' runs to the closing quote ('""' = escaped quote); an unquoted field may not contain
    '"', CR or LF.  Returns (fields, status)."""
    fields: List[bytes] = []
    i, n = 0, len(rec)
    d = delim[0]
    while True:
        if use_quote_delim and i < n and rec[i] == 0x22:
            i += 1
            out = bytearray()
            closed = False
            while i < n:
                c = rec[i]
                if c == 0x22:
                    if i + 1 < n and rec[i + 1] == 0x22:
                        out.append(0x22)
                        i += 2
                        continue
                    closed = True
                    i += 1
                    break
                out.append(c)
                i += 1
            if not closed or (i < n and rec[i] != d):
                return fields, REC_BAD_QUOTE
            fields.append(bytes(out))
        else:
            j = i
            while j < n and rec[j] != d:
                if (use_quote_delim and rec[j] == 0x22) or rec[j] in (0x0A, 0x0D):
                    return fields, REC_BAD_QUOTE
                j += 1
            fields.append(rec[i:j])
            i = j
        if i >= n:
            break
        i += 1  # skip delimiter
        if i == n:  # trailing delimiter → one more empty field
            fields.append(b"")
            break
    return fields, REC_OK


def decode_csv(data: np.ndarray, offsets: np.ndarray, kinds: Sequence[int], defaults: Optional[Sequence[Optional[float]]] = None,
               delim: str = ",", use_quote_delim: bool = True):
    """Batch decode: returns (out f32[n, n_float], spans i32[n, n_str, 2], status i32[n]).

    kinds per field: 0 skip, 1 float32, 2 string.  Rows with status != 0 are filled with NaN."""
    n = len(offsets) - 1
    nf = sum(1 for k in kinds if k == 1)
    ns = sum(1 for k in kinds if k == 2)
    out = np.full((n, nf), np.nan, dtype=np.float32)
    spans = np.zeros((n, ns, 2), dtype=np.int32)
    status = np.zeros(n, dtype=np.int32)
    raw = data.tobytes()
    d = delim.encode()
    for r in range(n):
        rec = raw[offsets[r]:offsets[r + 1]]
        fields, st = split_csv_record(rec, d, use_quote_delim)
        if st == REC_OK and len(fields) != len(kinds):
            st = REC_FIELD_COUNT
        if st != REC_OK:
            status[r] = st
            continue
        vals = []
        for i, (k, f) in enumerate(zip(kinds, fields)):
            if k != 1:
                continue
            if len(f) == 0:
                dv = None if defaults is None else defaults[i]
                if dv is None:
                    st = REC_EMPTY_FIELD
                    break
                vals.append(np.float32(dv))
                continue
            v, ok = strtof(f)
            if not ok:
                st = REC_BAD_NUMBER
                break
            vals.append(np.float32(v))
        status[r] = st
        if st == REC_OK:
            out[r] = vals
            # spans of string fields (byte offsets into `data`, quotes excluded)
            if ns:
                pos = int(offsets[r])
                si = 0
                for k, f in zip(kinds, fields):
                    quoted = use_quote_delim and pos < offsets[r + 1] and raw[pos] == 0x22
                    flen = len(f)
                    raw_len = (len(f.replace(b'"', b'""')) + 2) if quoted else flen
                    if k == 2:
                        b = pos + (1 if quoted else 0)
                        spans[r, si] = (b, b + raw_len - (2 if quoted else 0))
                        si += 1
                    pos += raw_len + 1
    return out, spans, status


# ----------------------------------------------------------------------------------------------
# Avro decode — tf.strings.substr(e, 5, -1) + kafka_io.decode_avro(e, schema, dtype=[...])
# python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:12-37 (Avro 1.x binary spec)
# ----------------------------------------------------------------------------------------------
T_INT, T_LONG, T_FLOAT, T_DOUBLE, T_STRING, T_BOOLEAN, T_BYTES, T_NULL = 1, 2, 3, 4, 5, 6, 7, 8


def _read_varint(buf: bytes, p: int, end: int, max_bytes: int) -> Tuple[Optional[int], int]:
    u, shift = 0, 0
    for k in range(max_bytes):
        if p >= end:
            return None, p
        b = buf[p]
        p += 1
        u |= (b & 0x7F) << shift
        shift += 7
        if not (b & 0x80):
            return (u >> 1) ^ -(u & 1), p
    return None, p


def decode_avro(data: np.ndarray, offsets: np.ndarray, prog_fields: Sequence[Tuple[int, int, int]], n_out: int,
                skip_prefix: int = 5, dtype=np.float32):
    """prog_fields: (type, nullable, out) per schema field.  Returns (out dtype[n,n_out], label u8[n], status i32[n]).

    dtype float32 (cardata-v1.py:16-37): numeric values are converted to float32 exactly once (int → float32 like tf.cast at
    cardata-v1.py:61-66, double → float32 round-to-nearest like Keras' input cast).  dtype float64 (cardata-v3.py:53-74 decodes
    with dtype=tf.float64 and casts the int fields to float64, :98-103): doubles as they are, ints exactly.
    Null numeric → 0 ([upstream] tfio default)."""
    n = len(offsets) - 1
    out = np.zeros((n, n_out), dtype=dtype)
    label = np.full(n, 2, dtype=np.uint8)
    status = np.zeros(n, dtype=np.int32)
    raw = data.tobytes()
    for r in range(n):
        p, end = int(offsets[r]) + skip_prefix, int(offsets[r + 1])
        st = REC_OK
        if p > end:
            status[r] = REC_TRUNCATED
            out[r] = np.nan
            continue
        for (typ, nullable, oc) in prog_fields:
            present = True
            if nullable:
                br, p = _read_varint(raw, p, end, 10)
                if br is None:
                    st = REC_TRUNCATED
                    break
                if br not in (0, 1):
                    st = REC_BAD_BRANCH
                    break
                null_branch = 0 if nullable == 1 else 1
                present = br != null_branch
            val = 0.0
            lab = 2
            if present:
                if typ in (T_INT, T_LONG):
                    v, p = _read_varint(raw, p, end, 5 if typ == T_INT else 10)
                    if v is None:
                        st = REC_TRUNCATED
                        break
                    if typ == T_INT:
                        v = (v + 2 ** 31) % 2 ** 32 - 2 ** 31
                    val = dtype(v)
                elif typ == T_FLOAT:
                    if p + 4 > end:
                        st = REC_TRUNCATED
                        break
                    val = np.frombuffer(raw, dtype="<f4", count=1, offset=p)[0]
                    p += 4
                elif typ == T_DOUBLE:
                    if p + 8 > end:
                        st = REC_TRUNCATED
                        break
                    with np.errstate(over="ignore"):
                        val = dtype(np.frombuffer(raw, dtype="<f8", count=1, offset=p)[0])
                    p += 8
                elif typ == T_BOOLEAN:
                    if p + 1 > end:
                        st = REC_TRUNCATED
                        break
                    val = dtype(1.0 if raw[p] else 0.0)
                    p += 1
                elif typ in (T_STRING, T_BYTES):
                    ln, p = _read_varint(raw, p, end, 10)
                    if ln is None or ln < 0 or p + ln > end:
                        st = REC_TRUNCATED
                        break
                    s = raw[p:p + ln]
                    p += ln
                    lab = 0 if s == b"false" else 1 if s == b"true" else 2 if s == b"" else 3
            if oc >= 0:
                out[r, oc] = val
            elif oc == -2:
                label[r] = lab
        status[r] = st
        if st != REC_OK:
            out[r] = np.nan
            label[r] = 2
    return out, label, status


# ----------------------------------------------------------------------------------------------
# normalize_fn — python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:40-131 (f32), cardata-v3.py:78-168 (f64)
# ----------------------------------------------------------------------------------------------
NORM_IDENTITY, NORM_ZERO, NORM_SCALE = 0, 1, 2


def normalize(x: np.ndarray, kind, lo, rng, dtype=np.float32) -> np.ndarray:
    """scale_fn(v, lo, hi) = (v - lo) / (hi - lo) * 2.0 - 1.0 (cardata-v1.py:68-69), four features forced to 0.0
    (:72,78,84,87); every intermediate rounded to `dtype` in the reference's operation order."""
    x = np.asarray(x, dtype=dtype)
    out = np.empty_like(x)
    for j in range(x.shape[1]):
        if kind[j] == NORM_SCALE:
            t = (x[:, j] - dtype(lo[j])).astype(dtype)
            t = (t / dtype(rng[j])).astype(dtype)
            t = (t * dtype(2.0)).astype(dtype)
            out[:, j] = (t - dtype(1.0)).astype(dtype)
        elif kind[j] == NORM_ZERO:
            out[:, j] = 0
        else:
            out[:, j] = x[:, j]
    return out


# ----------------------------------------------------------------------------------------------
# Dense autoencoder forward, reconstruction error, flag
#   model: cardata-v1.py:161-167 / Sensor-Kafka-Consumer…py:16-24;  weights [in,out] row-major, y = act(x @ K + b)
#   mse  : np.mean(np.power(test_x - test_x_predictions, 2), axis=1)    nb2.0:1015
#   flag : 1 if e > threshold_fixed else 0, threshold_fixed = 5         nb2.0:1129,1163
# ----------------------------------------------------------------------------------------------
ACT_LINEAR, ACT_TANH, ACT_RELU = 0, 1, 2


def unflatten(weights: np.ndarray, widths: Sequence[int]):
    ws, p = [], 0
    for i in range(len(widths) - 1):
        a, b = widths[i], widths[i + 1]
        k = weights[p:p + a * b].reshape(a, b)
        p += a * b
        bias = weights[p:p + b]
        p += b
        ws.append((k, bias))
    assert p == len(weights)
    return ws


def _act(z, a):
    if a == ACT_TANH:
        return np.tanh(z)
    if a == ACT_RELU:
        return np.maximum(z, 0)
    return z


def ae_forward(x: np.ndarray, weights: np.ndarray, widths: Sequence[int], acts: Sequence[int], dtype=np.float32,
               return_hidden: bool = False):
    h = np.asarray(x, dtype=dtype)
    hs = [h]
    for (k, b), a in zip(unflatten(np.asarray(weights, dtype=dtype), widths), acts):
        h = _act((h @ k + b).astype(dtype), a).astype(dtype)
        hs.append(h)
    return hs if return_hidden else h


def reconstruction_error(x: np.ndarray, y: np.ndarray, dtype=np.float64) -> np.ndarray:
    x = np.asarray(x, dtype=dtype)
    y = np.asarray(y, dtype=dtype)
    return np.mean(np.power(x - y, 2), axis=1)


def flags(mse: np.ndarray, threshold: float = 5.0) -> np.ndarray:
    return (mse > threshold).astype(np.uint8)


def score(x, weights, widths, acts, norm=None, threshold=5.0, dtype=np.float32):
    """normalise → forward → mse → flag; returns (mse, flag, recon, x_normalised)."""
    xn = normalize(x, *norm, dtype=dtype) if norm is not None else np.asarray(x, dtype=dtype)
    y = ae_forward(xn, weights, widths, acts, dtype=dtype)
    m = reconstruction_error(xn, y, dtype=dtype)
    return m, flags(m, threshold), y, xn


# ----------------------------------------------------------------------------------------------
# Training step: Model.compile('adam', 'mean_squared_error', ['accuracy']) + Model.fit
#   cardata-v1.py:169-171,192-194; .h5 training_config; op list of logs/plugins/profile/*/local.trace
# ----------------------------------------------------------------------------------------------
def ae_loss_and_grads(x: np.ndarray, weights: np.ndarray, widths, acts, l1: float, dtype=np.float64):
    """Loss L = mean_i mse_i + (l1/B) Σ|h1| and dL/dθ (flat, Keras order), plus the 'accuracy' count
    (rows whose argmax(ŷ) == argmax(x))."""
    x = np.asarray(x, dtype=dtype)
    w = np.asarray(weights, dtype=dtype)
    B, D = x.shape
    hs = ae_forward(x, w, widths, acts, dtype=dtype, return_hidden=True)
    y = hs[-1]
    loss = np.mean(np.mean((y - x) ** 2, axis=1)) + dtype(l1) * np.sum(np.abs(hs[1])) / B
    correct = int(np.sum(np.argmax(y, axis=1) == np.argmax(x, axis=1)))
    layers = unflatten(w, widths)
    grads: List[np.ndarray] = [None] * len(layers)  # type: ignore
    dh = (2.0 / (B * D)) * (y - x)
    for l in reversed(range(len(layers))):
        h_out, h_in = hs[l + 1], hs[l]
        if l == 0:
            dh = dh + dtype(l1) / B * np.sign(h_out)  # activity regulariser gradient enters before TanhGrad
        if acts[l] == ACT_TANH:
            dz = dh * (1 - h_out ** 2)
        elif acts[l] == ACT_RELU:
            dz = dh * (h_out > 0)
        else:
            dz = dh
        k, _ = layers[l]
        grads[l] = np.concatenate([(h_in.T @ dz).ravel(), dz.sum(axis=0)])
        dh = dz @ k.T
    return dtype(loss), np.concatenate(grads).astype(dtype), correct


def adam_step(w, m, v, g, t: int, lr=1e-3, b1=0.9, b2=0.999, eps=1e-7, dtype=np.float64):
    """[upstream] ResourceApplyAdam: α_t = lr·√(1-β2^t)/(1-β1^t); m += (1-β1)(g-m); v += (1-β2)(g²-v);
    θ -= α_t·m/(√v + ε)."""
    w, m, v, g = (np.asarray(a, dtype=dtype) for a in (w, m, v, g))
    alpha = dtype(lr) * np.sqrt(dtype(1) - dtype(b2) ** t) / (dtype(1) - dtype(b1) ** t)
    m = m + (dtype(1) - dtype(b1)) * (g - m)
    v = v + (dtype(1) - dtype(b2)) * (g * g - v)
    w = w - alpha * m / (np.sqrt(v) + dtype(eps))
    return w.astype(dtype), m.astype(dtype), v.astype(dtype)


def fit(x, weights, widths, acts, l1, batch, epochs=1, order=None, adam=None, iter0=0, dtype=np.float64):
    """Sequential mini-batch Adam like Model.fit (last batch of an epoch may be short). `order` optionally gives
    the row permutation per epoch (list of index arrays).  Returns (weights, m, v, iter, per-step losses, accs)."""
    lr, b1, b2, eps = adam or (1e-3, 0.9, 0.999, 1e-7)
    w = np.asarray(weights, dtype=dtype).copy()
    m = np.zeros_like(w)
    v = np.zeros_like(w)
    t = iter0
    losses, accs = [], []
    n = x.shape[0]
    for e in range(epochs):
        idx = np.arange(n) if order is None else order[e]
        for s in range(0, n, batch):
            xb = x[idx[s:s + batch]]
            loss, g, correct = ae_loss_and_grads(xb, w, widths, acts, l1, dtype=dtype)
            t += 1
            w, m, v = adam_step(w, m, v, g, t, lr, b1, b2, eps, dtype=dtype)
            losses.append(float(loss))
            accs.append(correct / xb.shape[0])
    return w, m, v, t, np.array(losses), np.array(accs)


# ----------------------------------------------------------------------------------------------
# LSTM stack forward — python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.py:165-178
# ----------------------------------------------------------------------------------------------
def _sigmoid(z):
    return 1.0 / (1.0 + np.exp(-z))


def lstm_layer(x: np.ndarray, kernel, recurrent, bias, return_sequences: bool, dtype=np.float32):
    """[upstream] tf.keras.layers.LSTM(units, activation='relu'), recurrent_activation sigmoid, gate order i,f,c,o:
    z = x_t·W + h·U + b; c = σ(z_f)·c + σ(z_i)·relu(z_c); h = σ(z_o)·relu(c); h0 = c0 = 0."""
    x = np.asarray(x, dtype=dtype)
    B, T, _ = x.shape
    u = recurrent.shape[0]
    W, U, b = (np.asarray(a, dtype=dtype) for a in (kernel, recurrent, bias))
    h = np.zeros((B, u), dtype=dtype)
    c = np.zeros((B, u), dtype=dtype)
    seq = np.empty((B, T, u), dtype=dtype)
    for t in range(T):
        z = (x[:, t] @ W + h @ U + b).astype(dtype)
        i, f, g, o = z[:, :u], z[:, u:2 * u], z[:, 2 * u:3 * u], z[:, 3 * u:]
        c = (_sigmoid(f) * c + _sigmoid(i) * np.maximum(g, 0)).astype(dtype)
        h = (_sigmoid(o) * np.maximum(c, 0)).astype(dtype)
        seq[:, t] = h
    return seq if return_sequences else h


def lstm_param_shapes(features: int, units: Sequence[int]):
    shapes = []
    inp = features
    for u in units:
        shapes += [(inp, 4 * u), (u, 4 * u), (4 * u,)]
        inp = u
    shapes += [(units[-1], features), (features,)]
    return shapes


def lstm_unflatten(weights: np.ndarray, features: int, units: Sequence[int]):
    out, p = [], 0
    for shp in lstm_param_shapes(features, units):
        n = int(np.prod(shp))
        out.append(weights[p:p + n].reshape(shp))
        p += n
    assert p == len(weights)
    return out


def lstm_forward(x: np.ndarray, weights: np.ndarray, features: int = 18, units=(32, 16, 16, 32), dtype=np.float32):
    """LSTM(32,seq) → LSTM(16,last) → RepeatVector(T) → LSTM(16,seq) → LSTM(32,seq) → TimeDistributed(Dense(features))."""
    ws = lstm_unflatten(np.asarray(weights, dtype=dtype), features, units)
    x = np.asarray(x, dtype=dtype)
    T = x.shape[1]
    h = lstm_layer(x, *ws[0:3], return_sequences=True, dtype=dtype)
    h = lstm_layer(h, *ws[3:6], return_sequences=False, dtype=dtype)
    h = np.repeat(h[:, None, :], T, axis=1)  # RepeatVector(look_back)
    h = lstm_layer(h, *ws[6:9], return_sequences=True, dtype=dtype)
    h = lstm_layer(h, *ws[9:12], return_sequences=True, dtype=dtype)
    return (h @ ws[12] + ws[13]).astype(dtype)


# ----------------------------------------------------------------------------------------------
# LSTM stack training step — python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.py:178 (compile: mean_squared_error, adam,
# metrics=['accuracy']) and :191-194 (zip((dataset_x, dataset_y)).batch(1).take(1000); fit(epochs=5)): back-propagation
# through time of the stack above.  Gradients are checked against torch autograd in tests/test_oracle_golden.py.
# ----------------------------------------------------------------------------------------------
def _lstm_layer_fwd_cache(x, W, U, b):
    """Forward of one layer keeping what BPTT needs: gates after their activations, cell states, outputs."""
    B, T, _ = x.shape
    u = U.shape[0]
    h = np.zeros((B, u), dtype=x.dtype)
    c = np.zeros((B, u), dtype=x.dtype)
    cache = dict(x=x, i=[], f=[], g=[], o=[], c=[], h=[], W=W, U=U)
    for t in range(T):
        z = x[:, t] @ W + h @ U + b
        i, f, g, o = _sigmoid(z[:, :u]), _sigmoid(z[:, u:2 * u]), np.maximum(z[:, 2 * u:3 * u], 0), _sigmoid(z[:, 3 * u:])
        c = f * c + i * g
        h = o * np.maximum(c, 0)
        for k, v in zip("ifgoch", (i, f, g, o, c, h)):
            cache[k].append(v)
    return np.stack(cache["h"], axis=1), cache


def _lstm_layer_bwd(dh_seq, cache):
    """dh_seq [B,T,u]: loss gradient w.r.t. every output h_t → (dx [B,T,in], dW, dU, db)."""
    x, W, U = cache["x"], cache["W"], cache["U"]
    B, T, _ = x.shape
    u = U.shape[0]
    dW, dU, db = np.zeros_like(W), np.zeros_like(U), np.zeros(4 * u, dtype=x.dtype)
    dx = np.zeros_like(x)
    dh_next = np.zeros((B, u), dtype=x.dtype)
    dc_next = np.zeros((B, u), dtype=x.dtype)
    for t in range(T - 1, -1, -1):
        i, f, g, o, c = (cache[k][t] for k in "ifgoc")
        c_prev = cache["c"][t - 1] if t > 0 else np.zeros_like(c)
        h_prev = cache["h"][t - 1] if t > 0 else np.zeros_like(c)
        dh = dh_seq[:, t] + dh_next
        do = dh * np.maximum(c, 0)
        dc = dh * o * (c > 0) + dc_next
        dz = np.concatenate([dc * g * i * (1 - i), dc * c_prev * f * (1 - f), dc * i * (g > 0), do * o * (1 - o)], axis=1)
        dW += x[:, t].T @ dz
        dU += h_prev.T @ dz
        db += dz.sum(axis=0)
        dx[:, t] = dz @ W.T
        dh_next = dz @ U.T
        dc_next = dc * f
    return dx, dW, dU, db


def lstm_loss_and_grads(x, y, weights, features: int = 18, units=(32, 16, 16, 32), dtype=np.float64):
    """Mean-squared-error loss of the stack on (x [B,T,F], y [B,T,F]) and its gradient (flat, Keras weight order), plus the
    'accuracy' Keras reports for this compile() call (argmax over the feature axis, averaged over batch and time)."""
    ws = lstm_unflatten(np.asarray(weights, dtype=dtype), features, units)
    x, y = np.asarray(x, dtype=dtype), np.asarray(y, dtype=dtype)
    B, T, _ = x.shape
    h1, c1 = _lstm_layer_fwd_cache(x, *ws[0:3])
    h2seq, c2 = _lstm_layer_fwd_cache(h1, *ws[3:6])
    rep = np.repeat(h2seq[:, -1:, :], T, axis=1)  # return_sequences=False → RepeatVector(T)
    h3, c3 = _lstm_layer_fwd_cache(rep, *ws[6:9])
    h4, c4 = _lstm_layer_fwd_cache(h3, *ws[9:12])
    out = h4 @ ws[12] + ws[13]
    diff = out - y
    loss = float(np.mean(diff * diff))
    acc = float(np.mean(np.argmax(out, axis=-1) == np.argmax(y, axis=-1)))
    dout = 2.0 * diff / diff.size
    dWd = h4.reshape(-1, h4.shape[-1]).T @ dout.reshape(-1, features)
    dbd = dout.reshape(-1, features).sum(axis=0)
    dh4 = dout @ ws[12].T
    dh3, dW4, dU4, db4 = _lstm_layer_bwd(dh4, c4)
    drep, dW3, dU3, db3 = _lstm_layer_bwd(dh3, c3)
    dh2 = np.zeros_like(h2seq)
    dh2[:, -1] = drep.sum(axis=1)
    dh1, dW2, dU2, db2 = _lstm_layer_bwd(dh2, c2)
    _, dW1, dU1, db1 = _lstm_layer_bwd(dh1, c1)
    grads = np.concatenate([a.ravel() for a in (dW1, dU1, db1, dW2, dU2, db2, dW3, dU3, db3, dW4, dU4, db4, dWd, dbd)])
    return loss, grads, acc, out


def lstm_fit(x, y, weights, batch: int = 1, epochs: int = 1, features: int = 18, units=(32, 16, 16, 32), adam=None, iter0: int = 0,
             dtype=np.float64):
    """Mini-batch Adam over consecutive (unshuffled) batches, like model.fit(dataset.batch(batch_size)) on a tf.data pipeline;
    → (weights, m, v, per-epoch mean loss, per-epoch accuracy, iter)."""
    w = np.asarray(weights, dtype=dtype).copy()
    m, v = (np.zeros_like(w), np.zeros_like(w)) if adam is None else (np.asarray(adam[0], dtype=dtype).copy(), np.asarray(adam[1], dtype=dtype).copy())
    t = iter0
    n = x.shape[0]
    losses, accs = [], []
    for _ in range(epochs):
        ls, ac, k = 0.0, 0.0, 0
        for s in range(0, n, batch):
            loss, g, acc, _ = lstm_loss_and_grads(x[s:s + batch], y[s:s + batch], w, features, units, dtype=dtype)
            t += 1
            w, m, v = adam_step(w, m, v, g, t, dtype=dtype)
            ls, ac, k = ls + loss, ac + acc, k + 1
        losses.append(ls / k)
        accs.append(ac / k)
    return w, m, v, losses, accs, t


# ----------------------------------------------------------------------------------------------
# Keras-style initialisers (model_config: GlorotUniform kernels, zero biases, orthogonal recurrent, unit_forget_bias)
# ----------------------------------------------------------------------------------------------
def glorot_uniform(rng: np.random.Generator, shape) -> np.ndarray:
    lim = np.sqrt(6.0 / (shape[0] + shape[1]))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def init_ae_weights(widths: Sequence[int], seed: int = 314) -> np.ndarray:
    g = np.random.Generator(np.random.PCG64(seed))
    parts = []
    for a, b in zip(widths[:-1], widths[1:]):
        parts += [glorot_uniform(g, (a, b)).ravel(), np.zeros(b, dtype=np.float32)]
    return np.concatenate(parts)


def init_lstm_weights(features: int = 18, units=(32, 16, 16, 32), seed: int = 314) -> np.ndarray:
    g = np.random.Generator(np.random.PCG64(seed))
    parts = []
    inp = features
    for u in units:
        parts.append(glorot_uniform(g, (inp, 4 * u)).ravel())
        q, _ = np.linalg.qr(g.standard_normal((4 * u, u)))
        parts.append(q.T.astype(np.float32).ravel())  # [u, 4u] orthogonal rows
        b = np.zeros(4 * u, dtype=np.float32)
        b[u:2 * u] = 1.0  # unit_forget_bias
        parts.append(b)
        inp = u
    parts += [glorot_uniform(g, (units[-1], features)).ravel(), np.zeros(features, dtype=np.float32)]
    return np.concatenate(parts)


# ---- result sink: np.array2string(row) (cardata-v1.py:218-224) ---------------------------------------------------------------
# PINNED: the arithmetic owner of this row of SURVEY §8 (a12) is NumPy itself, which IS installed here, so the oracle for the
# two current generations is the library call; only the NumPy 1.14-1.20 generation (what the reference's tensorflow:2.0.0
# images carried) has to be restated, from the Dragon4 primitives NumPy still exposes (they did not change).
def array2string_rows(x: np.ndarray, compat: int = 0) -> List[str]:
    """compat 0: NumPy >= 2.3 (the installed one), 1: NumPy 1.21 .. 2.2 (printoptions legacy='2.2'), 2: NumPy 1.14 .. 1.20."""
    x = np.asarray(x, dtype=np.float32)
    if compat == 0:
        if tuple(int(v) for v in np.__version__.split(".")[:2]) < (2, 3):
            raise RuntimeError("compat 0 needs NumPy >= 2.3 as the oracle")
        return [np.array2string(r) for r in x]
    if compat == 1:
        with np.printoptions(legacy="2.2"):
            return [np.array2string(r) for r in x]
    return [_array2string_np114(r) for r in x]


def _array2string_np114(row: np.ndarray) -> str:
    """numpy/core/arrayprint.py FloatingFormat + _formatArray of NumPy 1.14-1.20 for a 1-D float32 row, default options:
    like today's except (a) no min_digits, i.e. elements shorter than the row's precision are zero-filled and integer parts
    beyond the shortest form are zero-filled, (b) scalar comparisons against Python floats ran in float64, (c) the scientific
    threshold is 1e8 for every dtype."""
    fin = row[np.isfinite(row)]
    nz = np.abs(fin[fin != 0])
    sci = False
    if len(nz):
        mx, mn = nz.max(), nz.min()
        with np.errstate(over="ignore"):
            sci = bool(float(mx) >= 1e8 or float(mn) < 0.0001 or float(mx / mn) > 1000.0)
    if len(fin) == 0:
        pad_left = pad_right = 0
        prec = exp_size = 0
    elif sci:
        strs = [np.format_float_scientific(v, precision=8, unique=True, trim=".", sign=False) for v in fin]
        fracs, exps = zip(*((s.partition("e")[0], s.partition("e")[2]) for s in strs))
        ints, fr = zip(*(s.split(".") for s in fracs))
        exp_size = max(len(s) for s in exps) - 1
        prec = max(len(s) for s in fr)
        pad_left = max(len(s) for s in ints)
        pad_right = exp_size + 2 + prec
    else:
        strs = [np.format_float_positional(v, precision=8, fractional=True, unique=True, trim=".", sign=False) for v in fin]
        ints, fr = zip(*(s.split(".") for s in strs))
        pad_left = max(len(s) for s in ints)
        pad_right = max(len(s) for s in fr)
    if len(fin) != len(row):
        neginf = bool(np.any(row[np.isinf(row)] < 0))
        offset = pad_right + 1
        pad_left = max(pad_left, 3 - offset, 3 + neginf - offset)

    def fmt(v):
        if not np.isfinite(v):
            s = "nan" if np.isnan(v) else ("-inf" if v < 0 else "inf")
            return " " * (pad_left + pad_right + 1 - len(s)) + s
        if sci:  # trim='k' of that generation zero-filled the shortest digits up to `precision`; today that takes min_digits
            m, _, e = np.format_float_scientific(v, precision=prec, unique=True, trim=".", sign=False, exp_digits=exp_size).partition("e")
            ip, _, fp = m.partition(".")
            return " " * (pad_left - len(ip)) + ip + "." + fp + "0" * (prec - len(fp)) + "e" + e
        return np.format_float_positional(v, precision=8, unique=True, fractional=True, trim=".", sign=False, pad_left=pad_left,
                                          pad_right=pad_right)

    s, line = "", " "
    for i, v in enumerate(row):
        word = fmt(v)
        if len(line) + len(word) > 74 and len(line) > 1:
            s += line.rstrip() + "\n"
            line = " "
        line += word
        if i + 1 < len(row):
            line += " "
    s += line
    return "[" + s[1:] + "]"


# ---- JSON wire format of the raw sensor-data topic --------------------------------------------------------------
REC_BAD_JSON = 8


def decode_json(data: np.ndarray, offsets: np.ndarray, keys: Sequence[str], label_key: Optional[str] = None):
    """One flat JSON object per record (LSTM-…/cardata-v1.json; KSQL value_format='JSON',
    01_installConfluentPlatform.sh:235) → float32[N, len(keys)], label u8[N], status int32[N].
    The parser is Python's ``json`` (RFC 8259, numbers through binary64 like any JSON reader), then the float32 cast of
    cardata-v1.py:16-37; a member that is absent or null gives 0 (KSQL NULL → Avro null → 0), duplicate names: last wins.
    Bytes are taken as Latin-1 so that the grammar, not UTF-8 validity, decides."""
    import json

    def _reject(name):
        raise ValueError(name)  # NaN / Infinity / -Infinity are not JSON

    n = len(offsets) - 1
    out = np.zeros((n, len(keys)), dtype=np.float32)
    label = np.full(n, 2, dtype=np.uint8)
    status = np.zeros(n, dtype=np.int32)
    buf = bytes(np.asarray(data, dtype=np.uint8))
    for i in range(n):
        txt = buf[offsets[i]:offsets[i + 1]].decode("latin-1")
        if any(ch in txt for ch in "\x0b\x0c\x1c\x1d\x1e\x1f\x85\xa0"):
            # Python's str.strip-free scanner treats only ' \t\n\r' as white space, like RFC 8259 — nothing to adjust; kept explicit
            pass
        try:
            obj = json.loads(txt, parse_constant=_reject)
        except (ValueError, RecursionError):
            status[i] = REC_BAD_JSON
            continue
        if not isinstance(obj, dict):
            status[i] = REC_BAD_JSON
            continue
        st = 0
        for j, k in enumerate(keys):
            v = obj.get(k)
            if v is None:
                continue
            if isinstance(v, bool) or not isinstance(v, (int, float)):
                st = st or REC_BAD_NUMBER
                continue
            with np.errstate(over="ignore"):
                out[i, j] = np.float32(float(v))
        if label_key is not None and label_key in obj:
            v = obj[label_key]
            if v is None or v == "":
                label[i] = 2
            elif v is True or v == "true":
                label[i] = 1
            elif v is False or v == "false":
                label[i] = 0
            else:
                label[i] = 3
        status[i] = st
    return out, label, status


# ---- evaluation metrics of the notebooks (nb2.0 cells 23, 24, 27) ------------------------------------------------------------------
# PINNED: the arithmetic owner is scikit-learn, which is installed — the oracle is the library call.
def sk_roc_curve(y_true, y_score, **kw):
    from sklearn.metrics import roc_curve
    return roc_curve(y_true, y_score, **kw)


def sk_precision_recall_curve(y_true, y_score, **kw):
    from sklearn.metrics import precision_recall_curve
    return precision_recall_curve(y_true, y_score, **kw)


def sk_auc(x, y):
    from sklearn.metrics import auc
    return auc(x, y)


def sk_confusion_matrix(y_true, y_pred):
    from sklearn.metrics import confusion_matrix
    return confusion_matrix(y_true, y_pred, labels=[0, 1])
