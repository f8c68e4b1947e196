"""ctypes wrapper of oracle/liboracle.so (the C restatement in oracle.c). TEST INFRASTRUCTURE / CPU BASELINE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def lib():
    global _lib
    if _lib is None:
        name = f"liboracle_{_cpu_key()}.so"  # -march=native build: one library per host CPU type
        so = os.path.join(HERE, name)
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(os.path.join(HERE, "oracle.c")):
            subprocess.run(["make", "-C", HERE, "-s", f"OUT={name}"], check=True)
        _lib = C.CDLL(so)
        _lib.orc_score.restype = C.c_longlong
        _lib.orc_decode_score.restype = C.c_longlong
        _lib.orc_max_threads.restype = C.c_int
    return _lib


def _cpu_key() -> str:
    import hashlib

    try:
        flags = [l for l in open("/proc/cpuinfo") if l.startswith(("flags", "model name"))][:2]
    except OSError:
        flags = ["unknown"]
    return hashlib.md5("".join(flags).encode()).hexdigest()[:8]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _i32(v):
    return np.ascontiguousarray(v, dtype=np.int32)


def _f32(v):
    return np.ascontiguousarray(v, dtype=np.float32)


def max_threads() -> int:
    return lib().orc_max_threads()


def set_threads(n: int):
    lib().orc_set_threads(int(n))


def score(x, weights, widths, acts, norm=None, threshold=5.0, want_recon=False):
    x = _f32(x)
    n = x.shape[0]
    mse = np.empty(n, np.float32)
    flag = np.empty(n, np.uint8)
    recon = np.empty_like(x) if want_recon else None
    w, wd, ac = _f32(weights), _i32(widths), _i32(acts)
    if norm is not None:
        nk, nl, nr = _i32(norm[0]), _f32(norm[1]), _f32(norm[2])
    else:
        nk = nl = nr = None
    flagged = lib().orc_score(_p(x), C.c_longlong(n), len(acts), _p(wd), _p(ac), _p(w), 0 if norm is None else len(nk), _p(nk), _p(nl),
                              _p(nr), C.c_float(threshold), _p(mse), _p(flag), _p(recon))
    return mse, flag, recon, int(flagged)


def _csv_spec(kinds, defaults):
    k = _i32(kinds)
    hd = _i32([0 if (defaults is None or defaults[i] is None) else 1 for i in range(len(kinds))])
    df = _f32([0.0 if (defaults is None or defaults[i] is None) else defaults[i] for i in range(len(kinds))])
    return k, hd, df


def decode_csv(data, offsets, kinds, defaults=None, delim=",", use_quote_delim=True):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    offsets = _i32(offsets)
    n = len(offsets) - 1
    k, hd, df = _csv_spec(kinds, defaults)
    nf, ns = int((k == 1).sum()), int((k == 2).sum())
    out = np.empty((n, nf), np.float32)
    span = np.zeros((n, ns, 2), np.int32)
    status = np.empty(n, np.int32)
    lib().orc_decode_csv(_p(data), _p(offsets), C.c_longlong(n), len(k), _p(k), _p(hd), _p(df), ord(delim), int(use_quote_delim),
                         _p(out), _p(span), _p(status))
    return out, span, status


def _avro_fields(prog_fields):
    a = np.zeros((len(prog_fields), 4), dtype=np.int8)
    for i, (t, nl, o) in enumerate(prog_fields):
        a[i] = (t, nl, o, 0)
    return a


def decode_avro(data, offsets, prog_fields, n_out, skip_prefix=5):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    offsets = _i32(offsets)
    n = len(offsets) - 1
    f = _avro_fields(prog_fields)
    out = np.empty((n, n_out), np.float32)
    label = np.empty(n, np.uint8)
    status = np.empty(n, np.int32)
    lib().orc_decode_avro(_p(data), _p(offsets), C.c_longlong(n), skip_prefix, len(prog_fields), _p(f), n_out, _p(out), _p(label),
                          _p(status))
    return out, label, status


def decode_score(data, offsets, weights, widths, acts, norm=None, threshold=5.0, csv=None, avro=None, skip_prefix=5):
    """csv = (kinds, defaults, delim, use_quote) or avro = prog_fields; returns (mse, flag, status, flagged)."""
    data = np.ascontiguousarray(data, dtype=np.uint8)
    offsets = _i32(offsets)
    n = len(offsets) - 1
    mse = np.empty(n, np.float32)
    flag = np.empty(n, np.uint8)
    status = np.empty(n, np.int32)
    w, wd, ac = _f32(weights), _i32(widths), _i32(acts)
    if norm is not None:
        nk, nl, nr = _i32(norm[0]), _f32(norm[1]), _f32(norm[2])
    else:
        nk = nl = nr = None
    if csv is not None:
        kinds, defaults, delim, uq = csv
        k, hd, df = _csv_spec(kinds, defaults)
        nfields, av = len(k), None
    else:
        k = hd = df = None
        delim, uq = ",", True
        av = _avro_fields(avro)
        nfields = len(avro)
    flagged = lib().orc_decode_score(_p(data), _p(offsets), C.c_longlong(n), 0 if csv is not None else 1, skip_prefix, nfields, _p(k),
                                     _p(hd), _p(df), ord(delim), int(uq), _p(av), len(acts), _p(wd), _p(ac), _p(w),
                                     0 if norm is None else len(nk), _p(nk), _p(nl), _p(nr), C.c_float(threshold), _p(mse), _p(flag),
                                     _p(status))
    return mse, flag, status, int(flagged)
