"""ctypes binding of libb2sensor.so (the C ABI declared in include/b2_api.h).

There is deliberately NO fallback: if the CUDA library is missing or fails to load, or a call returns an
error, an exception is raised.  PyTorch tensors are used only to hand over device buffers
(``tensor.data_ptr()``) and the current CUDA stream.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2_LIB") or os.path.join(HERE, "libb2sensor.so")  # B2_LIB: A/B runs against another build

MAX_WIDTH = 64
MAX_FIELDS = 64
NOALLOC_SCRATCH_FLOATS = 160 * 840  # B2_NOALLOC_SCRATCH_FLOATS

ACT_CODES = {"linear": 0, "tanh": 1, "relu": 2}


class B2Error(RuntimeError):
    pass


class AeDesc(C.Structure):
    _fields_ = [("n_layers", C.c_int32), ("widths", C.c_int32 * 9), ("act", C.c_int32 * 8), ("l1_activity", C.c_float)]


class Norm(C.Structure):
    _fields_ = [("n", C.c_int32), ("kind", C.c_int32 * MAX_WIDTH), ("lo", C.c_float * MAX_WIDTH),
                ("range", C.c_float * MAX_WIDTH), ("order", C.c_int32)]


class CsvSpec(C.Structure):
    _fields_ = [("n_fields", C.c_int32), ("kind", C.c_int32 * MAX_FIELDS), ("has_default", C.c_int32 * MAX_FIELDS),
                ("dflt", C.c_float * MAX_FIELDS), ("delim", C.c_int32), ("use_quote_delim", C.c_int32)]


class AvroField(C.Structure):
    _fields_ = [("type", C.c_int8), ("nullable", C.c_int8), ("out", C.c_int8), ("pad", C.c_int8)]


class AvroProg(C.Structure):
    _fields_ = [("n_fields", C.c_int32), ("n_out", C.c_int32), ("f", AvroField * MAX_FIELDS)]


JSON_NAME_MAX = 32


class JsonSpec(C.Structure):
    _fields_ = [("n_keys", C.c_int32), ("n_out", C.c_int32), ("name", (C.c_char * JSON_NAME_MAX) * MAX_FIELDS),
                ("name_len", C.c_uint8 * MAX_FIELDS), ("out", C.c_int8 * MAX_FIELDS)]


class Adam(C.Structure):
    _fields_ = [("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("iter", C.c_int64)]


class LstmDesc(C.Structure):
    _fields_ = [("features", C.c_int32), ("units", C.c_int32 * 4)]


class Partition(C.Structure):
    _fields_ = [("bytes_dev", C.c_void_p), ("offs_dev", C.c_void_p), ("n", C.c_int64), ("mse_dev", C.c_void_p),
                ("flag_dev", C.c_void_p)]


class RunStats(C.Structure):
    _fields_ = [("records", C.c_ulonglong), ("flagged", C.c_ulonglong), ("decode_errors", C.c_ulonglong),
                ("bytes", C.c_ulonglong), ("elapsed_ms", C.c_float)]


# every symbol include/b2_api.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_PROTOTYPES = {
    "b2_ctx_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "b2_ctx_destroy": (None, [_P]),
    "b2_ctx_set_record_bytes_hint": (C.c_int, [_P, C.c_int]),
    "b2_ctx_sm_count": (C.c_int, [_P]),
    "b2_ctx_set_option": (C.c_int, [_P, C.c_char_p, C.c_int]),
    "b2_ctx_get_option": (C.c_int, [_P, C.c_char_p, C.POINTER(C.c_int)]),
    "b2_normalize": (C.c_int, [_P, _P, C.c_int64, C.c_int, C.POINTER(Norm), _P, _P]),
    "b2_selftest_normalize": (C.c_int, [_P, C.POINTER(Norm), C.c_int, C.c_uint64, C.c_uint64, _P, _P]),
    "b2_strerror": (C.c_char_p, [C.c_int]),
    "b2_last_error": (C.c_char_p, []),
    "b2_version": (C.c_char_p, []),
    "b2_ae_create": (C.c_int, [_P, C.POINTER(AeDesc), _P, C.POINTER(_P)]),
    "b2_ae_destroy": (None, [_P]),
    "b2_ae_param_count": (C.c_int, [_P]),
    "b2_ae_get_weights": (C.c_int, [_P, _P]),
    "b2_ae_set_weights": (C.c_int, [_P, _P]),
    "b2_ae_score": (C.c_int, [_P, _P, C.c_int64, C.POINTER(Norm), C.c_float, _P, _P, _P, _P, _P]),
    "b2_decode_csv": (C.c_int, [_P, _P, _P, C.c_int64, C.POINTER(CsvSpec), _P, _P, _P, _P]),
    "b2_decode_avro": (C.c_int, [_P, _P, _P, C.c_int64, C.c_int, C.POINTER(AvroProg), _P, _P, _P, _P]),
    "b2_avro_encode_sizes": (C.c_int, [_P, _P, _P, C.c_int64, C.POINTER(AvroProg), C.c_int, _P, _P]),
    "b2_avro_encode": (C.c_int, [_P, _P, _P, C.c_int64, C.POINTER(AvroProg), C.c_int, C.c_int, _P, _P, _P]),
    "b2_decode_json": (C.c_int, [_P, _P, _P, C.c_int64, C.POINTER(JsonSpec), _P, _P, _P, _P]),
    "b2_decode_score_json": (C.c_int, [_P, _P, _P, C.c_int64, C.POINTER(JsonSpec), C.POINTER(Norm), C.c_float, _P, _P, _P, _P, _P, _P]),
    "b2_decode_score": (C.c_int, [_P, _P, _P, C.c_int64, C.POINTER(CsvSpec), C.POINTER(AvroProg), C.c_int,
                                  C.POINTER(Norm), C.c_float, _P, _P, _P, _P, _P, _P]),
    "b2_ae_fit_steps": (C.c_int, [_P, _P, _P, C.c_int64, C.c_int, C.POINTER(Norm), C.POINTER(Adam), _P, _P, _P]),
    "b2_ae_grad_scratch_bytes": (C.c_int64, [_P, C.c_int64]),
    "b2_ae_grad": (C.c_int, [_P, _P, C.c_int64, C.POINTER(Norm), _P, _P, C.c_int64, _P]),
    "b2_ae_apply": (C.c_int, [_P, _P, C.POINTER(Adam), _P, _P]),
    "b2_ae_step_scratch_bytes": (C.c_int64, [_P, C.c_int64]),
    "b2_ae_exchange_bytes": (C.c_int64, []),
    "b2_ae_step_fused": (C.c_int, [_P, _P, C.c_int64, C.POINTER(Norm), C.POINTER(Adam), _P, C.c_int64, C.c_int64, C.POINTER(C.c_uint64),
                                   C.c_int, C.c_int, C.c_uint32, _P, _P]),
    "b2_ae_grad_noalloc": (C.c_int, [_P, _P, C.c_int64, C.POINTER(Norm), _P, _P]),
    "b2_ae_get_adam": (C.c_int, [_P, _P, _P]),
    "b2_ae_set_adam": (C.c_int, [_P, _P, _P]),
    "b2_lstm_create": (C.c_int, [_P, C.POINTER(LstmDesc), _P, C.POINTER(_P)]),
    "b2_lstm_destroy": (None, [_P]),
    "b2_lstm_param_count": (C.c_int, [_P]),
    "b2_lstm_forward": (C.c_int, [_P, _P, C.c_int64, C.c_int, _P, _P]),
    "b2_lstm_fit_scratch_bytes": (C.c_int64, [C.POINTER(LstmDesc), C.c_int, C.c_int]),
    "b2_lstm_fit_steps": (C.c_int, [_P, C.POINTER(LstmDesc), _P, _P, _P, _P, C.c_int64, _P, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                    C.POINTER(Adam), _P, _P, _P, C.c_int64, _P]),
    "b2_format_rows_bound": (C.c_int64, [C.c_int]),
    "b2_format_rows": (C.c_int, [_P, _P, C.c_int64, C.c_int, C.c_int, _P, C.c_int64, _P, _P, _P, _P]),
    "b2_decode_spans": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.POINTER(CsvSpec), C.POINTER(AvroProg), C.POINTER(JsonSpec), C.c_int, _P, _P, _P, _P, _P]),
    "b2_decode_score_spans": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.POINTER(CsvSpec), C.POINTER(AvroProg), C.POINTER(JsonSpec), C.c_int,
                                        C.POINTER(Norm), C.c_float, _P, _P, _P, _P, _P, _P]),
    "b2_kafka_scan_batches": (C.c_int64, [_P, C.c_int64, _P, _P, C.c_int64]),
    "b2_kafka_index_records": (C.c_int, [_P, _P, C.c_int64, _P, _P, C.c_int64, C.c_int, _P, _P, _P, _P]),
    "b2_kafka_inflate_sizes": (C.c_int, [_P, _P, C.c_int64, _P, C.c_int64, C.c_int, _P, _P, _P]),
    "b2_kafka_inflate": (C.c_int, [_P, _P, C.c_int64, _P, C.c_int64, _P, _P, _P, _P, C.c_int64, _P]),
    "b2_kafka_gather_values": (C.c_int, [_P, _P, _P, _P, C.c_int64, _P, _P, _P]),
    "b2_threshold_sweep_scratch_bytes": (C.c_int64, [C.c_int64]),
    "b2_threshold_sweep": (C.c_int, [_P, _P, _P, C.c_int64, _P, _P, _P, _P, _P, C.c_int64, _P]),
    "b2_confusion_matrix": (C.c_int, [_P, _P, _P, C.c_int64, _P, _P]),
    "b2_lstm_window_count": (C.c_int64, [C.c_int64, C.c_int, C.c_int]),
    "b2_lstm_forward_windows": (C.c_int, [_P, _P, C.c_int64, C.c_int, C.c_int, _P, _P]),
    "b2_partition_run": (C.c_int, [_P, C.POINTER(Partition), C.c_int, C.POINTER(CsvSpec), C.POINTER(AvroProg), C.c_int,
                                   C.POINTER(Norm), C.c_float, C.c_int, _P, C.POINTER(RunStats)]),
    "b2_partition_run_repeat": (C.c_int, [_P, C.POINTER(Partition), C.c_int, C.POINTER(CsvSpec), C.POINTER(AvroProg), C.c_int,
                                          C.POINTER(Norm), C.c_float, C.c_int, C.c_int, _P, C.POINTER(RunStats)]),
}

_lib: Optional[C.CDLL] = None


def declared_symbols():
    return sorted(_PROTOTYPES)


def lib() -> C.CDLL:
    """Load libb2sensor.so (built in-tree by build.py); raises if it is missing — no CPU fallback exists."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B2Error(f"{LIB_PATH} not found: run `python __graft_entry__.py build` (nvcc, sm_100a). "
                          "This package has no CPU fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOTYPES.items():
            fn = getattr(l, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        l = lib()
        raise B2Error(f"{what or 'b2 call'} failed: {l.b2_strerror(rc).decode()} — {l.b2_last_error().decode()}")


def ptr(t) -> Optional[int]:
    """Device pointer of a torch tensor (None → NULL)."""
    return None if t is None else t.data_ptr()


def make_norm(kind=None, lo=None, rng=None, order: int = 0) -> Norm:
    n = Norm()
    if kind is None:
        n.n = 0
        return n
    n.n = len(kind)
    n.order = int(order)   # 0: float32 order (cardata-v1.py), 1: float64 order (cardata-v3.py; fused Avro calls only)
    for i in range(n.n):
        n.kind[i] = int(kind[i])
        n.lo[i] = float(lo[i])
        n.range[i] = float(rng[i])
    return n


def make_avro_prog(prog) -> AvroProg:
    p = AvroProg()
    p.n_fields = len(prog.fields)
    p.n_out = prog.n_out
    for i, f in enumerate(prog.fields):
        p.f[i].type, p.f[i].nullable, p.f[i].out, p.f[i].pad = f.type, f.nullable, f.out, 0
    return p


def make_json_spec(keys, label_key=None) -> JsonSpec:
    """keys: member names routed to output columns 0..len(keys)-1 (in that order); label_key: member holding "true"/"false"."""
    names = [(k, j) for j, k in enumerate(keys)] + ([(label_key, -2)] if label_key else [])
    if not 1 <= len(names) <= MAX_FIELDS:
        raise ValueError(f"1..{MAX_FIELDS} member names, got {len(names)}")
    s = JsonSpec()
    s.n_keys, s.n_out = len(names), len(keys)
    for i, (k, out) in enumerate(names):
        kb = k.encode("latin-1") if isinstance(k, str) else bytes(k)
        if not 1 <= len(kb) <= JSON_NAME_MAX:
            raise ValueError(f"member name {k!r}: 1..{JSON_NAME_MAX} bytes")
        C.memmove(C.addressof(s.name[i]), kb, len(kb))
        s.name_len[i], s.out[i] = len(kb), out
    return s


def make_csv_spec(kinds, defaults=None, delim=",", use_quote_delim=True) -> CsvSpec:
    """kinds: per field 0 skip / 1 float32 / 2 string; defaults: per field float or None (required)."""
    s = CsvSpec()
    s.n_fields = len(kinds)
    if s.n_fields > MAX_FIELDS:
        raise ValueError(f"at most {MAX_FIELDS} CSV fields")
    for i, k in enumerate(kinds):
        s.kind[i] = int(k)
        d = None if defaults is None else defaults[i]
        s.has_default[i] = 0 if d is None else 1
        s.dflt[i] = 0.0 if d is None else float(d)
    s.delim = ord(delim)
    s.use_quote_delim = 1 if use_quote_delim else 0
    return s
