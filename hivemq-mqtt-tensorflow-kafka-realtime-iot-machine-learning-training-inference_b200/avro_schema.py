"""Avro schema → decode "field program", plus a vectorised Avro *encoder* used to build fixtures.

The reference decodes Kafka messages with ``tensorflow_io.kafka.decode_avro(e, schema=..., dtype=[...])``
after stripping the 5-byte Confluent framing with ``tf.strings.substr(e, 5, -1)``
(python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:12-37, cardata-v3.py:49-74,
python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.py:24-49).  Two record schemas are in play
(SURVEY.md §0 fact 4): ``testdata/cardata-v1.avsc`` (18 plain float/int fields) and
``python-scripts/*/cardata-v1.avsc`` (19 ``["null", T]`` unions incl. the FAILURE_OCCURRED string).
Rather than hard-coding either, the schema JSON is compiled into a flat program of at most
``MAX_FIELDS`` entries ``(type, nullable, out)`` that the CUDA decode kernel interprets per record.

Wire format (Avro 1.x binary spec): int/long = zig-zag varint, float/double = IEEE little-endian,
string/bytes = zig-zag varint length + payload, boolean = 1 byte, union = varint branch index + value.
"""
from __future__ import annotations

import json
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

MAX_FIELDS = 64

# field types understood by the kernel (values are part of the C ABI: include/b2_api.h)
T_INT, T_LONG, T_FLOAT, T_DOUBLE, T_STRING, T_BOOLEAN, T_BYTES, T_NULL = 1, 2, 3, 4, 5, 6, 7, 8
_TYPE_CODES = {"int": T_INT, "long": T_LONG, "float": T_FLOAT, "double": T_DOUBLE, "string": T_STRING,
               "boolean": T_BOOLEAN, "bytes": T_BYTES, "null": T_NULL}
# `out` special values
OUT_SKIP, OUT_LABEL = -1, -2
# label codes written by the decode kernel for the string field routed to OUT_LABEL
LABEL_FALSE, LABEL_TRUE, LABEL_NULL, LABEL_OTHER = 0, 1, 2, 3

CONFLUENT_PREFIX = 5  # magic 0x00 + 4-byte big-endian schema id; what substr(e, 5, -1) strips


@dataclass
class AvroField:
    name: str
    type: int
    nullable: int  # 0 plain, 1 = ["null", T] (null is branch 0), 2 = [T, "null"] (null is branch 1)
    out: int       # output column, OUT_SKIP or OUT_LABEL


@dataclass
class AvroProgram:
    fields: List[AvroField]
    n_out: int

    def as_array(self) -> np.ndarray:
        """int8[MAX_FIELDS, 4] rows (type, nullable, out, 0) — the layout of ``b2_avro_prog.f``."""
        a = np.zeros((MAX_FIELDS, 4), dtype=np.int8)
        for i, f in enumerate(self.fields):
            a[i] = (f.type, f.nullable, f.out, 0)
        return a

    @property
    def names(self) -> List[str]:
        return [f.name for f in self.fields]


def _field_type(t) -> Tuple[int, int]:
    if isinstance(t, dict):
        t = t.get("type")
    if isinstance(t, str):
        if t not in _TYPE_CODES:
            raise ValueError(f"unsupported Avro type {t!r}")
        return _TYPE_CODES[t], 0
    if isinstance(t, list):
        if len(t) != 2 or "null" not in t:
            raise ValueError(f"only two-branch unions with null are supported, got {t}")
        null_at = t.index("null")
        code, _ = _field_type(t[1 - null_at])
        return code, 1 if null_at == 0 else 2
    raise ValueError(f"unsupported Avro type spec {t!r}")


def compile_schema(schema_json: str, label_field: Optional[str] = "auto",
                   select: Optional[Sequence[str]] = None) -> AvroProgram:
    """Compile an Avro record schema into the kernel's field program.

    Numeric/boolean fields are routed to consecutive output columns in schema order (the order
    ``decode_avro``'s ``dtype=[...]`` list follows, cardata-v1.py:18-37); the first string field
    (``FAILURE_OCCURRED`` in the KSQL schema) becomes the label unless ``label_field`` names another
    one or is None.  ``select`` restricts/reorders the numeric outputs by (case-insensitive) name.
    """
    sch = json.loads(schema_json) if isinstance(schema_json, (str, bytes)) else schema_json
    if sch.get("type") != "record":
        raise ValueError("top-level Avro schema must be a record")
    if len(sch["fields"]) > MAX_FIELDS:
        raise ValueError(f"record has more than {MAX_FIELDS} fields")
    fields: List[AvroField] = []
    n_out = 0
    label_done = False
    sel = None if select is None else {s.lower(): i for i, s in enumerate(select)}
    for fd in sch["fields"]:
        code, nullable = _field_type(fd["type"])
        out = OUT_SKIP
        if code in (T_INT, T_LONG, T_FLOAT, T_DOUBLE, T_BOOLEAN):
            if sel is None:
                out = n_out
                n_out += 1
            elif fd["name"].lower() in sel:
                out = sel[fd["name"].lower()]
        elif code == T_STRING and not label_done:
            if label_field == "auto" or (label_field is not None and fd["name"].lower() == label_field.lower()):
                out, label_done = OUT_LABEL, True
        fields.append(AvroField(fd["name"], code, nullable, out))
    if sel is not None:
        n_out = len(sel)
    return AvroProgram(fields, n_out)


# ----------------------------------------------------------------------------------------------
# encoder (fixtures / synthetic Kafka batches; the reference's producer side is the device simulator + KSQL)
# ----------------------------------------------------------------------------------------------
def zigzag(v: np.ndarray) -> np.ndarray:
    v = v.astype(np.int64)
    return ((v << 1) ^ (v >> 63)).astype(np.uint64)


def _varint_cols(u: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """uint64[n] → (uint8[n,10] bytes, int32[n] lengths) little-endian base-128 groups."""
    n = u.shape[0]
    out = np.zeros((n, 10), dtype=np.uint8)
    ln = np.ones(n, dtype=np.int32)
    rest = u.copy()
    for k in range(10):
        byte = (rest & np.uint64(0x7F)).astype(np.uint8)
        rest = rest >> np.uint64(7)
        more = rest != 0
        out[:, k] = byte | (more.astype(np.uint8) << 7)
        ln = np.where(more, k + 2, ln)
        if not more.any():
            break
    return out, ln


def encode_records(prog: AvroProgram, columns: Dict[str, np.ndarray], n: int,
                   nulls: Optional[Dict[str, np.ndarray]] = None,
                   strings: Optional[Dict[str, Sequence[bytes]]] = None,
                   prefix_schema_id: Optional[int] = 1) -> Tuple[np.ndarray, np.ndarray]:
    """Vectorised Avro binary encoding of ``n`` records → ``(uint8 data[], int32 offsets[n+1])``.

    ``columns[name]`` holds numeric values per record; ``strings[name]`` a list of byte strings;
    ``nulls[name]`` an optional bool mask of null entries for nullable fields.  With
    ``prefix_schema_id`` each record gets the 5-byte Confluent framing (0x00 + big-endian id).
    """
    nulls = nulls or {}
    strings = strings or {}
    pieces: List[Tuple[np.ndarray, np.ndarray]] = []  # (bytes[n,w], len[n])

    def fixed(b: np.ndarray):
        pieces.append((b, np.full(n, b.shape[1], dtype=np.int32)))

    if prefix_schema_id is not None:
        pre = np.zeros((n, 5), dtype=np.uint8)
        pre[:, 1:] = np.frombuffer(int(prefix_schema_id).to_bytes(4, "big"), dtype=np.uint8)
        fixed(pre)
    for f in prog.fields:
        isnull = nulls.get(f.name)
        if isnull is None:
            isnull = np.zeros(n, dtype=bool)
        if f.nullable:
            null_branch = 0 if f.nullable == 1 else 1
            branch = np.where(isnull, null_branch, 1 - null_branch)
            b, ln = _varint_cols(zigzag(branch))
            pieces.append((b[:, :1], ln))
        elif isnull.any():
            raise ValueError(f"{f.name}: nulls given for a non-nullable field")
        present = ~isnull
        if f.type in (T_INT, T_LONG):
            b, ln = _varint_cols(zigzag(np.asarray(columns[f.name])))
            pieces.append((b, np.where(present, ln, 0).astype(np.int32)))
        elif f.type == T_FLOAT:
            b = np.ascontiguousarray(np.asarray(columns[f.name], dtype="<f4")).view(np.uint8).reshape(n, 4)
            pieces.append((b, np.where(present, 4, 0).astype(np.int32)))
        elif f.type == T_DOUBLE:
            b = np.ascontiguousarray(np.asarray(columns[f.name], dtype="<f8")).view(np.uint8).reshape(n, 8)
            pieces.append((b, np.where(present, 8, 0).astype(np.int32)))
        elif f.type == T_BOOLEAN:
            b = np.asarray(columns[f.name]).astype(np.uint8).reshape(n, 1)
            pieces.append((b, np.where(present, 1, 0).astype(np.int32)))
        elif f.type in (T_STRING, T_BYTES):
            vals = strings[f.name]
            lens = np.fromiter((len(s) for s in vals), dtype=np.int64, count=n)
            b, ln = _varint_cols(zigzag(lens))
            pieces.append((b, np.where(present, ln, 0).astype(np.int32)))
            w = int(lens.max()) if n else 0
            sb = np.zeros((n, max(w, 1)), dtype=np.uint8)
            for i, s in enumerate(vals):
                sb[i, :len(s)] = np.frombuffer(s, dtype=np.uint8)
            pieces.append((sb, np.where(present, lens, 0).astype(np.int32)))
        elif f.type == T_NULL:
            pass
    width = sum(p[0].shape[1] for p in pieces)
    big = np.zeros((n, width), dtype=np.uint8)
    mask = np.zeros((n, width), dtype=bool)
    c = 0
    for b, ln in pieces:
        w = b.shape[1]
        big[:, c:c + w] = b
        mask[:, c:c + w] = np.arange(w)[None, :] < ln[:, None]
        c += w
    data = big[mask]
    offsets = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(mask.sum(axis=1), out=offsets[1:])
    if offsets[-1] >= 2 ** 31:
        raise ValueError("batch exceeds int32 offsets")
    return data, offsets.astype(np.int32)
