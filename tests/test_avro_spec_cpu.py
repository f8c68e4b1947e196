"""The Avro specification's own worked examples (tests/golden/avro_spec_vectors.py) against the CPU oracle's decoder and the
Python encoder that builds every synthetic Avro batch — the anchor that neither shares with the other."""
import json
import os
import sys

import numpy as np

import b200_sensor as bs
from b200_sensor import avro_schema as av
from oracle import cpu_oracle as orc

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import avro_spec_vectors as sv  # noqa: E402


def _decode(schema, payload: bytes):
    prog = av.compile_schema(json.dumps(schema))
    data = np.frombuffer(payload, dtype=np.uint8)
    offs = np.array([0, len(payload)], dtype=np.int32)
    return orc.decode_avro(data, offs, [(f.type, f.nullable, f.out) for f in prog.fields], prog.n_out, skip_prefix=0), prog


def test_zigzag_table():
    for schema in (sv.SCHEMA_ZIGZAG_LONGS, sv.SCHEMA_ZIGZAG_INTS):
        (out, _, st), prog = _decode(schema, sv.h(sv.ZIGZAG_HEX))
        assert st[0] == 0 and out[0].tolist() == [float(v) for v in sv.ZIGZAG_VALUES]
        data, offs = av.encode_records(prog, {f"v{i}": np.array([v]) for i, v in enumerate(sv.ZIGZAG_VALUES)}, 1, prefix_schema_id=None)
        assert data.tobytes() == sv.h(sv.ZIGZAG_HEX) and offs.tolist() == [0, 8]
    assert av.zigzag(np.array(sv.ZIGZAG_VALUES)).tolist() == [0, 1, 2, 3, 4, 127, 128]


def test_spec_record_example():
    (out, lab, st), prog = _decode(sv.SCHEMA_SPEC_RECORD, sv.h(sv.SPEC_RECORD_HEX))
    assert st[0] == 0 and out[0].tolist() == [27.0] and lab[0] == av.LABEL_OTHER  # "foo" is neither true / false / empty
    data, _ = av.encode_records(prog, {"a": np.array([27])}, 1, strings={"b": [b"foo"]}, prefix_schema_id=None)
    assert data.tobytes() == sv.h(sv.SPEC_RECORD_HEX)
    # one byte short: the string runs past the record
    (_, _, st), _ = _decode(sv.SCHEMA_SPEC_RECORD, sv.h(sv.SPEC_RECORD_HEX)[:-1])
    assert st[0] == orc.REC_TRUNCATED


def test_union_examples():
    for hx, want in ((sv.UNION_NULL_HEX, av.LABEL_NULL), (sv.UNION_A_HEX, av.LABEL_OTHER), (sv.UNION_TRUE_HEX, av.LABEL_TRUE),
                     (sv.UNION_FALSE_HEX, av.LABEL_FALSE)):
        (_, lab, st), _ = _decode(sv.SCHEMA_UNION_STRING, sv.h(hx))
        assert st[0] == 0 and lab[0] == want
    prog = av.compile_schema(json.dumps(sv.SCHEMA_UNION_STRING))
    data, offs = av.encode_records(prog, {}, 3, nulls={"s": np.array([True, False, False])}, strings={"s": [b"", b"a", b"true"]},
                                   prefix_schema_id=None)
    assert data.tobytes() == sv.h(sv.UNION_NULL_HEX) + sv.h(sv.UNION_A_HEX) + sv.h(sv.UNION_TRUE_HEX) and offs.tolist() == [0, 1, 4, 10]
    (out, _, st), prog = _decode(sv.SCHEMA_UNION_NUMBERS, sv.h(sv.UNION_NUMBERS_HEX))
    assert st[0] == 0 and out[0].tolist() == sv.UNION_NUMBERS_VALUES
    data, _ = av.encode_records(prog, {"d": np.array([0.5]), "i": np.array([300]), "e": np.array([0.0])}, 1,
                                nulls={"e": np.array([True])}, prefix_schema_id=None)
    assert data.tobytes() == sv.h(sv.UNION_NUMBERS_HEX)


def test_float_double_boolean_null():
    (out, _, st), prog = _decode(sv.SCHEMA_FLOATS, sv.h(sv.FLOATS_HEX))
    assert st[0] == 0 and out[0].tolist() == sv.FLOATS_VALUES
    data, _ = av.encode_records(prog, {"f": np.array([1.0]), "d": np.array([-2.5]), "b": np.array([1]), "g": np.array([-0.15625])}, 1,
                                prefix_schema_id=None)
    assert data.tobytes() == sv.h(sv.FLOATS_HEX)


def test_confluent_framing_is_magic_zero_plus_big_endian_schema_id():
    """cardata-v1.py:13 strips 5 bytes: Confluent's wire format = magic byte 0, 4-byte big-endian schema id, Avro body."""
    prog = av.compile_schema(json.dumps(sv.SCHEMA_SPEC_RECORD))
    data, _ = av.encode_records(prog, {"a": np.array([27])}, 1, strings={"b": [b"foo"]}, prefix_schema_id=0x01020304)
    assert data.tobytes() == bytes([0, 1, 2, 3, 4]) + sv.h(sv.SPEC_RECORD_HEX)


def test_float64_decode_of_cardata_v3_keeps_doubles_and_ints_exact():
    """cardata-v3.py:53-74 decodes with dtype=tf.float64 (ints cast to float64, :98-103): the oracle's float64 mode returns a double
    as it is stored and an int exactly, where the float32 mode (cardata-v1.py) rounds both once."""
    import struct

    schema = {"type": "record", "name": "r", "fields": [{"name": "d", "type": ["null", "double"]}, {"name": "i", "type": ["null", "int"]},
                                                        {"name": "n", "type": ["null", "double"]}]}
    prog = av.compile_schema(json.dumps(schema))
    d, i = 0.1 + 2.0 ** -40, 16777217                       # neither is a float32
    zz = (i << 1) ^ (i >> 31)
    vi = bytearray()
    while zz >= 0x80:
        vi.append((zz & 0x7F) | 0x80)
        zz >>= 7
    vi.append(zz)
    payload = b"\x02" + struct.pack("<d", d) + b"\x02" + bytes(vi) + b"\x00"    # union branch 1, branch 1, null
    data, offs = np.frombuffer(payload, dtype=np.uint8), np.array([0, len(payload)], dtype=np.int32)
    fields = [(f.type, f.nullable, f.out) for f in prog.fields]
    o64, _, st64 = orc.decode_avro(data, offs, fields, prog.n_out, skip_prefix=0, dtype=np.float64)
    o32, _, st32 = orc.decode_avro(data, offs, fields, prog.n_out, skip_prefix=0)
    assert st64[0] == 0 and st32[0] == 0
    assert o64.dtype == np.float64 and o64[0].tolist() == [d, float(i), 0.0]
    assert o32.dtype == np.float32 and o32[0].tolist() == [float(np.float32(d)), float(np.float32(i)), 0.0]
    # and the float64 normalize_fn differs from the float32 one exactly where a rounding step does
    kind, lo, rg = np.array([2, 2, 1]), np.array([0.0, 0.0, 0.0]), np.array([7.0, 3.0, 1.0])
    n64 = orc.normalize(o64, kind, lo, rg, dtype=np.float64)
    assert n64[0, 0] == (d - 0.0) / 7.0 * 2.0 - 1.0 and n64[0, 1] == float(i) / 3.0 * 2.0 - 1.0 and n64[0, 2] == 0.0
