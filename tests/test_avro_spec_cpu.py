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
