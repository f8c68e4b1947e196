"""The Avro specification's worked examples (tests/golden/avro_spec_vectors.py) through the CUDA decoder (b2_decode_avro) and
the CUDA encoder (b2_avro_encode): bytes that no code of this repository produced."""
import json
import os
import sys

import numpy as np
import pytest
import torch

from b200_sensor import avro_schema as av, decode as dec

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import avro_spec_vectors as sv  # noqa: E402

pytestmark = pytest.mark.gpu


def _decode(schema, payloads):
    rb = dec.RecordBatch.from_messages(payloads)
    out, lab, st, _ = dec.decode_avro_matrix(rb, json.dumps(schema), skip_prefix=0, raise_on_error=False)
    return out.cpu().numpy(), lab.cpu().numpy(), st.cpu().numpy()


def _encode(schema, rows, label=None):
    rb = dec.encode_avro(np.asarray(rows, dtype=np.float32), json.dumps(schema), label, confluent_prefix=False)
    d, o = rb.data.cpu().numpy().tobytes(), rb.offsets.cpu().numpy()
    return [d[o[i]:o[i + 1]] for i in range(rb.n)]


def test_zigzag_table(cuda_device):
    for schema in (sv.SCHEMA_ZIGZAG_LONGS, sv.SCHEMA_ZIGZAG_INTS):
        out, _, st = _decode(schema, [sv.h(sv.ZIGZAG_HEX)] * 33)          # 33: a full warp chunk and a ragged one
        assert not st.any() and (out == np.array(sv.ZIGZAG_VALUES, dtype=np.float32)).all()
        assert _encode(schema, [sv.ZIGZAG_VALUES] * 33) == [sv.h(sv.ZIGZAG_HEX)] * 33


def test_spec_record_example(cuda_device):
    out, lab, st = _decode(sv.SCHEMA_SPEC_RECORD, [sv.h(sv.SPEC_RECORD_HEX), sv.h(sv.SPEC_RECORD_HEX)[:-1]])
    assert st.tolist() == [0, 5] and out[0].tolist() == [27.0] and lab[0] == av.LABEL_OTHER  # 5 = B2_REC_TRUNCATED


def test_union_examples(cuda_device):
    # the CUDA decoder wants at least one numeric column: the spec's union examples follow a long field holding 1 (zig-zag 02) —
    # a record is the plain concatenation of its fields
    schema = {"type": "record", "name": "u", "fields": [{"name": "n", "type": "long"}] + sv.SCHEMA_UNION_STRING["fields"]}
    msgs = [b"\x02" + sv.h(x) for x in (sv.UNION_NULL_HEX, sv.UNION_A_HEX, sv.UNION_TRUE_HEX, sv.UNION_FALSE_HEX)]
    out, lab, st = _decode(schema, msgs)
    assert not st.any() and lab.tolist() == [av.LABEL_NULL, av.LABEL_OTHER, av.LABEL_TRUE, av.LABEL_FALSE] and (out == 1.0).all()
    out, _, st = _decode(sv.SCHEMA_UNION_NUMBERS, [sv.h(sv.UNION_NUMBERS_HEX)])
    assert not st.any() and out[0].tolist() == sv.UNION_NUMBERS_VALUES
    assert _encode(sv.SCHEMA_UNION_NUMBERS, [[0.5, 300.0, float("nan")]]) == [sv.h(sv.UNION_NUMBERS_HEX)]


def test_float_double_boolean_null(cuda_device):
    out, _, st = _decode(sv.SCHEMA_FLOATS, [sv.h(sv.FLOATS_HEX)])
    assert not st.any() and out[0].tolist() == sv.FLOATS_VALUES
    assert _encode(sv.SCHEMA_FLOATS, [sv.FLOATS_VALUES]) == [sv.h(sv.FLOATS_HEX)]
