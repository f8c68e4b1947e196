"""K8: the Avro encoder kernel against the Python encoder of avro_schema.py (byte for byte), decode(encode(x)) round trips, and
the KSQL projection JSON → Avro (01_installConfluentPlatform.sh:235-242)."""
import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import avro_schema as av, decode as dec, schemas, synth

pytestmark = pytest.mark.gpu


def _cols(rows):
    return {name: (np.rint(rows[:, j]).astype(np.int64) if is_int else rows[:, j].astype(np.float32))
            for j, (name, is_int) in enumerate(schemas.FEATURES)}


@pytest.mark.parametrize("n", [1, 31, 32, 33, 50_007])
def test_schema_a_bytes_equal_python_encoder(cuda_device, n):
    rows = synth.sensor_rows(n, seed=n)
    rows[::5, 9] = 20000.0   # three-byte varint
    rows[::7, 17] = -3.0     # negative int
    want_d, want_o = synth.avro_a_batch(rows, schema_id=7)
    rb = dec.encode_avro(rows, schemas.schema_a_json(), schema_id=7)
    np.testing.assert_array_equal(rb.offsets.cpu().numpy(), want_o)
    np.testing.assert_array_equal(rb.data.cpu().numpy(), want_d)
    out, _, st, _ = dec.decode_avro_matrix(rb, schemas.schema_a_json())
    want = rows.copy()
    want[:, [9, 10, 11, 12, 17]] = np.rint(want[:, [9, 10, 11, 12, 17]])
    np.testing.assert_array_equal(out.cpu().numpy(), want)


def test_schema_b_nulls_and_labels(cuda_device):
    n = 20_011
    g = np.random.default_rng(4)
    rows = synth.sensor_rows(n, seed=4)
    rows[:, [9, 10, 11, 12, 17]] = np.rint(rows[:, [9, 10, 11, 12, 17]])
    nullmask = g.random(rows.shape) < 0.03
    lab = g.choice([0, 1, 2], size=n, p=[0.9, 0.07, 0.03]).astype(np.uint8)
    x = rows.copy()
    x[nullmask] = np.nan
    rb = dec.encode_avro(x, schemas.schema_b_json(), lab, schema_id=2)
    # the Python encoder on the same nulls / labels
    prog = av.compile_schema(schemas.schema_b_json())
    cols, nulls = {}, {}
    for j, f in enumerate(prog.fields[:18]):
        is_int = schemas.FEATURES[j][1]
        cols[f.name] = rows[:, j].astype(np.int64) if is_int else rows[:, j].astype(np.float64)
        nulls[f.name] = nullmask[:, j]
    lf = prog.fields[18].name
    nulls[lf] = lab == 2
    strings = {lf: [b"false" if v == 0 else (b"true" if v == 1 else b"") for v in lab]}
    want_d, want_o = av.encode_records(prog, cols, n, nulls=nulls, strings=strings, prefix_schema_id=2)
    np.testing.assert_array_equal(rb.offsets.cpu().numpy(), want_o)
    np.testing.assert_array_equal(rb.data.cpu().numpy(), want_d)
    out, lab2, st, _ = dec.decode_avro_matrix(rb, schemas.schema_b_json())
    want = np.where(nullmask, 0.0, rows).astype(np.float32)
    np.testing.assert_array_equal(out.cpu().numpy(), want)
    np.testing.assert_array_equal(lab2.cpu().numpy(), lab)


def test_ksql_projection_json_to_avro(cuda_device):
    """SENSOR_DATA_S (JSON) → SENSOR_DATA_S_AVRO: the Avro records decode to what the JSON said."""
    n = 30_000
    rows = synth.sensor_rows(n, seed=12)
    labels = (np.arange(n) % 11 == 0).astype(np.int64)
    d, o = synth.json_batch(rows, labels=labels)
    rb = dec.json_to_avro(dec.RecordBatch(d, o), list(schemas.FEATURE_NAMES), schemas.schema_b_json(), label_key="failure_occurred", schema_id=2)
    assert rb.skip_prefix == 5 and rb.n == n
    out, lab, st, _ = dec.decode_avro_matrix(rb, schemas.schema_b_json())
    want = rows.copy()
    want[:, [9, 10, 11, 12, 17]] = np.rint(want[:, [9, 10, 11, 12, 17]])
    np.testing.assert_array_equal(out.cpu().numpy(), want)
    np.testing.assert_array_equal(lab.cpu().numpy(), labels)
    assert dec.encode_avro(np.zeros((0, 18), np.float32), schemas.schema_a_json()).n == 0
