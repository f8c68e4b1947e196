import json
import os

import pytest

from b200_sensor import avro_schema as av, schemas


def test_generated_schemas_match_reference_files(have_reference):
    if not have_reference:
        pytest.skip("reference checkout not present on this box")
    a = json.load(open("/root/reference/testdata/cardata-v1.avsc"))
    mine = json.loads(schemas.schema_a_json())
    assert [(f["name"], f["type"]) for f in a["fields"]] == [(f["name"], f["type"]) for f in mine["fields"]]
    b = json.load(open("/root/reference/python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.avsc"))
    mine = json.loads(schemas.schema_b_json())
    assert [(f["name"], f["type"]) for f in b["fields"]] == [(f["name"], f["type"]) for f in mine["fields"]]
    hdr = open("/root/reference/testdata/car-sensor-data.csv").readline().strip()
    assert hdr == schemas.CSV_HEADER


def test_field_program():
    p = av.compile_schema(schemas.schema_a_json())
    assert p.n_out == 18 and [f.out for f in p.fields] == list(range(18)) and all(f.nullable == 0 for f in p.fields)
    p = av.compile_schema(schemas.schema_b_json())
    assert p.n_out == 18 and p.fields[-1].out == av.OUT_LABEL and all(f.nullable == 1 for f in p.fields)
    arr = p.as_array()
    assert arr.shape == (64, 4) and arr[18, 0] == av.T_STRING
    with pytest.raises(ValueError):
        av.compile_schema(json.dumps({"type": "record", "name": "x", "fields": [{"name": "a", "type": {"type": "array", "items": "int"}}]}))


def test_norm_table():
    kind, lo, rng = schemas.norm_arrays()
    assert kind.tolist().count(schemas.NORM_ZERO) == 4 and rng[1] == 25.0 and rng[13] == 7.0 and rng[9] == 15.0


def test_reference_table_has_the_shape_the_specialised_kernels_are_compiled_for():
    """csrc/ae_math.cuh compiles one instantiation for the reference's normalize_fn table: features 0, 2, 4, 5 forced to 0.0
    (cardata-v1.py:72, :78, :84, :87 — REF_ZERO_MASK = 0x35), every other feature scaled.  If the table here ever changes shape, the
    generic kernels take over silently: this test is the reminder."""
    import os
    import re

    from b200_sensor import schemas

    kind, lo, rg = schemas.norm_arrays()
    zero = sum(1 << i for i, k in enumerate(kind) if k == schemas.NORM_ZERO)
    assert zero == 0x35 and all(k in (schemas.NORM_ZERO, schemas.NORM_SCALE) for k in kind)
    src = open(os.path.join(os.path.dirname(os.path.abspath(schemas.__file__)), "csrc", "ae_math.cuh")).read()
    assert int(re.search(r"REF_ZERO_MASK = (0x[0-9a-fA-F]+)u", src).group(1), 16) == zero
    assert (rg[kind == schemas.NORM_SCALE] >= 1.0).all()   # reciprocals ≤ 1: no overflow on the way to the quotient
