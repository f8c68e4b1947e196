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
