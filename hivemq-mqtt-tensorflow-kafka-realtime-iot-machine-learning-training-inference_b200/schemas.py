"""The car-sensor record schema, its two Avro flavours and the normalisation table.

* Feature order: README.md:82 / testdata/car-sensor-data.csv:1 / ``tf.stack`` order in
  python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:113-131.
* Schema A = testdata/cardata-v1.avsc:6-96 (``com.hivemq.avro.CarData``: plain float/int fields).
* Schema B = python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.avsc:5-157
  (``KsqlDataSourceSchema``: every field ``["null", T]``, doubles, + FAILURE_OCCURRED string).
* Normalisation = ``normalize_fn`` cardata-v1.py:68-111.

The schema JSON is *generated* here from the field table (the reference files are not shipped with
this repo); tests/test_schemas.py checks the generated field lists against the reference files
whenever /root/reference is present.
"""
from __future__ import annotations

import json
from typing import List, Tuple

import numpy as np

# (csv/avro-A name, is_int)
FEATURES: List[Tuple[str, bool]] = [
    ("coolant_temp", False), ("intake_air_temp", False), ("intake_air_flow_speed", False),
    ("battery_percentage", False), ("battery_voltage", False), ("current_draw", False),
    ("speed", False), ("engine_vibration_amplitude", False), ("throttle_pos", False),
    ("tire_pressure_1_1", True), ("tire_pressure_1_2", True), ("tire_pressure_2_1", True),
    ("tire_pressure_2_2", True),
    ("accelerometer_1_1_value", False), ("accelerometer_1_2_value", False),
    ("accelerometer_2_1_value", False), ("accelerometer_2_2_value", False),
    ("control_unit_firmware", True),
]
N_FEATURES = len(FEATURES)  # 18
FEATURE_NAMES = [n for n, _ in FEATURES]
CSV_HEADER = "time,car," + ",".join(FEATURE_NAMES)
CSV_N_FIELDS = 2 + N_FEATURES

# KSQL upper-cases names; the tyre/accelerometer indices are fused (TIRE_PRESSURE11, ACCELEROMETER11_VALUE)
_KSQL_NAMES = [
    "COOLANT_TEMP", "INTAKE_AIR_TEMP", "INTAKE_AIR_FLOW_SPEED", "BATTERY_PERCENTAGE", "BATTERY_VOLTAGE",
    "CURRENT_DRAW", "SPEED", "ENGINE_VIBRATION_AMPLITUDE", "THROTTLE_POS",
    "TIRE_PRESSURE11", "TIRE_PRESSURE12", "TIRE_PRESSURE21", "TIRE_PRESSURE22",
    "ACCELEROMETER11_VALUE", "ACCELEROMETER12_VALUE", "ACCELEROMETER21_VALUE", "ACCELEROMETER22_VALUE",
    "CONTROL_UNIT_FIRMWARE",
]
LABEL_NAME_B = "FAILURE_OCCURRED"


def schema_a_json() -> str:
    """Simulator schema (plain ``float`` / ``int``), field-for-field like testdata/cardata-v1.avsc."""
    fields = [{"name": n, "type": "int" if is_int else "float"} for n, is_int in FEATURES]
    return json.dumps({"type": "record", "namespace": "com.hivemq.avro", "name": "CarData", "version": "1",
                       "fields": fields})


def schema_b_json() -> str:
    """KSQL schema (nullable unions, ``double``/``int`` + label string), like python-scripts/*/cardata-v1.avsc."""
    fields = [{"name": n, "type": ["null", "int" if is_int else "double"], "default": None}
              for n, (_, is_int) in zip(_KSQL_NAMES, FEATURES)]
    fields.append({"name": LABEL_NAME_B, "type": ["null", "string"], "default": None})
    return json.dumps({"type": "record", "name": "KsqlDataSourceSchema",
                       "namespace": "io.confluent.ksql.avro_schemas", "fields": fields})


# ---- normalisation table (cardata-v1.py:72-111) ---------------------------------------------------
NORM_IDENTITY, NORM_ZERO, NORM_SCALE = 0, 1, 2  # part of the C ABI (b2_norm.kind)

# per feature: (kind, lo, hi)
NORM_TABLE = [
    (NORM_ZERO, 0.0, 0.0),         # coolant_temp                :72
    (NORM_SCALE, 15.0, 40.0),      # intake_air_temp             :75
    (NORM_ZERO, 0.0, 0.0),         # intake_air_flow_speed       :78
    (NORM_SCALE, 0.0, 100.0),      # battery_percentage          :81
    (NORM_ZERO, 0.0, 0.0),         # battery_voltage             :84
    (NORM_ZERO, 0.0, 0.0),         # current_draw                :87
    (NORM_SCALE, 0.0, 50.0),       # speed                       :90
    (NORM_SCALE, 0.0, 7500.0),     # engine_vibration_amplitude  :93
    (NORM_SCALE, 0.0, 1.0),        # throttle_pos                :96
    (NORM_SCALE, 20.0, 35.0), (NORM_SCALE, 20.0, 35.0),   # tire_pressure_11/12   :99-100
    (NORM_SCALE, 20.0, 35.0), (NORM_SCALE, 20.0, 35.0),   # tire_pressure_21/22   :101-102
    (NORM_SCALE, 0.0, 7.0), (NORM_SCALE, 0.0, 7.0),       # accelerometer_11/12   :105-106
    (NORM_SCALE, 0.0, 7.0), (NORM_SCALE, 0.0, 7.0),       # accelerometer_21/22   :107-108
    (NORM_SCALE, 1000.0, 2000.0),  # control_unit_firmware       :111
]


def norm_arrays():
    """(kind int32[18], lo f32[18], range f32[18]) with range = hi - lo evaluated like the script (:69)."""
    kind = np.array([k for k, _, _ in NORM_TABLE], dtype=np.int32)
    lo = np.array([l for _, l, _ in NORM_TABLE], dtype=np.float32)
    rng = np.array([(h - l) if k == NORM_SCALE else 1.0 for k, l, h in NORM_TABLE], dtype=np.float32)
    return kind, lo, rng
