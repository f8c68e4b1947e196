"""Seeded synthetic record batches shaped like the reference's Kafka topics.

Kafka/HiveMQ are replaced offline by in-memory batches ``(uint8 data[], int32 offsets[n+1])`` — one Kafka
message per record, no trailing newline (python-scripts/autoencoder-anomaly-detection/
Sensor-Kafka-Producer-From-CSV.py:9-12).  Value distributions follow testdata/car-sensor-data.csv and the
normalisation ranges of cardata-v1.py:72-111; text formatting follows the simulator's Java
``Float.toString`` output seen in that CSV (shortest digits, ``E-4`` style exponents, ints for tyre
pressure / firmware).  Everything is driven by ``numpy.random.Generator(PCG64(seed))``; the notebook's
RANDOM_SEED = 314 (nb2.0:76) is the default base seed and partition ``p`` uses ``314 + p``.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from . import avro_schema as av
from . import schemas

BASE_SEED = 314


def rng_for(seed: int = BASE_SEED) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64(seed))


# ----------------------------------------------------------------------------------------------
# pre-decoded rows
# ----------------------------------------------------------------------------------------------
def sensor_rows(n: int, seed: int = BASE_SEED, outlier_frac: float = 0.005) -> np.ndarray:
    """Raw-domain car-sensor rows f32[n,18]: each feature U(lo, hi) widened by 10 %, integer tyre
    pressures, firmware in {1000, 2000} (79:21 like the CSV), ``outlier_frac`` rows with 3 features
    pushed 4-12 ranges out so that ``mse > 5`` fires."""
    g = rng_for(seed)
    x = np.empty((n, schemas.N_FEATURES), dtype=np.float32)
    raw_ranges = {0: (20.0, 60.0), 2: (0.0, 150.0), 4: (200.0, 260.0), 5: (0.0, 40.0)}  # features normalise_fn zeroes
    for j, (kind, lo, hi) in enumerate(schemas.NORM_TABLE):
        if kind != schemas.NORM_SCALE:
            lo, hi = raw_ranges[j]
        w = (hi - lo) * 0.05
        x[:, j] = g.uniform(lo - w, hi + w, size=n).astype(np.float32)
    for j in range(9, 13):
        x[:, j] = g.integers(20, 36, size=n).astype(np.float32)
    x[:, 17] = np.where(g.random(n) < 0.79, 2000.0, 1000.0).astype(np.float32)
    n_out = int(round(n * outlier_frac))
    if n_out:
        rows = g.choice(n, size=n_out, replace=False)
        scaled = [j for j, (k, _, _) in enumerate(schemas.NORM_TABLE) if k == schemas.NORM_SCALE and j not in (9, 10, 11, 12, 17)]
        for r in rows:
            for j in g.choice(scaled, size=3, replace=False):
                _, lo, hi = schemas.NORM_TABLE[j]
                x[r, j] = np.float32(lo + (hi - lo) * g.uniform(4.0, 12.0))
    return x


def credit_rows(n: int, seed: int = BASE_SEED, outlier_frac: float = 0.002) -> np.ndarray:
    """30-wide rows for the shipped ``.h5`` models: 28 PCA-like columns N(0, σ_j) with σ_j decaying
    1.9 → 0.3 plus two N(0,1) columns (Time, Amount after scaling, nb2.0:560-565); ``outlier_frac`` rows ×6."""
    g = rng_for(seed)
    sig = np.concatenate([[1.0], np.linspace(1.9, 0.3, 28), [1.0]]).astype(np.float32)
    x = (g.standard_normal((n, 30), dtype=np.float32) * sig[None, :]).astype(np.float32)
    n_out = int(round(n * outlier_frac))
    if n_out:
        rows = g.choice(n, size=n_out, replace=False)
        x[rows] *= np.float32(6.0)
    return x


# ----------------------------------------------------------------------------------------------
# CSV text
# ----------------------------------------------------------------------------------------------
def java_float_str(v: float) -> str:
    """``Float.toString``-like text of one float32 (decimal for 1e-3 <= |v| < 1e7, else d.dddE±n)."""
    v = np.float32(v)
    if v == 0:
        return "-0.0" if np.signbit(v) else "0.0"
    a = abs(float(v))
    if 1e-3 <= a < 1e7:
        s = np.format_float_positional(v, unique=True, trim="0")
        return s if "." in s else s + ".0"
    s = np.format_float_scientific(v, unique=True, trim="0", exp_digits=1)
    mant, ex = s.split("e")
    if "." not in mant:
        mant += ".0"
    return f"{mant}E{int(ex)}"


def _float_col_to_text(col: np.ndarray):
    import pyarrow as pa
    import pyarrow.compute as pc

    col = np.ascontiguousarray(col, dtype=np.float32)
    s = pc.cast(pa.array(col), pa.string())
    a = np.abs(col)
    special = ((a < 1e-3) | (a >= 1e7) | (col == np.floor(col))) & np.isfinite(col)
    if special.any():
        idx = np.nonzero(special)[0]
        rep = pa.array([java_float_str(col[i]) for i in idx], type=pa.string())
        s = pc.replace_with_mask(s, pa.array(special), rep)
    return s


def csv_batch(rows: np.ndarray, seed: int = BASE_SEED, t0: int = 1567606196, n_cars: int = 100
              ) -> Tuple[np.ndarray, np.ndarray]:
    """Format raw sensor rows f32[n,18] as Kafka CSV messages ``time,car<k>,f0,...,f17``.

    Returns ``(uint8 data[], int32 offsets[n+1])``.  Int-typed features (tyre pressures, firmware) are
    written as integers, the rest like Java ``Float.toString`` (car-sensor-data.csv:2)."""
    import pyarrow as pa
    import pyarrow.compute as pc

    n = rows.shape[0]
    g = rng_for(seed + 7919)
    t = pa.array((t0 + 10 * (np.arange(n) // n_cars)).astype(np.int64))
    car = pa.array((1 + g.integers(0, n_cars, size=n)).astype(np.int64))
    cols = [pc.cast(t, pa.string()), pc.binary_join_element_wise(pa.scalar("car"), pc.cast(car, pa.string()), "")]
    for j, (_, is_int) in enumerate(schemas.FEATURES):
        if is_int:
            cols.append(pc.cast(pa.array(np.rint(rows[:, j]).astype(np.int64)), pa.string()))
        else:
            cols.append(_float_col_to_text(rows[:, j]))
    joined = pc.binary_join_element_wise(*cols, ",")
    if isinstance(joined, pa.ChunkedArray):
        joined = joined.combine_chunks()
    return _arrow_strings_to_batch(joined)


def _arrow_strings_to_batch(arr) -> Tuple[np.ndarray, np.ndarray]:
    import pyarrow as pa

    if arr.type == pa.large_string():
        arr = arr.cast(pa.string())
    _, offs, data = arr.buffers()
    n = len(arr)
    offsets = np.frombuffer(offs, dtype=np.int32, count=n + 1, offset=arr.offset * 4).copy()
    base = int(offsets[0])
    d = np.frombuffer(data, dtype=np.uint8, count=int(offsets[-1]) - base, offset=base).copy()
    return d, (offsets - base).astype(np.int32)


def lines_to_batch(lines) -> Tuple[np.ndarray, np.ndarray]:
    """List of ``bytes``/``str`` messages → ``(data, offsets)``."""
    bs = [l.encode("utf-8") if isinstance(l, str) else bytes(l) for l in lines]
    offsets = np.zeros(len(bs) + 1, dtype=np.int64)
    np.cumsum([len(b) for b in bs], out=offsets[1:])
    data = np.frombuffer(b"".join(bs), dtype=np.uint8).copy()
    return data, offsets.astype(np.int32)


def batch_to_lines(data: np.ndarray, offsets: np.ndarray):
    raw = data.tobytes()
    return [raw[offsets[i]:offsets[i + 1]] for i in range(len(offsets) - 1)]


def credit_csv_batch(rows: np.ndarray, labels: Optional[np.ndarray] = None) -> Tuple[np.ndarray, np.ndarray]:
    """30 float fields with full ``repr`` digits + quoted class label, as consumed by
    Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-34 (``decode_csv`` of 30 floats + 1 string)."""
    import pyarrow as pa
    import pyarrow.compute as pc

    n = rows.shape[0]
    cols = [pc.cast(pa.array(rows[:, j].astype(np.float64)), pa.string()) for j in range(rows.shape[1])]
    lab = np.zeros(n, dtype=np.int64) if labels is None else np.asarray(labels, dtype=np.int64)
    q = pc.binary_join_element_wise(pa.scalar('"'), pc.cast(pa.array(lab), pa.string()), pa.scalar('"'), "")
    joined = pc.binary_join_element_wise(*cols, q, ",")
    if isinstance(joined, pa.ChunkedArray):
        joined = joined.combine_chunks()
    return _arrow_strings_to_batch(joined)


def json_batch(rows: np.ndarray, labels: Optional[np.ndarray] = None, camel_duplicates: bool = False
               ) -> Tuple[np.ndarray, np.ndarray]:
    """The simulator's JSON messages on the raw ``sensor-data`` topic (LSTM-…/cardata-v1.json): one flat object per record,
    members in schema order, floats as Java prints them, tyre pressures / firmware as integers; ``failure_occurred`` as a
    string when ``labels`` (0/1) is given; ``camel_duplicates`` repeats every member under its camelCase name like the fixture."""
    import pyarrow as pa
    import pyarrow.compute as pc

    def camel(name):
        parts = name.split("_")
        return parts[0] + "".join(p.capitalize() if not p.isdigit() else p for p in parts[1:])

    n = rows.shape[0]
    vals = []
    for j, (_, is_int) in enumerate(schemas.FEATURES):
        vals.append(pc.cast(pa.array(np.rint(rows[:, j]).astype(np.int64)), pa.string()) if is_int else _float_col_to_text(rows[:, j]))
    pieces = []
    first = True
    for names in ([schemas.FEATURE_NAMES] + ([[camel(k) for k in schemas.FEATURE_NAMES]] if camel_duplicates else [])):
        for k, v in zip(names, vals):
            pieces.append(pa.scalar(("{" if first else ",") + '"' + k + '":'))
            pieces.append(v)
            first = False
    if labels is not None:
        lab = pa.array(np.where(np.asarray(labels) != 0, "true", "false"))
        pieces += [pa.scalar(',"failure_occurred":"'), lab, pa.scalar('"')]
    pieces.append(pa.scalar("}"))
    joined = pc.binary_join_element_wise(*pieces, "")
    if isinstance(joined, pa.ChunkedArray):
        joined = joined.combine_chunks()
    return _arrow_strings_to_batch(joined)


# ----------------------------------------------------------------------------------------------
# Avro
# ----------------------------------------------------------------------------------------------
def avro_a_batch(rows: np.ndarray, schema_id: int = 1) -> Tuple[np.ndarray, np.ndarray]:
    """Confluent-framed Avro, simulator schema A (testdata/cardata-v1.avsc): 5 + 58 bytes per record."""
    prog = av.compile_schema(schemas.schema_a_json())
    cols = {}
    for j, (name, is_int) in enumerate(schemas.FEATURES):
        cols[name] = np.rint(rows[:, j]).astype(np.int64) if is_int else rows[:, j].astype(np.float32)
    return av.encode_records(prog, cols, rows.shape[0], prefix_schema_id=schema_id)


def avro_b_batch(rows: np.ndarray, seed: int = BASE_SEED, schema_id: int = 2, null_frac: float = 0.0,
                 label_probs=(0.98, 0.015, 0.005)) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Confluent-framed Avro, KSQL schema B (nullable doubles/ints + FAILURE_OCCURRED string).

    Returns ``(data, offsets, label_code[n])`` with label codes as the kernel reports them
    (0 "false", 1 "true", 2 null).  ``null_frac`` of the numeric cells are null (decoded as 0)."""
    g = rng_for(seed + 104729)
    n = rows.shape[0]
    prog = av.compile_schema(schemas.schema_b_json())
    cols, nulls = {}, {}
    for j, (f, (_, is_int)) in enumerate(zip(prog.fields[:-1], schemas.FEATURES)):
        cols[f.name] = np.rint(rows[:, j]).astype(np.int64) if is_int else rows[:, j].astype(np.float64)
        if null_frac > 0:
            nulls[f.name] = g.random(n) < null_frac
    lab = g.choice(3, size=n, p=np.asarray(label_probs) / np.sum(label_probs)).astype(np.uint8)
    strs = [(b"false", b"true", b"")[c] for c in lab]
    nulls[schemas.LABEL_NAME_B] = lab == 2
    data, offs = av.encode_records(prog, cols, n, nulls=nulls, strings={schemas.LABEL_NAME_B: strs},
                                   prefix_schema_id=schema_id)
    return data, offs, lab


# ----------------------------------------------------------------------------------------------
# malformed records (status-code paths)
# ----------------------------------------------------------------------------------------------
def corrupt_csv(data: np.ndarray, offsets: np.ndarray, frac: float = 1e-4, seed: int = BASE_SEED
                ) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Return a copy with ``frac`` of the records damaged (truncated / empty field / bad digit).
    Also returns the indices of the damaged records."""
    g = rng_for(seed + 15485863)
    lines = batch_to_lines(data, offsets)
    n = len(lines)
    k = max(1, int(round(n * frac)))
    bad = np.sort(g.choice(n, size=k, replace=False))
    for i in bad:
        parts = lines[i].split(b",")
        mode = int(g.integers(0, 3))
        if mode == 0:
            parts = parts[: int(g.integers(3, len(parts)))]          # truncated: too few fields
        elif mode == 1:
            parts[int(g.integers(2, len(parts)))] = b""              # empty numeric field, no default
        else:
            j = int(g.integers(2, len(parts)))
            parts[j] = parts[j][:1] + b"x" + parts[j][1:]            # junk inside a number
        lines[i] = b",".join(parts)
    d, o = lines_to_batch(lines)
    return d, o, bad


def corrupt_batch(data: np.ndarray, offsets: np.ndarray, fmt: str, frac: float = 1e-4, seed: int = BASE_SEED
                  ) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """``frac`` of the records of a batch malformed (truncated, empty field, bad digit / bad varint), for any of
    the wire formats ("csv", "avro_a", "avro_b", "json").  Only the damaged records are touched (the batch is re-packed around
    them), so this is cheap at 2^20 records.  Returns (data, offsets, indices of the damaged records)."""
    g = rng_for(seed + 32452843)
    n = len(offsets) - 1
    k = max(1, int(round(n * frac))) if frac > 0 else 0
    if k == 0:
        return data, offsets, np.zeros(0, dtype=np.int64)
    bad = np.sort(g.choice(n, size=k, replace=False))
    lens = np.diff(offsets.astype(np.int64))
    pieces, prev = [], 0
    for i in bad:
        b, e = int(offsets[i]), int(offsets[i + 1])
        rec = data[b:e].tobytes()
        mode = int(g.integers(0, 3))
        if fmt == "csv":
            parts = rec.split(b",")
            if mode == 0:
                parts = parts[: int(g.integers(3, len(parts)))]          # truncated: too few fields
            elif mode == 1:
                parts[int(g.integers(2, len(parts)))] = b""              # empty numeric field, no default
            else:
                j = int(g.integers(2, len(parts)))
                parts[j] = parts[j][:1] + b"x" + parts[j][1:]            # junk inside a number
            rec2 = b",".join(parts)
        elif fmt == "json":
            if mode == 0:
                rec2 = rec[: int(g.integers(2, len(rec) - 1))]           # cut inside the object
            else:
                import re

                digits = [m.start(1) for m in re.finditer(rb":-?(\d)", rec)]   # first digit of a number VALUE (not of a name)
                p = digits[int(g.integers(0, len(digits)))]
                rec2 = rec[:p] + b"x" + rec[p + 1:]                      # junk inside a number
        else:
            if mode == 0 or len(rec) < 48:
                rec2 = rec[: max(5, len(rec) - int(g.integers(1, 9)))]   # truncated inside the last fields
            elif fmt == "avro_b" and mode == 1:
                rec2 = rec[:5] + b"\x04" + rec[6:]                       # union branch index 2 of a two-branch union
            else:
                p = 5 + 36 if fmt == "avro_a" else len(rec) - 12
                rec2 = rec[:p] + b"\xff" * 6 + rec[p + 1:]               # an int varint that never ends
        pieces.append(data[prev:b])
        pieces.append(np.frombuffer(rec2, dtype=np.uint8))
        lens[i] = len(rec2)
        prev = e
    pieces.append(data[prev:])
    offs2 = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=offs2[1:])
    return np.concatenate(pieces), offs2.astype(np.int32), bad

