"""A small pass over every kernel for compute-sanitizer (memcheck / racecheck / synccheck): tiny sizes, ragged tails, a bad record,
the oversize retry path.  Usage: compute-sanitizer --tool memcheck python tools/sanitize_smoke.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200_sensor as bs
from b200_sensor import decode as dec, schemas, synth

k = bs.keras
dev = torch.device("cuda", 0)
m30 = k.models.load_model("tests/golden/ae30_5epochs.h5")
x30 = synth.credit_rows(1000 + 37, seed=1)
for tc in (-1, 0, 2):
    k.set_option("score_tc", tc)
    m30.score(x30)
k.set_option("score_tc", -1)
m30.predict(x30[:300])
inp = k.layers.Input(shape=(18,))
h = k.layers.Dense(14, activation="tanh")(inp); h = k.layers.Dense(7, activation="relu")(h); h = k.layers.Dense(7, activation="tanh")(h)
m18 = k.models.Model(inputs=inp, outputs=k.layers.Dense(18, activation="relu")(h), seed=3)
m18.set_normalization(*schemas.norm_arrays())
rows = synth.sensor_rows(700 + 13, seed=2)
m18.score(rows)
for fmt in ("csv", "avro_a", "avro_b", "json"):
    if fmt == "csv":
        d, o = synth.csv_batch(rows, seed=2); kw = dict(csv_kinds=[0, 0] + [1] * 18); skip = 0
    elif fmt == "avro_a":
        d, o = synth.avro_a_batch(rows); kw = dict(avro_schema=schemas.schema_a_json()); skip = 5
    elif fmt == "avro_b":
        d, o, _ = synth.avro_b_batch(rows, seed=2); kw = dict(avro_schema=schemas.schema_b_json()); skip = 5
    else:
        d, o = synth.json_batch(rows); kw = dict(json_keys=list(schemas.FEATURE_NAMES), json_label="failure_occurred"); skip = 0
    d, o, _ = synth.corrupt_batch(d, o, fmt, 0.01, seed=3)
    rb = dec.RecordBatch(d, o, skip_prefix=skip)
    for r2 in ((0, 2, 3) if fmt.startswith("avro") else (0,)):
        k.set_option("decode_r2", r2)
        dec.decode_score(m18, rb, **kw)
    k.set_option("decode_r2", 0)
lines = [l for l in synth.batch_to_lines(*synth.csv_batch(rows[:200], seed=5))]
f = lines[7].split(b",")
lines[7] = b",".join([f[0], b"q" * 9000] + f[2:])
dec.decode_csv_matrix(dec.RecordBatch.from_messages(lines), [0, 0] + [1] * 18, raise_on_error=False)
out = dec.decode_avro_matrix(dec.RecordBatch(*synth.avro_a_batch(rows), skip_prefix=5), schemas.schema_a_json())[0]
dec.encode_avro(out, schemas.schema_a_json())
k.normalize_fn(out, *schemas.norm_arrays())
m = k.models.Model(inputs=(i2 := k.layers.Input(shape=(30,))), outputs=k.layers.Dense(30, activation="relu")(k.layers.Dense(7, activation="tanh")(k.layers.Dense(7, activation="relu")(k.layers.Dense(14, activation="tanh")(i2)))), seed=4)
m.compile(optimizer="adam", loss="mean_squared_error", metrics=["accuracy"])
m.fit(x30[:300], epochs=1, batch_size=128, shuffle=False)
m.fit(x30, epochs=1, batch_size=512, shuffle=False)
for tcl in (1, 0):
    k.set_option("lstm_tc", tcl)
    ml = k.models.Sequential(seed=1)
    ml.add(k.layers.LSTM(32, activation="relu", input_shape=(3, 18), return_sequences=True))
    ml.add(k.layers.LSTM(16, activation="relu", return_sequences=False))
    ml.add(k.layers.RepeatVector(3))
    ml.add(k.layers.LSTM(16, activation="relu", return_sequences=True))
    ml.add(k.layers.LSTM(32, activation="relu", return_sequences=True))
    ml.add(k.layers.TimeDistributed(k.layers.Dense(18)))
    xl = np.random.default_rng(0).uniform(-1, 1, (150, 3, 18)).astype(np.float32)
    ml.predict(xl)
k.set_option("lstm_tc", 2)
ml.compile(metrics=["accuracy"], loss="mean_squared_error", optimizer="adam")
ml.fit(xl[:20], xl[:20], epochs=1, batch_size=4, shuffle=False)
bs.sink.format_rows(m30.predict(x30[:200]))
bs.kafka.values_from_segment(bs.kafka.write_segment([b"abc" * i for i in range(300)], records_per_batch=64, headers=True))
y = (torch.rand(5000, device=dev) < 0.1).to(torch.uint8)
bs.metrics.threshold_sweep(y, torch.rand(5000, device=dev))
torch.cuda.synchronize()
print("sanitize_smoke ok")
