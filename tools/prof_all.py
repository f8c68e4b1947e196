"""One launch of every kernel of the library at a profiling-friendly size (for `ncu --set full`): score d30 (tcgen05 K1t and
FFMA2) / d18, fused decode+score csv / avro_a / avro_b / json at 2^20 records, decode-only, normalize, fit epoch kernel, fused
data-parallel step, LSTM forward (tcgen05 + FP32) and LSTM training, result-sink formatter, Avro encoder, Kafka framing,
threshold sweep."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, b200_sensor as bs

dev = torch.device("cuda", 0)
k = bs.keras
for d, tc in ((30, -1), (30, 0), (18, -1)):
    k.set_option("score_tc", tc)
    wl = bench.ScoreWorkload(d, 22, 20)
    wl.setup(dev, 0)
    wl.step(0)
    wl.finish()
k.set_option("score_tc", -1)
for fmt in ("csv", "avro_a", "avro_b", "json"):
    wl = bench.DecodeScoreWorkload(fmt, batch_log2=20 if fmt != "json" else 19, n_batches=1, repeat=1)
    wl.setup(dev, 0)
    wl.step(0)
    rb = wl.batches[0]
    if fmt == "csv":
        bs.decode.decode_csv_matrix(rb, [0, 0] + [1] * 18, raise_on_error=False)
    elif fmt == "avro_a":
        out = bs.decode.decode_avro_matrix(rb, bs.schemas.schema_a_json(), raise_on_error=False)[0]
        bs.decode.encode_avro(torch.nan_to_num(out), bs.schemas.schema_a_json())           # avro_sizes + avro_encode kernels
        k.normalize_fn(out, *bs.schemas.norm_arrays())                                      # normalize_kernel
m = bench.ScoreWorkload(30, 20, 20).model()
m.compile(optimizer="adam", loss="mean_squared_error", metrics=["accuracy"])
x = bs.synth.credit_rows(128 * 64, seed=1, outlier_frac=0.0)
m.fit(x, epochs=1, batch_size=128, shuffle=False)          # ae_fit_kernel: 64 dependent steps
m.fit(bs.synth.credit_rows(1 << 18, seed=2), epochs=1, batch_size=1 << 18, shuffle=False)  # ae_grad_kernel with the fused tail
print(bench.aux_lstm(dev, 4)["value"])
k.set_option("lstm_tc", 0)
print(bench.aux_lstm(dev, 4)["value"])
k.set_option("lstm_tc", 2)
ml = k.models.Sequential()
ml.add(k.layers.LSTM(32, activation="relu", input_shape=(1, 18), return_sequences=True))
ml.add(k.layers.LSTM(16, activation="relu", return_sequences=False))
ml.add(k.layers.RepeatVector(1))
ml.add(k.layers.LSTM(16, activation="relu", return_sequences=True))
ml.add(k.layers.LSTM(32, activation="relu", return_sequences=True))
ml.add(k.layers.TimeDistributed(k.layers.Dense(18)))
ml.compile(metrics=["accuracy"], loss="mean_squared_error", optimizer="adam")
rows = np.random.default_rng(1).uniform(-1, 1, (257, 18)).astype(np.float32)
ml.fit(k.Windows(rows, 1, 1), epochs=1, batch_size=1, shuffle=False)                        # lstm_fit_kernel: 256 steps
print(bs.sink.format_rows(torch.relu(torch.randn((1 << 18, 18), device=dev) * 0.5 + 0.3)).text.shape)
vals = [bytes(np.random.default_rng(i).integers(0, 256, 63, dtype=np.uint8)) for i in range(20000)]
bs.kafka.values_from_segment(bs.kafka.write_segment(vals, records_per_batch=500))            # kafka_index + kafka_gather
y = (torch.rand(1 << 20, device=dev) < 0.002).to(torch.uint8)
bs.metrics.threshold_sweep(y, torch.randn(1 << 20, device=dev) ** 2 + y * 6.0)
torch.cuda.synchronize()
print("ok")
