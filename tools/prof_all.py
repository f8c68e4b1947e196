"""One launch of every hot kernel at a profiling-friendly size (for `ncu --set full`): score d30/d18, fused decode+score
csv/avro_a/avro_b/json, fit epoch kernel, large-batch grad, LSTM forward, result-sink formatter."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, b200_sensor as bs

dev = torch.device("cuda", 0)
for d in (30, 18):
    wl = bench.ScoreWorkload(d, 22, 20)
    wl.setup(dev, 0)
    wl.step(0)
for fmt in ("csv", "avro_a", "avro_b", "json"):
    wl = bench.DecodeScoreWorkload(fmt, batch_log2=18, n_batches=1, repeat=1)
    wl.setup(dev, 0)
    wl.step(0)
m = bench.ScoreWorkload(30, 20, 20).model()
m.compile(optimizer="adam", loss="mean_squared_error", metrics=["accuracy"])
x = bs.synth.credit_rows(128 * 64, seed=1, outlier_frac=0.0)
m.fit(x, epochs=1, batch_size=128, shuffle=False)          # ae_fit_kernel: 64 dependent steps
m.fit(bs.synth.credit_rows(1 << 18, seed=2), epochs=1, batch_size=1 << 18, shuffle=False)  # ae_grad + reduce + apply
print(bench.aux_lstm(dev, 4)["value"])
print(bs.sink.format_rows(torch.relu(torch.randn((1 << 18, 18), device=dev) * 0.5 + 0.3)).text.shape)
torch.cuda.synchronize()
print("ok")
