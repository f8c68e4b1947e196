"""The fused decode+score launches (csv, avro_a, avro_b, json at 2^20 / 2^19 records) and the D = 18 score launch of tools/prof_all.py
alone — for refreshing their entries of the committed ncu summary after a change that touches only these kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, b200_sensor as bs

dev = torch.device("cuda", 0)
wl = bench.ScoreWorkload(18, 22, 20)
wl.setup(dev, 0)
wl.step(0)
wl.finish()
for fmt in ("csv", "avro_a", "avro_b", "json"):
    wl = bench.DecodeScoreWorkload(fmt, batch_log2=20 if fmt != "json" else 19, n_batches=1, repeat=1)
    wl.setup(dev, 0)
    wl.step(0)
torch.cuda.synchronize()
print("ok")
