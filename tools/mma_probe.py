"""Fused decode+score with the score phase as warp-level TF32 MMAs (context option decode_mma) against the FFMA2 score phase:
agreement of the per-record results and throughput through the partition driver."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, b200_sensor as bs
dev = torch.device("cuda", 0)
fmts = [a for a in sys.argv[1:] if not a.isdigit()] or ["avro_a", "avro_b", "csv"]
modes = [int(a) for a in sys.argv[1:] if a.isdigit()] or [1]
for fmt in fmts:
    res = {}
    for mode in [0] + modes:
        bs.keras.set_option("decode_mma", mode)
        wl = bench.DecodeScoreWorkload(fmt, n_batches=4, repeat=16)
        wl.setup(dev, 0)
        for i in range(2):
            wl.step(i)
        wl.kernel_ms = 0.0
        wl.counters.zero_()
        n = 0
        for i in range(4):
            n += wl.step(i)
        rate = n / (wl.kernel_ms * 1e-3)
        rb = wl.batches[0]
        mse, flag, label, status = bs.decode.decode_score(wl.m, rb, threshold=5.0, **wl.kw)
        torch.cuda.synchronize()
        res[mode] = (mse.cpu().numpy().copy(), flag.cpu().numpy().copy(), status.cpu().numpy().copy(),
                     None if label is None else label.cpu().numpy().copy(), wl.counters.cpu().tolist())
        print(f"{fmt} decode_mma={mode}: {rate:.4g} records/s  hbm_frac {rate * wl.bytes_per_record / 6538.3e9:.3f}  counters {res[mode][4]}", flush=True)
        del wl
    for mode in modes:
        a, b = res[0], res[mode]
        ok = a[2] == 0
        rel = np.abs(a[0][ok] - b[0][ok]) / np.maximum(np.abs(a[0][ok]), 1e-2)
        nan_same = np.array_equal(np.isnan(a[0]), np.isnan(b[0]))
        print(f"{fmt} decode_mma={mode}: max relative difference of the score {rel.max():.3g} (mean {rel.mean():.3g}); NaN pattern same {nan_same}; "
              f"status same {np.array_equal(a[2], b[2])}; labels same {a[3] is None or np.array_equal(a[3], b[3])}; "
              f"flags differing {int((a[1] != b[1]).sum())}; counters {a[4]} vs {b[4]}", flush=True)
bs.keras.set_option("decode_mma", 0)
print("ok")
