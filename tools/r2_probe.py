"""Fused decode+score: one vs two records per lane (context option decode_r2) — bit equality of the results and throughput."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, b200_sensor as bs
dev = torch.device("cuda", 0)
fmts = sys.argv[1:] or ["avro_a", "avro_b"]
for fmt in fmts:
    res = {}
    for mode in (0, 2, 3):
        bs.keras.set_option("decode_r2", mode)
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
        # per-record outputs with status, through the Python entry point
        rb = wl.batches[0]
        mse, flag, label, status = bs.decode.decode_score(wl.m, rb, threshold=5.0, **wl.kw)
        torch.cuda.synchronize()
        res[mode] = (mse.cpu().numpy().view(np.uint32).copy(), flag.cpu().numpy().copy(), status.cpu().numpy().copy(),
                     None if label is None else label.cpu().numpy().copy(), wl.counters.cpu().tolist())
        print(f"{fmt} decode_r2={mode}: {rate:.4g} records/s  hbm_frac {rate * wl.bytes_per_record / 6538.3e9:.3f}  counters {res[mode][4]}", flush=True)
        del wl
    for mode in (2, 3):
        same = all(np.array_equal(a, b) for a, b in zip(res[0][:3], res[mode][:3])) and res[0][4] == res[mode][4]
        lab_same = res[0][3] is None or np.array_equal(res[0][3], res[mode][3])
        print(f"{fmt}: decode_r2={mode} bit-identical to one record per lane: {same and lab_same}", flush=True)
bs.keras.set_option("decode_r2", -1)
print("ok")
