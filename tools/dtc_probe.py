"""Fused decode+score with the score phase on tcgen05 (context option decode_tc) against the FFMA2 kernel: per-record agreement
(mse within the parity tolerance, identical flags away from the threshold, identical statuses / labels / counters) and throughput."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, b200_sensor as bs
dev = torch.device("cuda", 0)
fmts = sys.argv[1:] or ["avro_a", "avro_b", "csv", "json"]
for fmt in fmts:
    res = {}
    for tc in (0, 1):
        bs.keras.set_option("decode_tc", tc)
        for ctas in ((0,) if tc == 0 else (0, 4, 3)):
            bs.keras.set_option("tc_ctas", ctas)
            wl = bench.DecodeScoreWorkload(fmt, batch_log2=20 if fmt != "json" else 19, n_batches=4, repeat=16)
            wl.setup(dev, 0)
            for i in range(2):
                wl.step(i)
            wl.kernel_ms = 0.0
            wl.counters.zero_()
            n = 0
            for i in range(4):
                n += wl.step(i)
            rate = n / (wl.kernel_ms * 1e-3)
            print(f"{fmt} decode_tc={tc} tc_ctas={ctas}: {rate:.4g} records/s  hbm_frac {rate * wl.bytes_per_record / 6538.3e9:.3f}", flush=True)
            if ctas == 0:
                rb = wl.batches[0]
                mse, flag, label, status = bs.decode.decode_score(wl.m, rb, threshold=5.0, **wl.kw)
                torch.cuda.synchronize()
                res[tc] = (mse.cpu().numpy().astype(np.float64), flag.cpu().numpy(), status.cpu().numpy(),
                           None if label is None else label.cpu().numpy(), wl.counters.cpu().tolist())
            del wl
    bs.keras.set_option("tc_ctas", 0)
    a, b = res[0], res[1]
    ok = a[2] == 0
    rel = np.abs(a[0][ok] - b[0][ok]) / np.maximum(a[0][ok], 1e-2)
    safe = ok & (np.abs(a[0] - 5.0) > 1e-4)
    print(f"{fmt}: max rel diff of mse (tc vs ffma2) {rel.max():.3g}; statuses equal {np.array_equal(a[2], b[2])}; nan pattern equal "
          f"{np.array_equal(np.isnan(a[0]), np.isnan(b[0]))}; flags equal away from threshold {np.array_equal(a[1][safe], b[1][safe])}; "
          f"labels equal {a[3] is None or np.array_equal(a[3], b[3])}; counters {a[4]} vs {b[4]}", flush=True)
bs.keras.set_option("decode_tc", 0)
print("ok")
