"""K1t probe: the tcgen05 score kernel against the float64 oracle and against the FP32 kernel (error statistics, throughput)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import b200_sensor as bs
from oracle import cpu_oracle as orc
A = [1, 2, 1, 2]
RTOL, ATOL = 1e-5, 1e-7

def stats(mse, ref):
    err = np.abs(mse.astype(np.float64) - ref)
    tol = np.maximum(RTOL * ref, ATOL)
    return dict(bad=int((err > tol).sum()), worst_tol_frac=float((err / tol).max()), max_rel_ge1e2=float((err / np.maximum(ref, 1e-2)).max()))

def timeit(m, xd, mse, flag, n_it=10):
    for _ in range(3): m.score_device(xd, 5.0, mse=mse, flag=flag)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_it): m.score_device(xd, 5.0, mse=mse, flag=flag)
    e1.record(); torch.cuda.synchronize()
    return xd.shape[0] * n_it / (e0.elapsed_time(e1) * 1e-3)

w30 = np.load("tests/golden/ae30_5epochs.npz")["weights"]
m30 = bs.keras.models.load_model("tests/golden/ae30_5epochs.h5")
x30 = bs.synth.credit_rows(1 << 20, seed=5, outlier_frac=0.01)
ref30 = orc.score(x30, w30, [30, 14, 7, 7, 30], A, dtype=np.float64)[0]
k = bs.keras
inp = k.layers.Input(shape=(18,))
h = k.layers.Dense(14, activation="tanh")(inp); h = k.layers.Dense(7, activation="relu")(h); h = k.layers.Dense(7, activation="tanh")(h)
m18 = k.models.Model(inputs=inp, outputs=k.layers.Dense(18, activation="relu")(h), seed=314)
nm = bs.schemas.norm_arrays(); m18.set_normalization(*nm)
x18 = bs.synth.sensor_rows(1 << 20, seed=11)
ref18 = orc.score(x18, m18.get_flat_weights(), [18, 14, 7, 7, 18], A, norm=nm, dtype=np.float64)[0]
for tc in (0, 1, 2):  # 0 = FFMA2 kernel, 1 = tcgen05 with the middle layers in FP32 registers, 2 = all four layers on tcgen05
    bs.keras.set_option("score_tc", tc)
    for name, m, x, ref in (("d30", m30, x30, ref30), ("d18+norm", m18, x18, ref18)):
        mse, flag = m.score(x)
        print(f"score_tc={tc} {name}: {stats(mse, ref)}  flagged {int(flag.sum())} (oracle {(ref > 5).sum()})", flush=True)
    for name, m, d in (("d30", m30, 30), ("d18+norm", m18, 18)):
        n = 1 << 25
        g = torch.Generator(device="cuda").manual_seed(1)
        xd = torch.randn(n, d, device="cuda", generator=g)
        mse = torch.empty(n, device="cuda"); flag = torch.empty(n, dtype=torch.uint8, device="cuda")
        for ctas in ((0,) if tc == 0 else (2, 3, 4, 5, 6, 8)):
            bs.keras.set_option("tc_ctas", ctas)
            r = timeit(m, xd, mse, flag)
            print(f"score_tc={tc} tc_ctas={ctas} {name}: {r:.4g} rows/s  hbm_frac {r * (4 * d + 5) / 6538.3e9:.3f}", flush=True)
        bs.keras.set_option("tc_ctas", 0)
        del xd, mse, flag
print("ok")
