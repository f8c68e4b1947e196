"""GPU diagnostic: error of the fused score kernel against the float64 oracle (accurate vs fast tanh)."""
import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import b200_sensor as bs
from oracle import cpu_oracle as orc
A = [1, 2, 1, 2]
w30 = np.load("tests/golden/ae30_5epochs.npz")["weights"]
x30 = bs.synth.credit_rows(1 << 20, seed=5)
m30 = bs.keras.models.load_model("tests/golden/ae30_5epochs.h5")
ref = orc.score(x30, w30, [30, 14, 7, 7, 30], A, dtype=np.float64)[0]
np32 = orc.score(x30, w30, [30, 14, 7, 7, 30], A, dtype=np.float32)[0]
mse, _ = m30.score(x30)
rel = np.abs(mse - ref) / np.maximum(ref, 1e-2)
print("FAST_TANH=%s d30: max rel %.3g  mean rel %.3g  (numpy f32 restatement: max rel %.3g)" % (
    os.environ.get("B2_FAST_TANH", "0"), rel.max(), rel.mean(), (np.abs(np32 - ref) / np.maximum(ref, 1e-2)).max()))
