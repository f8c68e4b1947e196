"""K1 parity: fused normalise → Dense stack → MSE → flag against the CPU oracle, through the C ABI.

Tolerances (BASELINE.json north_star): reconstruction error within 1e-5 RELATIVE of the oracle (float64 shadow)
— for records whose mse is below 1e-2 FP32 rounding alone exceeds that, so those are held to 1e-7 ABSOLUTE
instead; flags identical except for records within 1e-5*5 of the threshold; predict rows within 2e-6 absolute.
"""
import math
import os

import numpy as np
import pytest
import torch

import b200_sensor as bs
from oracle import cpu_oracle as orc

pytestmark = pytest.mark.gpu

ACTS = [orc.ACT_TANH, orc.ACT_RELU, orc.ACT_TANH, orc.ACT_RELU]
RTOL, ATOL_SMALL, THRESH = 1e-5, 1e-7, 5.0


def _check(mse, flag, ref64, thr=THRESH):
    err = np.abs(mse.astype(np.float64) - ref64)
    tol = np.maximum(RTOL * ref64, ATOL_SMALL)
    bad = err > tol
    assert not bad.any(), f"{bad.sum()} records out of tolerance; worst rel {np.max(err / np.maximum(ref64, 1e-30)):.3g}"
    safe = np.abs(ref64 - thr) > RTOL * thr
    np.testing.assert_array_equal(flag[safe], (ref64 > thr).astype(np.uint8)[safe])


def _model30(ae30, golden_dir):
    return bs.keras.models.load_model(os.path.join(golden_dir, "ae30_5epochs.h5"))


def _model18(seed=314):
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    e = k.layers.Dense(14, activation="tanh", activity_regularizer=k.regularizers.l1(1e-7))(inp)
    e = k.layers.Dense(7, activation="relu")(e)
    d = k.layers.Dense(7, activation="tanh")(e)
    d = k.layers.Dense(18, activation="relu")(d)
    return k.models.Model(inputs=inp, outputs=d, seed=seed)


def test_appendix_g_known_answers(cuda_device, ae30, golden_dir):
    m = _model30(ae30, golden_dir)
    x = np.stack([np.zeros(30), np.ones(30), (np.arange(30) - 15) / 10]).astype(np.float32)
    mse, flag = m.score(x)
    np.testing.assert_allclose(mse, [0.00108109054, 1.38512659, 0.901749313], rtol=0, atol=1e-6)
    assert flag.tolist() == [0, 0, 0]


@pytest.mark.parametrize("n", [1, 2, 31, 255, 256, 257, 511, 513, 4096 + 77, 200_003])
def test_score30_sizes_vs_oracle(cuda_device, ae30, golden_dir, n):
    m = _model30(ae30, golden_dir)
    x = bs.synth.credit_rows(n, seed=100 + n)
    mse, flag = m.score(x)
    ref = orc.score(x, ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float64)[0]
    _check(mse, flag, ref)


def test_score30_flags_fire_and_threshold_band(cuda_device, ae30, golden_dir):
    m = _model30(ae30, golden_dir)
    x = bs.synth.credit_rows(1 << 17, seed=7, outlier_frac=0.01)
    ref = orc.score(x, ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float64)[0]
    assert (ref > 5).sum() > 100, "synthetic outliers must exercise the flag path"
    # 64 rows constructed to sit within ±1e-4 of the threshold: scale a flagged row until mse ≈ 5 (bisection on the oracle)
    rows = np.nonzero(ref > 8)[0][:64]
    for r in rows:
        lo, hi, base = 0.0, 1.0, x[r].copy()
        for _ in range(40):
            mid = (lo + hi) / 2
            v = orc.score((base * mid)[None], ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float64)[0][0]
            lo, hi = (mid, hi) if v < 5 else (lo, mid)
        x[r] = (base * hi).astype(np.float32)
    ref = orc.score(x, ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float64)[0]
    assert (np.abs(ref[rows] - 5) < 1e-3).sum() >= 48
    mse, flag = m.score(x)
    _check(mse, flag, ref)
    # counters
    dev = torch.device("cuda")
    cnt = torch.zeros(4, dtype=torch.int64, device=dev)
    xd = torch.from_numpy(x).to(dev)
    m.score_device(xd, 5.0, counters=cnt)
    c = cnt.cpu().numpy()
    assert c[0] == x.shape[0] and c[1] == int(flag.sum())


def test_score18_with_fused_normalisation(cuda_device):
    m = _model18()
    kind, lo, rng = bs.schemas.norm_arrays()
    m.set_normalization(kind, lo, rng)
    x = bs.synth.sensor_rows(100_001, seed=11)
    mse, flag = m.score(x)
    w = m.get_flat_weights()
    ref, _, _, xn = orc.score(x, w, [18, 14, 7, 7, 18], ACTS, norm=(kind, lo, rng), dtype=np.float64)
    # the normalised features themselves must be BIT-exact with the float32 reference order of operations
    xn32 = orc.normalize(x, kind, lo, rng, dtype=np.float32)
    ref32in = orc.score(xn32.astype(np.float64), w, [18, 14, 7, 7, 18], ACTS, dtype=np.float64)[0]
    _check(mse, flag, ref32in)
    assert flag.sum() > 0


def test_score18_zero_set_instantiation_is_bit_identical(cuda_device):
    """b2_ae_score with the reference's table runs the kernel that leaves out the four features normalize_fn forces to 0.0
    (cardata-v1.py:72-87); context option zero_skip = 0 runs the generic one.  Same bits, also with NaN / inf in those columns,
    and the generic kernel is taken for a table with another zero set."""
    m = _model18()
    kind, lo, rng = bs.schemas.norm_arrays()
    m.set_normalization(kind, lo, rng)
    x = bs.synth.sensor_rows(70_003, seed=12)
    x[::7, 0] = np.nan
    x[::11, 2] = np.inf
    x[::13, 4] = -np.inf
    x[::17, 5] = 3e38
    res = {}
    try:
        for zs in (1, 0):
            bs.keras.set_option("zero_skip", zs)
            mse, flag = m.score(x)
            res[zs] = (np.asarray(mse, dtype=np.float32).view(np.uint32).copy(), np.asarray(flag).copy())
    finally:
        bs.keras.set_option("zero_skip", 1)
    np.testing.assert_array_equal(res[1][0], res[0][0])
    np.testing.assert_array_equal(res[1][1], res[0][1])
    assert np.isfinite(res[1][0].view(np.float32)).all()
    w = m.get_flat_weights()
    xn32 = orc.normalize(np.nan_to_num(x, nan=0.0, posinf=0.0, neginf=0.0), kind, lo, rng, dtype=np.float32)
    ref = orc.score(xn32.astype(np.float64), w, [18, 14, 7, 7, 18], ACTS, dtype=np.float64)[0]
    _check(res[1][0].view(np.float32), res[1][1], ref)


def test_predict_matches_oracle_and_callbacks(cuda_device, ae30, golden_dir):
    m = _model30(ae30, golden_dir)
    x = bs.synth.credit_rows(3000 + 13, seed=5)

    class Out(bs.keras.callbacks.Callback):
        def __init__(self):
            self.rows = []

        def on_predict_batch_end(self, batch, logs=None):
            index = batch * 32
            for outputs in logs["outputs"]:
                for output in outputs:
                    self.rows.append((index, np.array2string(output)))
                    index += 1

    cb = Out()
    y = m.predict(x, callbacks=[cb])
    ref = orc.ae_forward(x, ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float64)
    # reconstruction rows: hold the CUDA result to (a multiple of) the float32 evaluation noise itself — the error of the
    # float32 NumPy restatement against its own float64 shadow on the same rows — with a floor of 2e-6
    noise = np.abs(orc.ae_forward(x, ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float32).astype(np.float64) - ref).max()
    assert np.abs(y - ref).max() <= max(4 * noise, 2e-6), (np.abs(y - ref).max(), noise)
    assert [i for i, _ in cb.rows] == list(range(x.shape[0]))
    # float64 input is cast to float32 on entry like Keras does
    y2 = m.predict(x.astype(np.float64))
    np.testing.assert_array_equal(y, y2)


def test_generic_stack_fallback_kernel(cuda_device):
    k = bs.keras
    inp = k.layers.Input(shape=(12,))
    h = k.layers.Dense(9, activation="relu")(inp)
    h = k.layers.Dense(5, activation="tanh")(h)
    h = k.layers.Dense(12, activation=None)(h)
    m = k.models.Model(inputs=inp, outputs=h, seed=3)
    x = np.random.default_rng(2).standard_normal((5000, 12)).astype(np.float32)
    mse, flag = m.score(x, threshold=1.0)
    ref = orc.score(x, m.get_flat_weights(), [12, 9, 5, 12], [orc.ACT_RELU, orc.ACT_TANH, orc.ACT_LINEAR], dtype=np.float64)[0]
    _check(mse, flag, ref, thr=1.0)


def test_unaligned_input_takes_the_plain_load_path(cuda_device, ae30, golden_dir):
    m = _model30(ae30, golden_dir)
    x = bs.synth.credit_rows(4001, seed=9)
    xd = torch.from_numpy(x).cuda()
    view = xd[1:]  # row offset 1 → 120-byte shift: 8-byte aligned only
    assert view.data_ptr() % 16 == 8
    mse, flag = m.score_device(view)
    ref = orc.score(x[1:], ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float64)[0]
    _check(mse.cpu().numpy(), flag.cpu().numpy(), ref)


def test_bad_arguments_raise(cuda_device, ae30, golden_dir):
    m = _model30(ae30, golden_dir)
    with pytest.raises(ValueError):
        m.score(np.zeros((4, 18), np.float32))
    with pytest.raises(bs._ffi.B2Error):
        m.set_normalization(*bs.schemas.norm_arrays())  # 18-feature table on a 30-wide model
        m.score(np.zeros((4, 30), np.float32))


def test_full_size_properties(cuda_device, ae30, golden_dir):
    """At benchmark scale (2^24 rows here) parity is checked through size-independent properties: the same rows
    scored in one launch, in two halves and after a permutation give bit-identical per-record results, and the flag
    count equals the counter."""
    m = _model30(ae30, golden_dir)
    n = 1 << 24
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(n, 30, device="cuda", generator=g) * 1.3
    mse, flag = m.score_device(x)
    h = n // 2 + 1  # odd split → second half starts unaligned for the bulk path
    m1, f1 = m.score_device(x[:h])
    m2, f2 = m.score_device(x[h:])
    assert torch.equal(mse, torch.cat([m1, m2])) and torch.equal(flag, torch.cat([f1, f2]))
    perm = torch.randperm(n, device="cuda", generator=g)
    mp, fp = m.score_device(x[perm].contiguous())
    assert torch.equal(mp, mse[perm]) and torch.equal(fp, flag[perm])
    sub = torch.arange(0, n, 4099, device="cuda")
    ref = orc.score(x[sub].cpu().numpy(), ae30["weights"], [30, 14, 7, 7, 30], ACTS, dtype=np.float64)[0]
    _check(mse[sub].cpu().numpy(), flag[sub].cpu().numpy(), ref)


def test_score_host_overlapped_path_equals_plain_path(cuda_device, ae30, golden_dir):
    """Model.score(pinned rows, reuse_output=True): chunked H2D / kernel / D2H overlap — same bits as the plain call; sizes around the
    chunk arithmetic; the D = 18 model with its fused normalisation too."""
    m30 = _model30(ae30, golden_dir)
    m18 = _model18()
    m18.set_normalization(*bs.schemas.norm_arrays())
    for m, gen, d in ((m30, bs.synth.credit_rows, 30), (m18, bs.synth.sensor_rows, 18)):
        for n in (4096, 4097, 65_537, 200_003):
            x = np.ascontiguousarray(gen(n, seed=n % 97), dtype=np.float32)
            ref_mse, ref_flag = m.score(x)
            hx = torch.from_numpy(x).pin_memory()
            for _ in range(2):   # second call reuses the buffers
                mse, flag = m.score(hx, reuse_output=True)
                np.testing.assert_array_equal(np.asarray(mse).view(np.uint32), np.asarray(ref_mse, dtype=np.float32).view(np.uint32))
                np.testing.assert_array_equal(flag, ref_flag)
