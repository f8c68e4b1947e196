"""K3 parity: fused forward/backward/Adam against the float64 CPU oracle (oracle/cpu_oracle.py, itself checked against
torch autograd), through the C ABI.  Keras' reduction order cannot be reproduced, so parity is by tolerance:
per-step loss within 2e-4 relative, weights within 5e-5 absolute after tens of dependent Adam steps."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import _ffi, schemas, synth
from oracle import cpu_oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

pytestmark = pytest.mark.gpu
ACTS = [orc.ACT_TANH, orc.ACT_RELU, orc.ACT_TANH, orc.ACT_RELU]
L1 = float(np.float32(1e-7))


def _model(d, seed=314):
    k = bs.keras
    inp = k.layers.Input(shape=(d,))
    e = k.layers.Dense(14, activation="tanh", activity_regularizer=k.regularizers.l1(1e-7))(inp)
    e = k.layers.Dense(7, activation="relu")(e)
    dd = k.layers.Dense(7, activation="tanh")(e)
    dd = k.layers.Dense(d, activation="relu")(dd)
    m = k.models.Model(inputs=inp, outputs=dd, seed=seed)
    m.compile(optimizer="adam", loss="mean_squared_error", metrics=["accuracy"])
    return m


@pytest.mark.parametrize("d,batch,n", [(18, 32, 512), (30, 128, 1000), (30, 100, 777), (18, 1, 40), (30, 256, 600)])
def test_fit_steps_match_oracle(cuda_device, d, batch, n):
    m = _model(d, seed=11)
    w0 = m.get_flat_weights()
    g = np.random.default_rng(d + batch)
    x = g.uniform(-1, 1, size=(n, d)).astype(np.float32)
    hist = m.fit(x, x, epochs=2, batch_size=batch, shuffle=False)
    widths = [d, 14, 7, 7, d]
    w, mm, vv, t, losses, accs = orc.fit(x.astype(np.float64), w0, widths, ACTS, L1, batch=batch, epochs=2, dtype=np.float64)
    steps = (n + batch - 1) // batch
    sizes = np.full(steps, batch, dtype=np.float64)
    sizes[-1] = n - batch * (steps - 1)
    for ep in range(2):
        ref_loss = (losses[ep * steps:(ep + 1) * steps] * sizes).sum() / n
        ref_acc = (accs[ep * steps:(ep + 1) * steps] * sizes).sum() / n
        assert abs(hist.history["loss"][ep] - ref_loss) <= 2e-4 * ref_loss
        assert abs(hist.history["accuracy"][ep] - ref_acc) <= 0.02
    np.testing.assert_allclose(m.get_flat_weights(), w, rtol=0, atol=5e-5)
    assert m._adam["iter"] == t == 2 * steps


def test_fit_with_order_and_normalisation(cuda_device):
    m = _model(18, seed=5)
    kind, lo, rng = schemas.norm_arrays()
    m.set_normalization(kind, lo, rng)
    w0 = m.get_flat_weights()
    x = synth.sensor_rows(640, seed=3, outlier_frac=0.0)
    m.fit(x, epochs=1, batch_size=64, shuffle=True, seed=123)
    perm = np.random.Generator(np.random.PCG64(123)).permutation(640)
    xn = orc.normalize(x, kind, lo, rng, dtype=np.float32).astype(np.float64)
    w, *_ = orc.fit(xn, w0, [18, 14, 7, 7, 18], ACTS, L1, batch=64, epochs=1, order=[perm], dtype=np.float64)
    np.testing.assert_allclose(m.get_flat_weights(), w, rtol=0, atol=5e-5)


def test_large_batch_grad_apply_matches_oracle(cuda_device, ae30):
    l = _ffi.lib()
    m = bs.keras.models.load_model(os.path.join(os.path.dirname(__file__), "golden", "ae30_5epochs.h5"))
    m._h5_optimizer = None
    x = synth.credit_rows(70_001, seed=4)
    xd = torch.from_numpy(x).cuda()
    h = m._handle()
    P = m.n_params
    nbytes = int(l.b2_ae_grad_scratch_bytes(h, x.shape[0]))
    assert nbytes == 4 * min(148, (x.shape[0] + 255) // 256) * (P + 3) or nbytes > 0
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    grads = torch.empty(P + 3, dtype=torch.float32, device="cuda")
    nm = _ffi.make_norm()
    _ffi.check(l.b2_ae_grad(h, xd.data_ptr(), x.shape[0], C.byref(nm), grads.data_ptr(), scratch.data_ptr(), nbytes, None))
    gh = grads.cpu().numpy().astype(np.float64)
    loss, gref, correct = orc.ae_loss_and_grads(x.astype(np.float64), ae30["weights"], [30, 14, 7, 7, 30], ACTS, L1, dtype=np.float64)
    n = x.shape[0]
    assert gh[P + 2] == n and gh[P + 1] == correct
    assert abs(gh[P] / n - loss) <= 1e-5 * loss
    np.testing.assert_allclose(gh[:P] / n, gref, rtol=2e-4, atol=2e-7)
    # scratch too small is refused, not overrun
    rc = l.b2_ae_grad(h, xd.data_ptr(), x.shape[0], C.byref(nm), grads.data_ptr(), scratch.data_ptr(), 16, None)
    assert rc == -1
    # one Adam step from the reduced gradient
    adam = _ffi.Adam(1e-3, 0.9, 0.999, 1e-7, 0)
    stats = torch.empty(2, dtype=torch.float32, device="cuda")
    _ffi.check(l.b2_ae_apply(h, grads.data_ptr(), C.byref(adam), stats.data_ptr(), None))
    w1, *_ = orc.adam_step(ae30["weights"].astype(np.float64), np.zeros(P), np.zeros(P), gref, 1)
    np.testing.assert_allclose(m.get_flat_weights(), w1, rtol=0, atol=2e-6)
    assert adam.iter == 1 and abs(float(stats[0]) - loss) <= 1e-5 * loss


def test_keras_style_fit_large_batch_and_history(cuda_device):
    m = _model(18, seed=2)
    x = np.random.default_rng(0).uniform(-1, 1, size=(20_000, 18)).astype(np.float32)
    h = m.fit(x, epochs=6, batch_size=4096, shuffle=True, validation_data=(x[:2000], x[:2000]))
    assert len(h.history["loss"]) == 6 and h.history["loss"][-1] < h.history["loss"][0]
    assert "val_loss" in h.history and "accuracy" in h.history


def test_save_load_roundtrip_with_adam_slots_and_checkpoint(cuda_device, tmp_path):
    m = _model(30, seed=9)
    x = synth.credit_rows(2048, seed=1, outlier_frac=0.0)
    cp = bs.keras.callbacks.ModelCheckpoint(str(tmp_path / "best.h5"), save_best_only=True)
    m.fit(x, epochs=3, batch_size=128, shuffle=False, validation_data=(x[:256],), callbacks=[cp])
    p = str(tmp_path / "m.h5")
    m.save(p)
    m2 = bs.keras.models.load_model(p)
    np.testing.assert_array_equal(m.get_flat_weights(), m2.get_flat_weights())
    assert m2._adam["iter"] == m._adam["iter"] == 48
    np.testing.assert_array_equal(m.predict(x[:100]), m2.predict(x[:100]))
    # resuming from the file continues exactly like continuing in memory (Adam slots and iter restored)
    m.fit(x, epochs=1, batch_size=128, shuffle=False)
    m2.fit(x, epochs=1, batch_size=128, shuffle=False)
    np.testing.assert_array_equal(m.get_flat_weights(), m2.get_flat_weights())
    assert os.path.exists(tmp_path / "best.h5")
    f = bs.h5lite.H5File(p)
    assert "optimizer_weights" in f and f.attrs["keras_version"] == "2.2.4-tf"
    # tf.keras' loader walks optimizer_weights.attrs['weight_names']: the reference's own file lists iter, then every slot m
    # (kernel, bias per layer), then every slot v (tests/test_h5lite.py checks that against models/*.h5); same here, with this
    # model's layer names, and every name resolves to a dataset
    wn = [str(s) for s in np.atleast_1d(f["optimizer_weights"].attrs["weight_names"])]
    ours = [l.name for l in m.layers if isinstance(l, bs.keras_compat.Dense)]
    want = ["training/Adam/iter:0"] + [f"training/Adam/{n}/{k}/{slot}:0" for slot in ("m", "v") for n in ours for k in ("kernel", "bias")]
    assert wn == want and len(wn) == 17
    for a in wn:
        assert f["optimizer_weights/" + a].read() is not None


def test_fit_rejects_what_the_reference_never_does(cuda_device):
    m = _model(18)
    with pytest.raises(NotImplementedError):
        m.fit(np.zeros((8, 18), np.float32), np.ones((8, 18), np.float32))
    k = bs.keras
    m2 = k.models.Model(inputs=(i := k.layers.Input(shape=(18,))), outputs=k.layers.Dense(18)(i))
    with pytest.raises(RuntimeError):
        m2.fit(np.zeros((8, 18), np.float32))
    with pytest.raises(NotImplementedError):
        m2.compile(optimizer="sgd")
