"""K4b: training of the reference LSTM stack (LSTM-TensorFlow-IO-Kafka/cardata-v1.py:178, :191-194) — back-propagation through
time + Adam in one launch per epoch — against the float64 restatement in oracle/cpu_oracle.py (whose gradients are checked
against torch autograd in tests/test_oracle_golden.py).  Tolerances as for the autoencoder's fit (tests/test_train_gpu.py): a
single step's weights to 2e-6 absolute (Adam's first steps move every weight by ~lr whatever the gradient's size, so this is a
sign test of every one of the 18 642 gradients), losses to 2e-4 relative, weights after a chain of dependent steps to 1e-4."""
import numpy as np
import pytest
import torch

import b200_sensor as bs
from oracle import cpu_oracle as orc

pytestmark = pytest.mark.gpu


def _model(T, seed=314):
    k = bs.keras
    m = k.models.Sequential(seed=seed)
    m.add(k.layers.LSTM(32, activation="relu", input_shape=(T, 18), return_sequences=True))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=False))
    m.add(k.layers.RepeatVector(T))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=True))
    m.add(k.layers.LSTM(32, activation="relu", return_sequences=True))
    m.add(k.layers.TimeDistributed(k.layers.Dense(18)))
    m.compile(metrics=["accuracy"], loss="mean_squared_error", optimizer="adam")
    return m


def _data(n, T, seed):
    g = np.random.default_rng(seed)
    x = g.uniform(-1, 1, (n, T, 18)).astype(np.float32)
    y = (0.5 * x + 0.1 * g.standard_normal((n, T, 18))).astype(np.float32)
    return x, y


@pytest.mark.parametrize("n,T,batch", [(1, 1, 1), (7, 1, 3), (6, 4, 2), (5, 3, 5), (40, 2, 16)])
def test_one_epoch_matches_float64_bptt(cuda_device, n, T, batch):
    m = _model(T, seed=n + T)
    w0 = m.get_flat_weights().astype(np.float64)
    # make the recurrent path and every gate matter
    w0 = w0 + np.random.default_rng(1).normal(size=w0.shape) * 0.05
    m.set_flat_weights(w0.astype(np.float32))
    w0 = m.get_flat_weights().astype(np.float64)
    x, y = _data(n, T, seed=n)
    h = m.fit(x, y, epochs=1, batch_size=batch, shuffle=False)
    w_ref, _, _, losses, accs, it = orc.lstm_fit(x, y, w0, batch=batch, epochs=1)
    steps = (n + batch - 1) // batch
    # Keras' epoch value is the sample-weighted mean of the per-batch values; the oracle returns the plain mean over steps
    sizes = np.full(steps, batch, dtype=np.float64)
    sizes[-1] = n - batch * (steps - 1)
    assert m._adam["iter"] == it == steps
    got = m.get_flat_weights().astype(np.float64)
    tol = 2e-6 if steps == 1 else 1e-4
    assert np.abs(got - w_ref).max() <= tol, np.abs(got - w_ref).max()
    if steps == 1:
        assert abs(h.history["loss"][0] - losses[0]) <= 2e-4 * losses[0]
        assert abs(h.history["accuracy"][0] - accs[0]) < 1e-6


def test_reference_usage_windows_batch1_lookback1_five_epochs(cuda_device):
    """The script's own loop: windows of look_back 1 over the normalised row stream, targets = the next row, batch_size 1,
    epochs=5 (take(200) here instead of take(1000) to keep the float64 oracle quick); the loss must go down, the trained model
    must predict with the new weights, and training resumes from where it stopped."""
    rows = np.random.default_rng(5).uniform(-1, 1, (201, 18)).astype(np.float32)
    rows[:, [0, 2, 4, 5]] = 0.0  # the four features normalize_fn forces to zero
    m = _model(1, seed=7)
    w0 = m.get_flat_weights().astype(np.float64)
    win = bs.keras.Windows(rows, 1, 1)
    h = m.fit(win, epochs=5, batch_size=1, shuffle=False, verbose=0)
    x = win.numpy()[:200]
    y = win.targets()[:200]
    w_ref, m_ref, v_ref, losses, accs, it = orc.lstm_fit(x, y, w0, batch=1, epochs=5)
    assert it == m._adam["iter"] == 1000
    np.testing.assert_allclose(h.history["loss"], losses, rtol=2e-3)
    assert h.history["loss"][-1] < h.history["loss"][0]
    got = m.get_flat_weights().astype(np.float64)
    assert np.abs(got - w_ref).max() <= 5e-3 and np.abs(got - w_ref).mean() <= 2e-4   # 1000 dependent float32 steps
    # predict runs on the trained weights (forward kernels repack them)
    yp = m.predict(x[:50])
    ref = orc.lstm_forward(x[:50], got, dtype=np.float64)
    assert np.abs(yp - ref).max() < 1e-4
    # resume: one more epoch continues the Adam state (iteration count, slots)
    h2 = m.fit(win, epochs=1, batch_size=1, shuffle=False)
    w_ref2, _, _, l2, _, it2 = orc.lstm_fit(x, y, w_ref, batch=1, epochs=1, adam=(m_ref, v_ref), iter0=it)
    assert m._adam["iter"] == it2 == 1200 and abs(h2.history["loss"][0] - l2[0]) <= 5e-3 * l2[0]


def test_fit_argument_checks(cuda_device):
    m = _model(2)
    x, y = _data(4, 2, 0)
    with pytest.raises(ValueError):
        m.fit(x, None)
    with pytest.raises(ValueError):
        m.fit(x, y[:, :1])
    with pytest.raises(NotImplementedError):
        m.fit(bs.keras.Windows(x.reshape(-1, 18), 2, 1))
    h = m.fit(x, y, epochs=2, batch_size=2, shuffle=True, seed=3)
    assert len(h.history["loss"]) == 2
