"""K4 parity: LSTM stack forward against the CPU oracle (float64 shadow), tolerance 1e-4 (BASELINE.json north_star)."""
import numpy as np
import pytest

import b200_sensor as bs
from oracle import cpu_oracle as orc

pytestmark = pytest.mark.gpu


def _lstm_model(look_back, features=18, seed=314):
    k = bs.keras
    m = k.models.Sequential(seed=seed)
    m.add(k.layers.LSTM(32, activation="relu", input_shape=(look_back, features), return_sequences=True))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=False))
    m.add(k.layers.RepeatVector(look_back))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=True))
    m.add(k.layers.LSTM(32, activation="relu", return_sequences=True))
    m.add(k.layers.TimeDistributed(k.layers.Dense(features)))
    return m


@pytest.mark.parametrize("path", ["tcgen05", "tcgen05x2", "fp32"])
@pytest.mark.parametrize("batch,T", [(1, 1), (200, 1), (127, 4), (129, 16), (5000, 3)])
def test_lstm_forward_matches_oracle(cuda_device, lstm_path, batch, T, path):
    """Both kernels: the default tensor-core one (tcgen05.mma kind::tf32, split precision) and the FP32 FMA one."""
    lstm_path(path)
    m = _lstm_model(T)
    assert m.count_params() == 18642
    w = m.get_flat_weights()
    # make the recurrent path matter: scale up so that gates are not all near 0.5
    g = np.random.default_rng(batch + T)
    w = (w * g.uniform(0.5, 2.0, size=w.shape)).astype(np.float32)
    m.set_flat_weights(w)
    x = g.uniform(-1, 1, size=(batch, T, 18)).astype(np.float32)
    x[:, :, [0, 2, 4, 5]] = 0.0  # the four features normalize_fn zeroes
    y = m.predict(x)
    ref = orc.lstm_forward(x, w, dtype=np.float64)
    assert y.shape == (batch, T, 18)
    np.testing.assert_allclose(y, ref, rtol=1e-4, atol=1e-4)  # BASELINE.json: LSTM outputs within 1e-4
    if path == "fp32":  # and tighter against what float32 itself can do
        assert np.abs(y - ref).max() < 2e-5
    print(f"lstm {path} B={batch} T={T}: max |err| = {np.abs(y - ref).max():.3g}")


def test_lstm_reference_usage_batch1_lookback1_with_callback(cuda_device):
    """The script's shape: look_back = 1, batch_size = 1, predict 200 windows with the Kafka OutputCallback."""
    m = _lstm_model(1)
    x = np.random.default_rng(0).uniform(-1, 1, size=(200, 1, 18)).astype(np.float32)
    got = []

    class Out(bs.keras.callbacks.Callback):
        def on_predict_batch_end(self, batch, logs=None):
            index = batch * 1
            for outputs in logs["outputs"]:
                for output in outputs:
                    got.append((index, np.array2string(output)))
                    index += 1

    y = m.predict(x, batch_size=1, callbacks=[Out()])
    assert [i for i, _ in got] == list(range(200))
    np.testing.assert_allclose(y, orc.lstm_forward(x, m.get_flat_weights(), dtype=np.float64), atol=1e-4)


def test_lstm_rejects_other_stacks(cuda_device):
    k = bs.keras
    m = k.models.Sequential()
    m.add(k.layers.LSTM(8, activation="tanh", input_shape=(1, 18)))
    with pytest.raises(NotImplementedError):
        m.predict(np.zeros((2, 1, 18), np.float32))


@pytest.mark.parametrize("path", ["tcgen05", "tcgen05x2", "fp32"])
@pytest.mark.parametrize("n_rows,look_back,shift", [(1200, 1, 1), (777, 4, 1), (1000, 16, 3), (5, 8, 1)])
def test_lstm_sliding_windows_read_in_place(cuda_device, lstm_path, n_rows, look_back, shift, path):
    """dataset.window(look_back, shift, drop_remainder=True) fused into the forward pass (LSTM-…/cardata-v1.py:184-185): the
    kernel reads window i at rows [i*shift, i*shift+look_back) of the stream; result == predict on the materialised windows."""
    lstm_path(path)
    m = _lstm_model(look_back)
    g = np.random.default_rng(n_rows)
    rows = g.uniform(-1, 1, size=(n_rows, 18)).astype(np.float32)
    win = bs.keras.Windows(rows, look_back, shift)
    n_win = 0 if n_rows < look_back else (n_rows - look_back) // shift + 1
    assert len(win) == n_win
    y = m.predict(win)
    assert y.shape == (n_win, look_back, 18)
    if n_win == 0:
        return
    xw = win.numpy()
    assert xw.shape == (n_win, look_back, 18)
    np.testing.assert_array_equal(xw[-1], rows[(n_win - 1) * shift:(n_win - 1) * shift + look_back])
    np.testing.assert_array_equal(y, m.predict(xw))  # same kernel, same arithmetic: bit-identical
    np.testing.assert_allclose(y, orc.lstm_forward(xw, m.get_flat_weights(), dtype=np.float64), rtol=1e-4, atol=1e-4)
    # the script's (x, y) pairing: y_i = row i + look_back, zip truncates to the shorter side (:189-191)
    t = win.targets()
    assert t.shape[0] == min(n_rows - look_back, n_win) and t.shape[1:] == (1, 18)
    if shift == 1 and t.shape[0]:
        np.testing.assert_array_equal(t[0, 0], rows[look_back])


def test_lstm_two_tiles_per_cta_equals_one_tile_on_a_large_batch(cuda_device, lstm_path):
    """More tiles than SMs (the default then runs two tiles per CTA, taking turns at one TMEM accumulator): bit-identical to the
    one-tile kernel — the arithmetic per sequence is the same — including an odd tile count and a ragged last tile."""
    m = _lstm_model(2)
    g = np.random.default_rng(9)
    x = g.uniform(-1, 1, size=(148 * 128 * 2 + 128 * 3 + 57, 2, 18)).astype(np.float32)
    idx = np.concatenate([np.arange(0, x.shape[0], 997), np.arange(x.shape[0] - 700, x.shape[0])])  # incl. the last round's tiles
    ref = orc.lstm_forward(x[idx], m.get_flat_weights(), dtype=np.float64)
    lstm_path("tcgen05")
    y1 = m.predict(x)
    assert np.abs(y1[idx] - ref).max() < 1e-4, "one tile per CTA"
    bs.keras.set_option("lstm_tc", 2)
    y2 = m.predict(x)
    assert np.abs(y2[idx] - ref).max() < 1e-4, "two tiles per CTA"
    np.testing.assert_array_equal(y1, y2)
