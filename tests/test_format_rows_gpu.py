"""K6 parity: b2_format_rows (np.array2string per row on the GPU) against NumPy, through the C ABI."""
import json
import os

import numpy as np
import pytest

import b200_sensor as bs
from oracle import cpu_oracle as orc

pytestmark = pytest.mark.gpu
NPY = tuple(int(v) for v in np.__version__.split(".")[:2])


def test_golden_rows(cuda_device, golden_dir):
    doc = json.load(open(os.path.join(golden_dir, "array2string_rows.json")))
    by_width = {}
    for r in doc["rows"]:
        by_width.setdefault(len(r["bits"]), []).append(r)
    for d, rs in by_width.items():
        x = np.array([r["bits"] for r in rs], dtype=np.uint32).view(np.float32)
        for compat, key in ((0, "np23"), (1, "np121"), (2, "np114")):
            got = bs.sink.format_rows(x, numpy_compat=compat).messages()
            assert got == [r[key] for r in rs], f"width {d} compat {compat}"


@pytest.mark.parametrize("d", [1, 7, 18, 30, 64])
def test_random_rows_match_numpy(cuda_device, d):
    compat = bs.sink.installed_numpy_compat()
    g = np.random.default_rng(d)
    parts = [np.maximum(g.normal(0.3, 0.5, (3000, d)), 0), g.uniform(-1, 1, (1000, d)) * 10.0 ** g.integers(-6, 9, (1000, d)),
             np.zeros((1000, d))]
    x = np.concatenate(parts).astype(np.float32)
    x[4000:] = g.integers(0, 2 ** 32, (1000, d), dtype=np.uint64).astype(np.uint32).view(np.float32)  # exact bit patterns (NaN payloads too)
    fr = bs.sink.format_rows(x)
    assert fr.messages() == orc.array2string_rows(x, compat)
    # arena bookkeeping: the texts tile the used bytes exactly once
    o, n = fr.offsets.cpu().numpy(), fr.lengths.cpu().numpy()
    order = np.argsort(o)
    assert o[order][0] == 0 and np.all(o[order][1:] == (o[order] + n[order])[:-1]) and int((o + n).max()) == fr.text.shape[0]


def test_large_batch_sampled(cuda_device):
    compat = bs.sink.installed_numpy_compat()
    g = np.random.default_rng(99)
    x = np.maximum(g.normal(0.2, 0.6, (1 << 18, 18)), 0).astype(np.float32)
    msgs = bs.sink.format_rows(x).messages()
    idx = g.integers(0, x.shape[0], 3000)
    ref = orc.array2string_rows(x[idx], compat)
    assert [msgs[i] for i in idx] == ref


def test_output_callback_through_predict(cuda_device, golden_dir):
    """The reference's predict call with its OutputCallback (cardata-v1.py:228-233): messages keyed by the running row index."""
    m = bs.keras.models.load_model(os.path.join(golden_dir, "ae30_5epochs.h5"))
    g = np.random.default_rng(3)
    x = g.normal(0, 1, (1000, 30)).astype(np.float32)
    cb = bs.sink.OutputCallback(32, "anomaly-topic", "localhost:9092")
    y = m.predict(x, batch_size=32, callbacks=[cb])
    cb.flush()
    assert cb.sent == orc.array2string_rows(y, bs.sink.installed_numpy_compat())
    # the plain Keras callback protocol (one call per batch) gives the same messages
    cb2 = bs.sink.OutputCallback(32)
    for b in range(0, 1000, 32):
        cb2.on_predict_batch_end(b // 32, {"outputs": [y[b:b + 32]]})
    cb2.flush()
    assert cb2.sent == cb.sent


def test_arguments_are_checked(cuda_device):
    with pytest.raises(ValueError):
        bs.sink.format_rows(np.zeros((4, 65), np.float32))
    with pytest.raises(ValueError):
        bs.sink.format_rows(np.zeros((4,), np.float32))
    assert len(bs.sink.format_rows(np.zeros((0, 18), np.float32))) == 0
