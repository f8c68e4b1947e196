"""The drop-in copy of the reference's cardata-v1.py flow runs end to end on the GPU path."""
import importlib.util
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cardata_v1_flow(cuda_device, tmp_path):
    spec = importlib.util.spec_from_file_location("cardata_v1_b200", os.path.join(ROOT, "scripts", "cardata_v1_b200.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    history, predict, seq, mse, flag = mod.main(32 * 200, str(tmp_path / "m.h5"), verbose=0)
    assert len(history["loss"]) == 5 and history["loss"][-1] < history["loss"][0]
    assert predict.shape == (3200, 18) and sorted(seq) == list(range(3200))
    assert seq[0].startswith("[") and np.isfinite(mse).all() and flag.dtype == np.uint8


def test_lstm_cardata_v1_flow(cuda_device):
    spec = importlib.util.spec_from_file_location("lstm_cardata_v1_b200", os.path.join(ROOT, "scripts", "lstm_cardata_v1_b200.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    history, predict, seq = mod.main(700, take=400, predict_take=200, verbose=0)
    assert len(history["loss"]) == 5 and history["loss"][-1] < history["loss"][0] and "accuracy" in history
    assert predict.shape == (200, 1, 18) and sorted(seq) == list(range(200)) and seq[0].startswith("[")
