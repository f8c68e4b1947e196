"""HDF5 reader/writer: anchors from SURVEY.md App. A (real files, only when the reference checkout is present) and
round trips through our own writer (always)."""
import json
import os

import numpy as np
import pytest

import b200_sensor as bs
from b200_sensor import h5lite


def test_golden_h5_written_by_our_writer_loads(golden_dir, ae30):
    f = h5lite.H5File(os.path.join(golden_dir, "ae30_5epochs.h5"))
    assert f.attrs["keras_version"] == "2.2.4-tf" and f.attrs["backend"] == "tensorflow"
    cfg = json.loads(f.attrs["model_config"])
    assert [l["class_name"] for l in cfg["config"]["layers"]] == ["InputLayer"] + ["Dense"] * 4
    k = f["model_weights/dense/dense/kernel:0"].read()
    assert k.shape == (30, 14) and k.dtype == np.float32
    np.testing.assert_array_equal(k.ravel(), ae30["weights"][:420])


def test_writer_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    tree = {"g1": {"a": rng.standard_normal((3, 5)).astype(np.float32), "b": np.arange(7, dtype=np.int64),
                   "@attrs": {"names": np.array(["x/y:0", "longer_name:0"])}},
            "g2": {"deep": {"deeper": {"s": np.array(42, dtype=np.int64)}}},
            **{f"many{i:02d}": {"v": np.full(2, i, dtype=np.float32)} for i in range(20)}}
    p = str(tmp_path / "t.h5")
    h5lite.write_h5(p, tree, {"keras_version": "2.2.4-tf", "model_config": json.dumps({"k": "v" * 5000})})
    f = h5lite.H5File(p)
    np.testing.assert_array_equal(f["g1/a"].read(), tree["g1"]["a"])
    np.testing.assert_array_equal(f["g1/b"].read(), tree["g1"]["b"])
    assert list(f["g1"].attrs["names"]) == ["x/y:0", "longer_name:0"]
    assert int(f["g2/deep/deeper/s"].read()) == 42 and f["g2/deep/deeper/s"].read().shape == ()
    assert len(json.loads(f.attrs["model_config"])["k"]) == 5000
    assert sorted(f.keys()) == sorted(tree.keys())
    for i in range(20):
        np.testing.assert_array_equal(f[f"many{i:02d}/v"].read(), np.full(2, i, np.float32))


def test_bad_files(tmp_path):
    p = tmp_path / "bad.h5"
    p.write_bytes(b"not hdf5 at all")
    with pytest.raises(h5lite.H5Error):
        h5lite.H5File(str(p))


@pytest.mark.parametrize("fn,first,iter_", [
    ("autoencoder_sensor_anomaly_detection.h5", [-0.0519325, 0.03871651, -0.8109186, 0.0207838], 35545),
    ("autoencoder_sensor_anomaly_detection_fully_trained_100_epochs.h5", [0.01263126, 0.07147288, -0.02311895, -0.00401752], 167132),
])
def test_reference_files_match_survey_anchors(have_reference, fn, first, iter_):
    if not have_reference:
        pytest.skip("reference checkout not present on this box")
    f = h5lite.H5File("/root/reference/models/" + fn)
    mw = f["model_weights"]
    names = [str(s) for s in mw.attrs["layer_names"]]
    g = mw[names[1]]
    wn = [str(s) for s in g.attrs["weight_names"]]
    k = g[wn[0]].read()
    assert k.shape == (30, 14)
    np.testing.assert_allclose(k[0, :4], first, rtol=0, atol=5e-8)  # survey prints 7-8 digits
    found = {}
    f["optimizer_weights"].visit(lambda p, o: found.__setitem__(p, o) if not o.is_group else None)
    it = [o for p, o in found.items() if p.endswith("iter:0")][0]
    assert int(it.read()) == iter_
    tc = json.loads(f.attrs["training_config"])
    assert tc["optimizer_config"]["class_name"] == "Adam" and tc["loss"] == "mean_squared_error"
    # the list Keras' loader walks: iter, then slot m of (kernel, bias) per Dense layer, then slot v — what Model.save() writes too
    wn = [str(s) for s in np.atleast_1d(f["optimizer_weights"].attrs["weight_names"])]
    prefix = wn[0].rsplit("/", 1)[0]
    dense = [w.split("/")[2] for w in wn[1:9:2]]  # the optimizer's own names of the four Dense layers (e.g. dense_4_1 in the second file)
    assert len(set(dense)) == 4
    assert wn == [prefix + "/iter:0"] + [f"{prefix}/{n}/{k}/{slot}:0" for slot in ("m", "v") for n in dense for k in ("kernel", "bias")]


def test_reference_offsets(have_reference):
    if not have_reference:
        pytest.skip("reference checkout not present on this box")
    f = h5lite.H5File("/root/reference/models/autoencoder_sensor_anomaly_detection.h5")
    assert f["model_weights/dense/dense/kernel:0"].file_offset == 10216
    assert f["model_weights/dense_3/dense_3/bias:0"].file_offset == 16248


def test_golden_npz_matches_reference(have_reference, ae30):
    if not have_reference:
        pytest.skip("reference checkout not present on this box")
    f = h5lite.H5File("/root/reference/models/autoencoder_sensor_anomaly_detection.h5")
    np.testing.assert_array_equal(f["model_weights/dense/dense/kernel:0"].read().ravel(), ae30["weights"][:420])
    np.testing.assert_array_equal(f["model_weights/dense_3/dense_3/bias:0"].read(), ae30["weights"][-30:])
