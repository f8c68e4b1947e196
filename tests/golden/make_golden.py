"""Regenerate the fixtures under tests/golden/ from the reference checkout (run in the authoring container only).

    python tests/golden/make_golden.py

Inputs (read-only): /root/reference/models/*.h5, /root/reference/testdata/car-sensor-data.csv.
Outputs:
  ae30_5epochs.npz / ae30_100epochs.npz   flat weights (Keras order), Adam m/v/iter, model_config + training_config JSON
  ae30_5epochs.h5                          the same model re-written by OUR h5lite writer (load_model fixture on the GPU box)
  car_sensor_sample.csv                    header + first 512 data lines of testdata/car-sensor-data.csv
  csv_tokens.npz                           all unique numeric tokens of that CSV with their glibc-strtof float32 bit patterns
  appg_vectors.json                        SURVEY.md App. G known-answer vectors recomputed with oracle/cpu_oracle.py (float32)
  cardata_v1_sample.json                   the 10 simulator messages of python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.json
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import b200_sensor as bs  # noqa: E402
from oracle import cpu_oracle as orc  # noqa: E402

REF = "/root/reference"
MODELS = {"ae30_5epochs": "autoencoder_sensor_anomaly_detection.h5",
          "ae30_100epochs": "autoencoder_sensor_anomaly_detection_fully_trained_100_epochs.h5"}


def extract(h5path):
    f = bs.h5lite.H5File(h5path)
    mw = f["model_weights"]
    parts, inner = [], []
    for ln in [str(s) for s in np.atleast_1d(mw.attrs["layer_names"])]:
        g = mw[ln]
        wn = [str(s) for s in np.atleast_1d(g.attrs.get("weight_names", []))]
        if not wn:
            continue
        inner.append(wn[0].split("/")[0])
        by = {("kernel" if "kernel" in w else "bias"): g[w].read() for w in wn}
        parts += [by["kernel"].ravel(), by["bias"].ravel()]
    flat = np.concatenate(parts).astype(np.float32)
    ow = f["optimizer_weights"]
    found = {}
    ow.visit(lambda p, o: found.__setitem__(p, o.read()) if not o.is_group else None)
    it = [v for k, v in found.items() if k.endswith("iter:0")][0]
    m, v = [], []
    for nm in inner:
        for kind in ("kernel", "bias"):
            m.append([val for k, val in found.items() if k.endswith(f"Adam/{nm}/{kind}/m:0")][0].ravel())
            v.append([val for k, val in found.items() if k.endswith(f"Adam/{nm}/{kind}/v:0")][0].ravel())
    return dict(weights=flat, adam_m=np.concatenate(m).astype(np.float32), adam_v=np.concatenate(v).astype(np.float32),
                adam_iter=np.int64(np.asarray(it).ravel()[0]), model_config=f.attrs["model_config"],
                training_config=f.attrs["training_config"])


def main():
    with open(os.path.join(REF, 'python-scripts', 'LSTM-TensorFlow-IO-Kafka', 'cardata-v1.json')) as f:
        open(os.path.join(HERE, 'cardata_v1_sample.json'), 'w').write(''.join(f.readlines()[:10]))
    widths, acts = [30, 14, 7, 7, 30], [orc.ACT_TANH, orc.ACT_RELU, orc.ACT_TANH, orc.ACT_RELU]
    appg = {}
    for key, fn in MODELS.items():
        d = extract(os.path.join(REF, "models", fn))
        np.savez(os.path.join(HERE, key + ".npz"), **d)
        vecs = {"zeros": np.zeros((1, 30), np.float32), "ones": np.ones((1, 30), np.float32),
                "ramp": ((np.arange(30, dtype=np.float32) - 15) / 10).reshape(1, 30)}
        appg[key] = {}
        for nm, x in vecs.items():
            m, fl, y, _ = orc.score(x, d["weights"], widths, acts, dtype=np.float32)
            appg[key][nm] = {"mse": float(m[0]), "flag": int(fl[0]), "y": [float(t) for t in y[0]]}
    json.dump(appg, open(os.path.join(HERE, "appg_vectors.json"), "w"), indent=1)

    # same model through OUR writer (needs no GPU: weights are set on the Python object only)
    d = np.load(os.path.join(HERE, "ae30_5epochs.npz"))
    cfg = json.loads(str(d["model_config"]))
    tree = {"model_weights": {"@attrs": {"layer_names": np.array([l["name"] for l in cfg["config"]["layers"]])}}}
    p = 0
    for l in cfg["config"]["layers"]:
        nm = l["name"]
        if l["class_name"] != "Dense":
            tree["model_weights"][nm] = {"@attrs": {"weight_names": np.array([], dtype="S1")}}
            continue
        i = widths[len([k for k in tree["model_weights"] if k.startswith("dense")])]
        o = l["config"]["units"]
        k = d["weights"][p:p + i * o].reshape(i, o)
        b = d["weights"][p + i * o:p + i * o + o]
        p += i * o + o
        tree["model_weights"][nm] = {nm: {"kernel:0": k, "bias:0": b},
                                     "@attrs": {"weight_names": np.array([f"{nm}/kernel:0", f"{nm}/bias:0"])}}
    bs.h5lite.write_h5(os.path.join(HERE, "ae30_5epochs.h5"), tree,
                       {"keras_version": "2.2.4-tf", "backend": "tensorflow", "model_config": str(d["model_config"]),
                        "training_config": str(d["training_config"])})

    lines = open(os.path.join(REF, "testdata", "car-sensor-data.csv")).read().split("\n")
    open(os.path.join(HERE, "car_sensor_sample.csv"), "w").write("\n".join(lines[:513]) + "\n")
    toks = sorted({t for ln in lines[1:] if ln for t in ln.split(",")[2:]})
    bits = np.array([np.float32(orc.strtof(t.encode())[0]).view(np.uint32) for t in toks], dtype=np.uint32)
    np.savez_compressed(os.path.join(HERE, "csv_tokens.npz"), tokens=np.array(toks), bits=bits)
    print("golden fixtures written:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
