"""B200-native sensor-record scoring path (decode → normalise → Dense autoencoder → MSE → flag, fit, LSTM forward).

The directory name carries the reference's name and so is not a valid Python identifier; the repo-root module
``b200_sensor`` registers this package under that importable alias::

    import b200_sensor as bs
    model = bs.keras.models.load_model("autoencoder_sensor_anomaly_detection.h5")
    mse, flag = model.score(x)

Compute lives in ``libb2sensor.so`` (csrc/*.cu, sm_100a); see include/b2_api.h for the C ABI.
"""
from . import _ffi, avro_schema, h5lite, schemas  # noqa: F401  (pure-Python, importable without a GPU)

__all__ = ["_ffi", "avro_schema", "h5lite", "schemas", "keras", "decode", "synth", "partition", "sink", "metrics", "kafka", "build_native"]


def build_native(force: bool = False, verbose: bool = False) -> str:
    from .build import build

    return build(force=force, verbose=verbose)


def __getattr__(name):
    # torch-dependent modules are imported lazily so that `import b200_sensor` stays cheap
    import importlib

    if name == "keras":
        mod = importlib.import_module(".keras_compat", __name__)
    elif name in ("decode", "synth", "partition", "sink", "metrics", "kafka", "keras_compat"):
        mod = importlib.import_module("." + name, __name__)
    else:
        raise AttributeError(name)
    globals()[name] = mod
    return mod
