"""Importable alias for the package directory (whose name contains hyphens): ``import b200_sensor``."""
import importlib.util as _u
import os as _os
import sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)),
                     "hivemq-mqtt-tensorflow-kafka-realtime-iot-machine-learning-training-inference_b200")
_spec = _u.spec_from_file_location("b200_sensor", _os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = _u.module_from_spec(_spec)
_sys.modules["b200_sensor"] = _mod
_spec.loader.exec_module(_mod)
