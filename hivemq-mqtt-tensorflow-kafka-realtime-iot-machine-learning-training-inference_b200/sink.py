"""Result sink: the reference's ``OutputCallback`` / ``KafkaOutputSequence`` pair with the row formatting on the GPU.

Reference (cardata-v1.py:212-226, same in LSTM-…/cardata-v1.py:201-216)::

    class OutputCallback(tf.keras.callbacks.Callback):
        def __init__(self, batch_size, topic, servers):
            self._sequence = kafka_io.KafkaOutputSequence(topic=topic, servers=servers)
            self._batch_size = batch_size
        def on_predict_batch_end(self, batch, logs=None):
            index = batch * self._batch_size
            for outputs in logs['outputs']:
                for output in outputs:
                    self._sequence.setitem(index, np.array2string(output))
                    index += 1
        def flush(self):
            self._sequence.flush()

``format_rows`` does the ``np.array2string(output)`` of every row in one kernel launch (csrc/format_rows.cu); the texts come
back as a byte arena + (offset, length) per row.  Kafka itself is replaced offline by an in-memory sequence (BASELINE.json).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np
import torch

from . import _ffi
from .keras_compat import Callback, _stream_ptr, _to_device_f32, get_ctx

NUMPY_COMPAT = {"2.3": 0, "1.21": 1, "1.14": 2}


def installed_numpy_compat() -> int:
    """The ``numpy_compat`` level whose text equals ``np.array2string`` of the NumPy imported in this process."""
    major, minor = (int(v) for v in np.__version__.split(".")[:2])
    if (major, minor) >= (2, 3):
        return 0
    if (major, minor) >= (1, 21):
        return 1
    return 2


class FormattedRows:
    """Texts of N rows: ``text`` uint8[bytes_used] (device), ``offsets`` int64[N], ``lengths`` int32[N] (device).
    Row i is ``text[offsets[i] : offsets[i] + lengths[i]]`` (rows are packed tile by tile, not in row order)."""

    def __init__(self, text: torch.Tensor, offsets: torch.Tensor, lengths: torch.Tensor):
        self.text, self.offsets, self.lengths = text, offsets, lengths

    def __len__(self):
        return int(self.offsets.shape[0])

    def messages(self) -> List[str]:
        t = self.text.cpu().numpy().tobytes()
        o = self.offsets.cpu().numpy()
        n = self.lengths.cpu().numpy()
        return [t[a:a + b].decode("ascii") for a, b in zip(o.tolist(), n.tolist())]


def format_rows(x, numpy_compat: Optional[int] = None, device: Optional[int] = None) -> FormattedRows:
    """``[np.array2string(row) for row in x]`` for float32 ``x[N, D]`` (D ≤ 64), formatted on the GPU.
    ``numpy_compat``: 0 NumPy ≥ 2.3, 1 NumPy 1.21–2.2, 2 NumPy 1.14–1.20; default: the installed NumPy's generation."""
    compat = installed_numpy_compat() if numpy_compat is None else int(numpy_compat)
    if isinstance(x, torch.Tensor) and x.is_cuda:
        dev = x.device
    else:
        dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
    xd = _to_device_f32(x, dev)
    if xd.dim() != 2:
        raise ValueError(f"expected rows of shape (N, D), got {tuple(xd.shape)}")
    n, d = int(xd.shape[0]), int(xd.shape[1])
    lib = _ffi.lib()
    bound = int(lib.b2_format_rows_bound(d))
    if bound == 0:
        raise ValueError(f"row width {d} not supported (1..64)")
    offs = torch.empty(n, dtype=torch.int64, device=dev)
    lens = torch.empty(n, dtype=torch.int32, device=dev)
    cursor = torch.zeros(2, dtype=torch.int64, device=dev)
    cap = max(n * bound, 1)
    text = torch.empty(cap, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _ffi.check(lib.b2_format_rows(get_ctx(dev.index), _ffi.ptr(xd), n, d, compat, _ffi.ptr(text), cap, _ffi.ptr(offs),
                                      _ffi.ptr(lens), _ffi.ptr(cursor), _stream_ptr()), "b2_format_rows")
    used, dropped = (int(v) for v in cursor.cpu().tolist())
    if dropped:
        raise RuntimeError(f"b2_format_rows dropped {dropped} rows: text arena too small")
    return FormattedRows(text[:used], offs, lens)


class KafkaOutputSequence:
    """In-memory stand-in for ``kafka_io.KafkaOutputSequence`` (cardata-v1.py:215): ``setitem(index, message)`` buffers,
    ``flush()`` moves the buffered messages, ordered by index, to ``self.sent`` (what a producer would have written)."""

    def __init__(self, topic: str = "", servers: str = "", configuration=None):
        self.topic, self.servers, self.configuration = topic, servers, configuration
        self._pending: Dict[int, str] = {}
        self.sent: List[str] = []

    def setitem(self, index: int, message: str):
        self._pending[int(index)] = message

    def flush(self):
        for k in sorted(self._pending):
            self.sent.append(self._pending[k])
        self._pending.clear()


class OutputCallback(Callback):
    """The reference's callback; the per-row ``np.array2string`` runs as one GPU launch per predict batch, or — when
    ``Model.predict`` sees this class — as ONE launch over all rows (``consume_rows``)."""

    def __init__(self, batch_size, topic="", servers="", numpy_compat: Optional[int] = None):
        self._sequence = KafkaOutputSequence(topic=topic, servers=servers)
        self._batch_size = batch_size
        self._compat = numpy_compat

    def on_predict_batch_end(self, batch, logs=None):
        index = batch * self._batch_size
        for outputs in (logs or {}).get("outputs", []):
            rows = np.asarray(outputs, dtype=np.float32)
            if rows.ndim != 2:  # the LSTM script's outputs are [T, F] blocks per element: np.array2string of a 2-D array
                raise NotImplementedError("OutputCallback formats 1-D rows ([B, D] outputs)")
            rows2 = rows
            for msg in format_rows(rows2, self._compat).messages():
                self._sequence.setitem(index, msg)
                index += 1

    def consume_rows(self, rows_dev: torch.Tensor):
        """All rows of a predict call at once (device tensor [N, D]); message index = row index, as in the batch loop."""
        for i, msg in enumerate(format_rows(rows_dev, self._compat).messages()):
            self._sequence.setitem(i, msg)

    def flush(self):
        self._sequence.flush()

    @property
    def sent(self) -> List[str]:
        return self._sequence.sent
