"""Kafka record-batch (message format v2) framing: broker bytes → the ``RecordBatch`` of message values the decoders take.

In the reference this is tensorflow-io's ``KafkaDataset`` / librdkafka (cardata-v1.py:10).  ``scan_batches`` walks the batch
chain on the host (12 bytes per batch; ``b2_kafka_scan_batches``), ``index_segment`` lets the GPU inflate compressed batches
(gzip / Snappy / LZ4; ``b2_kafka_inflate_sizes`` + ``b2_kafka_inflate``), check every batch's CRC-32C and walk its records
(``b2_kafka_index_records``, one warp per batch) and returns the values as ``RecordSpans`` — byte ranges inside the segment, which
``decode.decode_score`` consumes where they lie.  ``values_from_segment`` additionally packs the values back to back
(``b2_kafka_gather_values``) into a ``RecordBatch`` for the other decoders.  ``write_segment`` / ``write_segment_uniform`` are the
inverse, host side, for tests and synthetic input (Kafka itself is replaced offline by in-memory bytes, BASELINE.json).
"""
from __future__ import annotations

import struct
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _ffi
from .decode import RecordBatch, RecordSpans
from .keras_compat import _stream_ptr, get_ctx

FRAME_STATUS = {0: "ok", 1: "not message format v2", 2: "CRC-32C mismatch", 3: "truncated", 4: "compressed batch", 5: "control batch"}
_HEADER = 61


class FrameError(ValueError):
    pass


def _crc32c_table() -> np.ndarray:
    t = np.arange(256, dtype=np.uint32)
    for _ in range(8):
        t = np.where(t & 1, (t >> 1) ^ np.uint32(0x82F63B78), t >> 1).astype(np.uint32)
    return t


_TAB = _crc32c_table()


def crc32c(data: bytes) -> int:
    """CRC-32C (Castagnoli), the checksum of Kafka record batches.  crc32c(b"123456789") == 0xE3069283 (RFC 3720 B.4)."""
    c = 0xFFFFFFFF
    tab = _TAB.tolist()
    for b in data:
        c = tab[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def _varint(v: int) -> bytes:
    u = (v << 1) ^ (v >> 63)
    out = bytearray()
    while u >= 0x80:
        out.append((u & 0x7F) | 0x80)
        u >>= 7
    out.append(u)
    return bytes(out)


_CODECS = {None: 0, "none": 0, "gzip": 1, "snappy": 2, "snappy-xerial": 2, "lz4": 3, "zstd": 4, "zstd-stream": 4}


def _compress(records: bytes, compression) -> bytes:
    """The records section of a batch under ``compression.type`` (test / synthetic input only: gzip from the standard library,
    Snappy, LZ4 frames and Zstandard from pyarrow; "snappy-xerial" adds the Java client's SnappyOutputStream framing)."""
    if compression in (None, "none"):
        return records
    if compression == "gzip":
        import gzip

        return gzip.compress(records)
    import pyarrow as pa

    if compression == "snappy":
        return pa.compress(records, codec="snappy", asbytes=True)
    if compression == "zstd":          # one frame with its content size, as ZSTD_compress writes it
        return pa.compress(records, codec="zstd", asbytes=True)
    if compression == "zstd-stream":   # the streaming API (what the Java client's ZstdOutputStream uses): window descriptor, no content size
        sink = pa.BufferOutputStream()
        with pa.CompressedOutputStream(sink, "zstd") as f:
            for i in range(0, len(records), 40_000):
                f.write(records[i:i + 40_000])
        return sink.getvalue().to_pybytes()
    if compression == "snappy-xerial":
        out = b"\x82SNAPPY\x00" + struct.pack(">ii", 1, 1)
        for i in range(0, len(records), 32768):
            z = pa.compress(records[i:i + 32768], codec="snappy", asbytes=True)
            out += struct.pack(">i", len(z)) + z
        return out
    if compression == "lz4":
        return pa.compress(records, codec="lz4", asbytes=True)
    raise ValueError(f"unknown compression {compression!r}")


def write_segment(values: Sequence[bytes], records_per_batch: int = 500, keys: Optional[Sequence[Optional[bytes]]] = None,
                  base_offset: int = 0, headers: bool = False, compression=None) -> bytes:
    """Message values → log-segment bytes: v2 record batches of ``records_per_batch`` records each, their records sections
    compressed as a whole when ``compression`` is "gzip" / "snappy" / "snappy-xerial" / "lz4" / "zstd" / "zstd-stream" (a list gives one per batch)."""
    out = bytearray()
    n = len(values)
    for b0 in range(0, n, records_per_batch):
        recs = bytearray()
        cnt = min(records_per_batch, n - b0)
        for i in range(cnt):
            v = values[b0 + i]
            k = keys[b0 + i] if keys is not None else None
            body = bytearray(b"\x00") + _varint(i * 3) + _varint(i)          # attributes, timestampDelta, offsetDelta
            body += _varint(-1) if k is None else _varint(len(k)) + k
            body += _varint(-1) if v is None else _varint(len(v)) + v
            if headers and i % 7 == 0:
                body += _varint(1) + _varint(2) + b"hk" + _varint(3) + b"hv!"
            else:
                body += _varint(0)
            recs += _varint(len(body)) + body
        comp = compression[(b0 // records_per_batch) % len(compression)] if isinstance(compression, (list, tuple)) else compression
        after_crc = struct.pack(">hiqqqhii", _CODECS[comp], cnt - 1, 1567606196000, 1567606196000 + 3 * cnt, -1, -1, -1, cnt) + _compress(bytes(recs), comp)
        batch_len = 4 + 1 + 4 + len(after_crc)                               # leader epoch, magic, crc, rest
        out += struct.pack(">qiib", base_offset + b0, batch_len, 0, 2) + struct.pack(">I", crc32c(after_crc)) + after_crc
    return bytes(out)


def scan_batches(segment) -> Tuple[np.ndarray, np.ndarray]:
    """Host walk of the batch chain → (batch_pos int64[nb], record_count int64[nb]).  A trailing partial batch (a fetch response
    may end in one) is ignored, like consumers do."""
    buf = segment.numpy() if isinstance(segment, torch.Tensor) else np.frombuffer(memoryview(segment).cast("B"), dtype=np.uint8)
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    lib = _ffi.lib()
    p = buf.ctypes.data if buf.size else 0
    if not buf.size:
        return np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.int64)
    nb = int(lib.b2_kafka_scan_batches(p, int(buf.size), None, None, 0))
    if nb < 0:
        raise ValueError("b2_kafka_scan_batches refused its arguments")
    bp, rc = np.zeros(nb, dtype=np.int64), np.zeros(nb, dtype=np.int64)
    if nb:
        lib.b2_kafka_scan_batches(p, int(buf.size), bp.ctypes.data, rc.ctypes.data, nb)
    return bp, rc


def write_segment_uniform(values: np.ndarray, records_per_batch: int = 256, base_offset: int = 0) -> np.ndarray:
    """Vectorised ``write_segment`` for n values of one length (uint8[n, L]; n a multiple of ``records_per_batch``): uncompressed
    batches, null keys, no headers.  Every batch then has the same layout, so one batch is laid out in Python and the rest is NumPy
    (the CRC-32C of all batches advances together, one byte position per step).  Returns the segment as uint8[...]."""
    values = np.ascontiguousarray(values, dtype=np.uint8)
    n, L = values.shape
    if n == 0 or n % records_per_batch:
        raise ValueError("write_segment_uniform wants a whole number of batches")
    nb = n // records_per_batch
    recs, vpos = bytearray(), []
    for i in range(records_per_batch):
        body = b"\x00" + _varint(3 * i) + _varint(i) + _varint(-1) + _varint(L)
        recs += _varint(len(body) + L + 1) + body
        vpos.append(len(recs))
        recs += bytes(L) + _varint(0)
    cnt = records_per_batch
    after_crc = struct.pack(">hiqqqhii", 0, cnt - 1, 1567606196000, 1567606196000 + 3 * cnt, -1, -1, -1, cnt) + bytes(recs)
    blen = 4 + 1 + 4 + len(after_crc)
    one = np.frombuffer(struct.pack(">qiib", 0, blen, 0, 2) + b"\0\0\0\0" + after_crc, dtype=np.uint8)
    seg = np.tile(one, (nb, 1))
    cols = (np.asarray(vpos, dtype=np.int64) + _HEADER)[:, None] + np.arange(L, dtype=np.int64)[None, :]  # value bytes inside a batch
    seg[:, cols.reshape(-1)] = values.reshape(nb, records_per_batch * L)
    offs = (base_offset + np.arange(nb, dtype=np.int64) * records_per_batch).astype(">i8")
    seg[:, 0:8] = offs.view(np.uint8).reshape(nb, 8)
    tab = _crc32c_table()
    c = np.full(nb, 0xFFFFFFFF, dtype=np.uint32)
    for j in range(21, seg.shape[1]):
        c = tab[(c ^ seg[:, j]) & 0xFF] ^ (c >> 8)
    seg[:, 17:21] = (c ^ 0xFFFFFFFF).astype(">u4").view(np.uint8).reshape(nb, 4)
    return seg.reshape(-1)


def _index(segment, verify_crc: bool, device: Optional[torch.device], raise_on_error: bool):
    """Shared front half: host chain walk, inflate if needed, CRC + record walk → (segment on the device, begin, end, n, status)."""
    if isinstance(segment, torch.Tensor):
        host = segment.cpu() if segment.is_cuda else segment
        seg_t = segment
    else:
        host = torch.from_numpy(np.frombuffer(segment, dtype=np.uint8).copy() if isinstance(segment, (bytes, bytearray, memoryview))
                                else np.ascontiguousarray(segment, dtype=np.uint8))
        seg_t = host
    dev = device or (seg_t.device if seg_t.is_cuda else torch.device("cuda", torch.cuda.current_device()))
    pos, cnt = scan_batches(host)
    nb, n = len(pos), int(cnt.sum())
    base = np.zeros(nb, dtype=np.int64)
    if nb:
        np.cumsum(cnt[:-1], out=base[1:])
    d = seg_t.to(dev).contiguous()
    lib, ctx = _ffi.lib(), get_ctx(dev.index)
    pre_status = None
    hb = host.numpy()
    if nb and (hb[pos + 22] & 7).any():
        # compressed batches: inflate into a rebuilt, uncompressed copy of the segment (two passes, one thread per batch), then go on
        with torch.cuda.device(dev):
            pos_d = torch.from_numpy(pos).to(dev)
            sizes = torch.zeros(nb, dtype=torch.int64, device=dev)
            pre_status = torch.zeros(nb, dtype=torch.int32, device=dev)
            _ffi.check(lib.b2_kafka_inflate_sizes(ctx, _ffi.ptr(d), int(d.numel()), _ffi.ptr(pos_d), nb, 1 if verify_crc else 0, _ffi.ptr(sizes),
                                                  _ffi.ptr(pre_status), _stream_ptr()), "b2_kafka_inflate_sizes")
            out_pos = torch.zeros(nb + 1, dtype=torch.int64, device=dev)
            torch.cumsum(sizes + _HEADER, 0, out=out_pos[1:])
            total_out = int(out_pos[-1])
            if total_out >= 2 ** 31:
                raise ValueError("inflated segment exceeds int32 offsets")
            rebuilt = torch.zeros(max(total_out, 1), dtype=torch.uint8, device=dev)
            _ffi.check(lib.b2_kafka_inflate(ctx, _ffi.ptr(d), int(d.numel()), _ffi.ptr(pos_d), nb, _ffi.ptr(sizes), _ffi.ptr(out_pos),
                                            _ffi.ptr(pre_status), _ffi.ptr(rebuilt), total_out, _stream_ptr()), "b2_kafka_inflate")
            d, pos, verify_crc = rebuilt[:max(total_out, 1)], out_pos[:nb].cpu().numpy(), False
    begin = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)
    end = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)
    status = torch.zeros(max(nb, 1), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        pos_d, base_d = torch.from_numpy(np.ascontiguousarray(pos)).to(dev), torch.from_numpy(base).to(dev)
        _ffi.check(lib.b2_kafka_index_records(ctx, _ffi.ptr(d), int(d.numel()), _ffi.ptr(pos_d), _ffi.ptr(base_d), nb, 1 if verify_crc else 0,
                                              _ffi.ptr(begin), _ffi.ptr(end), _ffi.ptr(status), _stream_ptr()), "b2_kafka_index_records")
        status = status[:nb]
        if pre_status is not None:  # a batch refused before / while inflating: its status wins and its records are blanked
            refused = pre_status != 0
            status = torch.where(refused, pre_status, status)
            if n and bool(refused.any()):
                cnt_d = torch.from_numpy(cnt).to(dev)
                rec_bad = torch.repeat_interleave(refused, cnt_d)
                at = torch.repeat_interleave(pos_d.to(torch.int32), cnt_d)  # empty, at the batch's position: the ranges stay ascending
                begin[:n] = torch.where(rec_bad, at, begin[:n])
                end[:n] = torch.where(rec_bad, at, end[:n])
        if raise_on_error and nb:
            bad = torch.nonzero(status).flatten()
            if bad.numel():
                b = int(bad[0])
                raise FrameError(f"record batch {b} at byte {int(pos[b])}: {FRAME_STATUS.get(int(status[b]), '?')} "
                                 f"({bad.numel()} of {nb} batches refused)")
    return d, begin, end, n, status


def index_segment(segment, verify_crc: bool = True, device: Optional[torch.device] = None, skip_prefix: int = 0,
                  raise_on_error: bool = True):
    """Log-segment / fetch-response bytes → ``(RecordSpans, batch status int32[nb] on the device)``: the message values as byte
    ranges inside the segment (inside its inflated copy when batches were compressed).  Records of a refused batch are empty."""
    d, begin, end, n, status = _index(segment, verify_crc, device, raise_on_error)
    spans = RecordSpans(d, begin, end, n, skip_prefix=skip_prefix)
    spans.avg_record_bytes = (int(d.numel()) + max(n, 1) - 1) // max(n, 1)
    return spans, status


def values_from_segment(segment, verify_crc: bool = True, device: Optional[torch.device] = None, skip_prefix: int = 0,
                        raise_on_error: bool = True):
    """Log-segment / fetch-response bytes (``bytes``, NumPy uint8, or a CPU / CUDA uint8 tensor; the batch chain is walked on
    the host copy) → ``(RecordBatch of the message values, batch status int32[nb] on the device)``.  Records of a refused batch
    come out empty; with ``raise_on_error`` a refused batch raises :class:`FrameError` instead."""
    d, begin, end, n, status = _index(segment, verify_crc, device, raise_on_error)
    dev = d.device
    lib, ctx = _ffi.lib(), get_ctx(dev.index)
    with torch.cuda.device(dev):
        offs = torch.zeros(n + 1, dtype=torch.int64, device=dev)
        if n:
            torch.cumsum((end[:n] - begin[:n]).to(torch.int64), 0, out=offs[1:])
        total = int(offs[-1])
        if total >= 2 ** 31:
            raise ValueError("values exceed int32 offsets")
        offs32 = offs.to(torch.int32)
        out = torch.empty(max(total, 1), dtype=torch.uint8, device=dev)
        _ffi.check(lib.b2_kafka_gather_values(ctx, _ffi.ptr(d), _ffi.ptr(begin), _ffi.ptr(end), n, _ffi.ptr(offs32), _ffi.ptr(out),
                                              _stream_ptr()), "b2_kafka_gather_values")
    return RecordBatch(out[:total], offs32, dev, skip_prefix=skip_prefix), status
