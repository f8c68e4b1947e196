"""Kafka ingest on the GPU, event-timed: record-batch framing (CRC-32C + record walk, one warp per batch), the value gather, and
the fused decode + score fed (a) with the packed values and (b) with the byte ranges inside the segment (no gather).

    python tools/kafka_probe.py [log2_records=20] [records_per_batch=256]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import b200_sensor as bs
from b200_sensor import _ffi, decode as dec, schemas, synth
from b200_sensor.keras_compat import _stream_ptr, get_ctx


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e-3


def main():
    lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    per = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    n = 1 << lg
    dev = torch.device("cuda", 0)
    rows = synth.sensor_rows(n, seed=5)
    data, offs = synth.avro_a_batch(rows)
    L = int(offs[1] - offs[0])
    assert (np.diff(offs) == L).all()
    seg = bs.kafka.write_segment_uniform(data.reshape(n, L), records_per_batch=per)
    d = torch.from_numpy(seg).to(dev)
    pos, cnt = bs.kafka.scan_batches(seg)
    nb = len(pos)
    base = np.zeros(nb, dtype=np.int64)
    np.cumsum(cnt[:-1], out=base[1:])
    pos_d, base_d = torch.from_numpy(pos).to(dev), torch.from_numpy(base).to(dev)
    begin, end = torch.zeros(n, dtype=torch.int32, device=dev), torch.zeros(n, dtype=torch.int32, device=dev)
    status = torch.zeros(nb, dtype=torch.int32, device=dev)
    lib, ctx = _ffi.lib(), get_ctx(0)

    def index(crc=1):
        _ffi.check(lib.b2_kafka_index_records(ctx, _ffi.ptr(d), int(d.numel()), _ffi.ptr(pos_d), _ffi.ptr(base_d), nb, crc, _ffi.ptr(begin),
                                              _ffi.ptr(end), _ffi.ptr(status), _stream_ptr()), "index")

    t_idx = timed(index)
    t_idx0 = timed(lambda: index(0))
    assert not status.any()
    o32 = torch.zeros(n + 1, dtype=torch.int32, device=dev)
    o32[1:] = torch.cumsum(end - begin, 0)
    out = torch.empty(int(o32[-1]), dtype=torch.uint8, device=dev)

    def gather():
        _ffi.check(lib.b2_kafka_gather_values(ctx, _ffi.ptr(d), _ffi.ptr(begin), _ffi.ptr(end), n, _ffi.ptr(o32), _ffi.ptr(out), _stream_ptr()), "gather")

    t_g = timed(gather)
    assert np.array_equal(out.cpu().numpy(), data)
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    h = k.layers.Dense(14, activation="tanh")(inp)
    h = k.layers.Dense(7, activation="relu")(h)
    h = k.layers.Dense(7, activation="tanh")(h)
    m = k.models.Model(inputs=inp, outputs=k.layers.Dense(18, activation="relu")(h))
    m.set_normalization(*schemas.norm_arrays())
    rb = dec.RecordBatch(out, o32, dev, skip_prefix=5)
    spans = dec.RecordSpans(d, begin, end, n, skip_prefix=5)
    spans.avg_record_bytes = (int(d.numel()) + n - 1) // n
    bufs = [torch.empty(n, dtype=torch.float32, device=dev), torch.empty(n, dtype=torch.uint8, device=dev),
            torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.int32, device=dev)]
    kw = dict(avro_schema=schemas.schema_a_json(), mse=bufs[0], flag=bufs[1], label=bufs[2], status=bufs[3])
    t_packed = timed(lambda: dec.decode_score(m, rb, **kw))
    ref = [b.clone() for b in bufs]
    t_spans = timed(lambda: dec.decode_score(m, spans, **kw))
    same = all(torch.equal(a.view(torch.uint8), b.view(torch.uint8)) for a, b in zip(ref, bufs))

    def chain_gather():
        index()
        gather()
        dec.decode_score(m, rb, **kw)

    def chain_spans():
        index()
        dec.decode_score(m, spans, **kw)

    t_cg, t_cs = timed(chain_gather), timed(chain_spans)
    gb = d.numel() / 1e9
    print(f"segment: {n} Avro-A records of {L} B in {nb} batches of {per} ({d.numel() / 2**20:.1f} MiB, {d.numel() / n:.1f} B/record)")
    print(f"kafka_index (CRC-32C + record walk): {t_idx * 1e6:8.1f} us  {gb / t_idx:7.1f} GB/s  {n / t_idx:.3g} records/s")
    print(f"kafka_index (record walk, CRC unchecked): {t_idx0 * 1e6:8.1f} us  {gb / t_idx0:7.1f} GB/s")
    print(f"kafka_gather:                        {t_g * 1e6:8.1f} us  {gb / t_g:7.1f} GB/s")
    print(f"decode+score, packed values:         {t_packed * 1e6:8.1f} us  {n / t_packed:.3g} records/s")
    print(f"decode+score, ranges in the segment: {t_spans * 1e6:8.1f} us  {n / t_spans:.3g} records/s   bit-identical: {same}")
    print(f"segment -> scores, index + gather + decode: {t_cg * 1e6:8.1f} us  {n / t_cg:.3g} records/s")
    print(f"segment -> scores, index + decode in place: {t_cs * 1e6:8.1f} us  {n / t_cs:.3g} records/s")


def variable_lengths():
    """Text values (CSV lines, lengths differ from record to record): the record walk cannot guess a stride."""
    n, per = 1 << 17, 256
    dev = torch.device("cuda", 0)
    data, offs = synth.csv_batch(synth.sensor_rows(n, seed=6), seed=6)
    seg = np.frombuffer(bs.kafka.write_segment([data[offs[i]:offs[i + 1]].tobytes() for i in range(n)], records_per_batch=per), dtype=np.uint8)
    d = torch.from_numpy(seg.copy()).to(dev)
    pos, cnt = bs.kafka.scan_batches(seg)
    nb = len(pos)
    base = np.zeros(nb, dtype=np.int64)
    np.cumsum(cnt[:-1], out=base[1:])
    pos_d, base_d = torch.from_numpy(pos).to(dev), torch.from_numpy(base).to(dev)
    begin, end = torch.zeros(n, dtype=torch.int32, device=dev), torch.zeros(n, dtype=torch.int32, device=dev)
    status = torch.zeros(nb, dtype=torch.int32, device=dev)
    lib, ctx = _ffi.lib(), get_ctx(0)
    t = timed(lambda: _ffi.check(lib.b2_kafka_index_records(ctx, _ffi.ptr(d), int(d.numel()), _ffi.ptr(pos_d), _ffi.ptr(base_d), nb, 1,
                                                            _ffi.ptr(begin), _ffi.ptr(end), _ffi.ptr(status), _stream_ptr()), "index"))
    assert not status.any()
    print(f"segment: {n} CSV records ({d.numel() / n:.1f} B/record) in {nb} batches of {per}")
    print(f"kafka_index (CRC-32C + record walk): {t * 1e6:8.1f} us  {d.numel() / 1e9 / t:7.1f} GB/s  {n / t:.3g} records/s")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "csv":
        variable_lengths()
    else:
        main()
