"""K9: Kafka record-batch (v2) framing on the GPU — values out of log-segment bytes, CRC-32C check, refused batches — and the
decode kernels running on the result (segment bytes → scores)."""
import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import decode as dec, schemas, synth

pytestmark = pytest.mark.gpu


def test_crc32c_known_answers():
    assert bs.kafka.crc32c(b"123456789") == 0xE3069283 and bs.kafka.crc32c(bytes(32)) == 0x8A9136AA  # RFC 3720 B.4


def test_values_keys_headers_nulls(cuda_device):
    g = np.random.default_rng(1)
    vals = [bytes(g.integers(0, 256, int(g.integers(0, 300)), dtype=np.uint8)) for _ in range(2000)]
    vals[5], vals[77] = None, b""
    keys = [None if i % 3 else b"car%d" % i for i in range(2000)]
    seg = bs.kafka.write_segment(vals, records_per_batch=97, keys=keys, headers=True)
    rb, st = bs.kafka.values_from_segment(seg)
    assert rb.n == 2000 and not st.any() and st.numel() == 21
    d, o = rb.data.cpu().numpy().tobytes(), rb.offsets.cpu().numpy()
    assert [d[o[i]:o[i + 1]] for i in range(2000)] == [v or b"" for v in vals]


def test_bad_batches_are_refused_not_fatal(cuda_device):
    vals = [b"v%04d" % i for i in range(300)]
    seg = bytearray(bs.kafka.write_segment(vals, records_per_batch=100))
    pos, _ = bs.kafka.scan_batches(bytes(seg))
    seg[pos[1] + 80] ^= 0x01      # payload bit flip → CRC mismatch
    seg[pos[2] + 22] |= 0x05      # attributes: a codec nobody knows; the stored CRC no longer matches either
    with pytest.raises(bs.kafka.FrameError):
        bs.kafka.values_from_segment(bytes(seg))
    rb, st = bs.kafka.values_from_segment(bytes(seg), raise_on_error=False)
    assert st.cpu().tolist() == [0, 2, 2]
    assert bs.kafka.values_from_segment(bytes(seg), raise_on_error=False, verify_crc=False)[1].cpu().tolist() == [0, 0, 4]
    d, o = rb.data.cpu().numpy().tobytes(), rb.offsets.cpu().numpy()
    got = [d[o[i]:o[i + 1]] for i in range(300)]
    assert got[:100] == vals[:100] and all(g == b"" for g in got[100:])
    rb2, st2 = bs.kafka.values_from_segment(bytes(seg[:pos[2] + 40]), raise_on_error=False, verify_crc=False)  # partial trailing batch
    assert rb2.n == 200 and st2.cpu().tolist() == [0, 0]


def test_segment_to_scores(cuda_device):
    """Broker bytes → framing → fused Avro decode + score == decode + score of the bare record batch."""
    rows = synth.sensor_rows(20_000, seed=3)
    data, offs = synth.avro_a_batch(rows)
    vals = [data[offs[i]:offs[i + 1]].tobytes() for i in range(len(offs) - 1)]
    seg = bs.kafka.write_segment(vals, records_per_batch=512)
    rb, _ = bs.kafka.values_from_segment(seg, skip_prefix=5)
    np.testing.assert_array_equal(rb.offsets.cpu().numpy(), offs)
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    h = k.layers.Dense(14, activation="tanh")(inp)
    h = k.layers.Dense(7, activation="relu")(h)
    h = k.layers.Dense(7, activation="tanh")(h)
    m = k.models.Model(inputs=inp, outputs=k.layers.Dense(18, activation="relu")(h))
    m.set_normalization(*schemas.norm_arrays())
    a = dec.decode_score(m, rb, avro_schema=schemas.schema_a_json())
    b = dec.decode_score(m, dec.RecordBatch(data, offs, skip_prefix=5), avro_schema=schemas.schema_a_json())
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def _golden():
    import json
    import os

    return json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kafka_v2_batches.json")))


def test_hand_assembled_protocol_batches(cuda_device):
    """Batches packed field by field from the protocol guide (tests/golden/make_kafka_golden.py, bitwise CRC-32C): keyed records,
    headers (incl. a null header value), null and empty values, a 130-byte key (two-byte length varint), a control batch."""
    fx = _golden()
    seg = bytes.fromhex(fx["segment_hex"])
    rb, st = bs.kafka.values_from_segment(seg, raise_on_error=False)
    assert st.cpu().tolist() == fx["batch_status"]  # [ok, control batch, ok]
    d, o = rb.data.cpu().numpy().tobytes(), rb.offsets.cpu().numpy()
    got = [d[o[i]:o[i + 1]] for i in range(rb.n)]
    assert got == [b"" if h is None else bytes.fromhex(h) for h in fx["values_hex"]]
    with pytest.raises(bs.kafka.FrameError, match="control batch"):
        bs.kafka.values_from_segment(seg)


def test_hostile_lengths_are_truncation_not_a_fault(cuda_device):
    """Key / value lengths near 2^62 and negative key lengths, each inside a batch whose CRC is valid: refused as truncated,
    no range of the batch survives, and the context is still usable afterwards."""
    fx = _golden()
    for name, hx in fx["hostile_hex"].items():
        for crc in (True, False):
            rb, st = bs.kafka.values_from_segment(bytes.fromhex(hx), raise_on_error=False, verify_crc=crc)
            assert st.cpu().tolist() == [3], name
            o = rb.offsets.cpu().numpy()
            assert (o == 0).all(), name
    torch.cuda.synchronize()
    rb, st = bs.kafka.values_from_segment(bs.kafka.write_segment([b"still", b"alive"]))
    assert rb.n == 2 and not st.any()


@pytest.mark.parametrize("compression", ["gzip", "snappy", "snappy-xerial", "lz4", "zstd", "zstd-stream", ["gzip", None, "lz4", "snappy", "zstd"]])
def test_compressed_batches_are_inflated(cuda_device, compression):
    """Producer-side compression.type (records section compressed as a whole): the values come out as from the uncompressed
    segment, CRC-32C is checked over the bytes as stored, a damaged stream / a batch of an unknown codec is refused without touching the others."""
    g = np.random.default_rng(3)
    rows = synth.sensor_rows(3000, seed=8)
    data, offs = synth.avro_a_batch(rows)
    vals = [data[offs[i]:offs[i + 1]].tobytes() for i in range(3000)]
    vals[17], vals[18] = None, b""
    keys = [None if i % 5 else b"car%d" % i for i in range(3000)]
    seg = bs.kafka.write_segment(vals, records_per_batch=250, keys=keys, headers=True, compression=compression)
    plain = bs.kafka.write_segment(vals, records_per_batch=250, keys=keys, headers=True)
    rb, st = bs.kafka.values_from_segment(seg)
    ref, _ = bs.kafka.values_from_segment(plain)
    assert not st.any() and torch.equal(rb.offsets, ref.offsets) and torch.equal(rb.data, ref.data)
    # damage: flip a byte inside the compressed payload of batch 1 (CRC mismatch), then with the CRC check off (stream error or
    # different bytes — either way the batch must not take the others down)
    pos, _ = bs.kafka.scan_batches(seg)
    bad = bytearray(seg)
    bad[pos[1] + 61 + 20] ^= 0x5A
    rb2, st2 = bs.kafka.values_from_segment(bytes(bad), raise_on_error=False)
    s2 = st2.cpu().tolist()
    assert s2[1] == 2 and s2[0] == 0 and all(v == 0 for v in s2[2:])
    o, o2 = ref.offsets.cpu().numpy(), rb2.offsets.cpu().numpy()
    assert (np.diff(o2)[250:500] == 0).all() and np.array_equal(np.diff(o2)[:250], np.diff(o)[:250]) and np.array_equal(np.diff(o2)[500:], np.diff(o)[500:])
    rb3, st3 = bs.kafka.values_from_segment(bytes(bad), raise_on_error=False, verify_crc=False)
    assert st3.cpu().tolist()[0] == 0 and rb3.n == 3000
    # a codec number nobody knows (5) is reported, not inflated
    z = bytearray(seg)
    z[pos[2] + 22] = (z[pos[2] + 22] & ~7) | 5
    _, st4 = bs.kafka.values_from_segment(bytes(z), raise_on_error=False, verify_crc=False)
    assert st4.cpu().tolist()[2] == 4


def _ae18():
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    h = k.layers.Dense(14, activation="tanh")(inp)
    h = k.layers.Dense(7, activation="relu")(h)
    h = k.layers.Dense(7, activation="tanh")(h)
    return k.models.Model(inputs=inp, outputs=k.layers.Dense(18, activation="relu")(h))


@pytest.mark.parametrize("per_batch", [1, 31, 32, 33, 700, 20_000])
def test_warp_per_batch_walk_all_batch_shapes(cuda_device, per_batch):
    """One warp per batch, 32 records per round, 4 KB stretches between CRC rows: batch sizes around every one of those edges,
    values from 0 to 600 bytes (records longer than a CRC row and shorter than a word), keys and headers in between."""
    g = np.random.default_rng(per_batch)
    n = 20_000 if per_batch > 33 else 40 * per_batch + 7
    vals = [bytes(g.integers(0, 256, int(g.integers(0, 600 if i % 50 == 0 else 90)), dtype=np.uint8)) for i in range(n)]
    keys = [None if i % 5 else b"k" * (i % 200) for i in range(n)]
    seg = bs.kafka.write_segment(vals, records_per_batch=per_batch, keys=keys, headers=True)
    rb, st = bs.kafka.values_from_segment(seg)
    assert rb.n == n and not st.any()
    d, o = rb.data.cpu().numpy().tobytes(), rb.offsets.cpu().numpy()
    assert [d[o[i]:o[i + 1]] for i in range(n)] == vals
    spans, st2 = bs.kafka.index_segment(seg)
    b, e = spans.begin.cpu().numpy()[:n], spans.end.cpu().numpy()[:n]
    assert (b[1:] >= e[:-1]).all() and (e >= b).all()
    assert [seg[b[i]:e[i]] for i in range(0, n, 17)] == vals[::17]


def test_crc_flags_a_flip_anywhere_in_the_checked_bytes(cuda_device):
    """The lane-interleaved CRC-32C: one flipped bit at the first checked byte, inside the front padding row, at row and tile
    borders, in the middle and at the very last byte must each turn exactly that batch into BAD_CRC; the two bytes in front of the
    checked range (magic is checked on its own, the epoch is not covered) must not."""
    g = np.random.default_rng(9)
    vals = [bytes(g.integers(0, 256, 63, dtype=np.uint8)) for _ in range(1000)]
    seg0 = bs.kafka.write_segment(vals, records_per_batch=250)
    pos, _ = bs.kafka.scan_batches(seg0)
    blen = int(pos[1] - pos[0])
    assert blen > 3 * 4096
    for b in (0, 2):
        for off in (21, 22, 24, 25, 60, 61, 21 + 127, 21 + 128, 21 + 129, 21 + 4095, 21 + 4096, 21 + 4097, blen // 2, blen - 5, blen - 4, blen - 1):
            seg = bytearray(seg0)
            seg[pos[b] + off] ^= 0x10
            st = bs.kafka.index_segment(bytes(seg), raise_on_error=False)[1].cpu().tolist()
            want = [0, 0, 0, 0]
            want[b] = 2
            if off == 22:
                assert st[b] != 0, off          # attributes: the codec bits change the verdict before the CRC is looked at
            else:
                assert st == want, (b, off, st)
    seg = bytearray(seg0)
    seg[pos[1] + 13] ^= 0x01                     # partitionLeaderEpoch: in front of the CRC'd range
    assert bs.kafka.index_segment(bytes(seg))[1].cpu().tolist() == [0, 0, 0, 0]


@pytest.mark.parametrize("fmt", ["avro_a", "csv", "json"])
def test_scores_straight_from_the_segment_equal_scores_of_the_packed_values(cuda_device, fmt):
    """index_segment → decode_score on the byte ranges inside the segment (no gather) == values_from_segment → decode_score, bit for
    bit, with a refused batch in the middle (its records are empty in both) and malformed records."""
    n = 30_000
    rows = synth.sensor_rows(n, seed=21)
    if fmt == "avro_a":
        data, offs = synth.avro_a_batch(rows)
        kw, skip = dict(avro_schema=schemas.schema_a_json()), 5
    elif fmt == "csv":
        data, offs = synth.csv_batch(rows, seed=21)
        kw, skip = dict(csv_kinds=[0, 0] + [1] * 18), 0
    else:
        data, offs = synth.json_batch(rows, labels=(np.arange(n) % 64 == 0))
        kw, skip = dict(json_keys=list(schemas.FEATURE_NAMES), json_label="failure_occurred"), 0
    data, offs, _ = synth.corrupt_batch(data, offs, fmt, 1e-3, seed=4)
    vals = [data[offs[i]:offs[i + 1]].tobytes() for i in range(n)]
    seg = bytearray(bs.kafka.write_segment(vals, records_per_batch=333))
    pos, _ = bs.kafka.scan_batches(bytes(seg))
    seg[pos[7] + 200] ^= 0x40
    seg = bytes(seg)
    m = _ae18()
    m.set_normalization(*schemas.norm_arrays())
    spans, st = bs.kafka.index_segment(seg, skip_prefix=skip, raise_on_error=False)
    rb, st2 = bs.kafka.values_from_segment(seg, skip_prefix=skip, raise_on_error=False)
    assert st.cpu().tolist() == st2.cpu().tolist() and st[7] == 2 and int((st != 0).sum()) == 1
    ca, cb = torch.zeros(4, dtype=torch.int64, device="cuda"), torch.zeros(4, dtype=torch.int64, device="cuda")
    a = dec.decode_score(m, spans, counters=ca, **kw)
    b = dec.decode_score(m, rb, counters=cb, **kw)
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        if x is None or y is None:
            assert x is None and y is None
            continue
        assert np.array_equal(x.cpu().numpy().view(np.uint8), y.cpu().numpy().view(np.uint8))
    assert ca[:3].tolist() == cb[:3].tolist() and int(ca[0]) > n * 0.98
    bad = a[3].cpu().numpy()
    assert (bad[7 * 333:8 * 333] != 0).all()


def test_decode_only_on_ranges_inside_the_segment(cuda_device):
    """b2_decode_spans: the three decode-only kernels on the value ranges of a segment == the same kernels on the packed values
    (matrix, label, status bit for bit; CSV string-field spans point at the same bytes)."""
    n = 5_000
    rows = synth.sensor_rows(n, seed=8)
    for fmt in ("avro_a", "csv", "json"):
        if fmt == "avro_a":
            data, offs = synth.avro_a_batch(rows)
        elif fmt == "csv":
            data, offs = synth.csv_batch(rows, seed=8)
        else:
            data, offs = synth.json_batch(rows, labels=(np.arange(n) % 3 == 0))
        data, offs, _ = synth.corrupt_batch(data, offs, fmt, 2e-3, seed=8)
        seg = bs.kafka.write_segment([data[offs[i]:offs[i + 1]].tobytes() for i in range(n)], records_per_batch=211, headers=True)
        skip = 5 if fmt == "avro_a" else 0
        spans, _ = bs.kafka.index_segment(seg, skip_prefix=skip)
        rb, _ = bs.kafka.values_from_segment(seg, skip_prefix=skip)
        if fmt == "avro_a":
            a = dec.decode_avro_matrix(spans, schemas.schema_a_json(), raise_on_error=False)[:3]
            b = dec.decode_avro_matrix(rb, schemas.schema_a_json(), raise_on_error=False)[:3]
        elif fmt == "json":
            a = dec.decode_json_matrix(spans, list(schemas.FEATURE_NAMES), "failure_occurred", raise_on_error=False)
            b = dec.decode_json_matrix(rb, list(schemas.FEATURE_NAMES), "failure_occurred", raise_on_error=False)
        else:
            kinds = [0, 2] + [1] * 18
            oa, sa, ta = dec.decode_csv_matrix(spans, kinds, raise_on_error=False)
            ob, sb_, tb = dec.decode_csv_matrix(rb, kinds, raise_on_error=False)
            a, b = (oa, ta), (ob, tb)
            good = (ta == 0).cpu().numpy()
            sa, sb_ = sa.cpu().numpy()[good, 0], sb_.cpu().numpy()[good, 0]
            packed = rb.data.cpu().numpy().tobytes()
            assert [seg[x:y] for x, y in sa[::97]] == [packed[x:y] for x, y in sb_[::97]]
            assert seg[sa[0][0]:sa[0][1]].startswith(b"car")
        for x, y in zip(a, b):
            assert np.array_equal(x.cpu().numpy().view(np.uint8), y.cpu().numpy().view(np.uint8)), fmt
        assert int((a[-1] != 0).sum()) >= 5
