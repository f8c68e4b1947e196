"""Kafka message-format-v2 fixtures (tests/golden/kafka_v2_batches.json, hand-assembled field by field from the protocol guide by
tests/golden/make_kafka_golden.py) against the host-side framing code: batch chain walk, CRC-32C, and the segment writer the GPU
tests use as their input generator — read back here by a plain walker that shares no code with it."""
import json
import os
import struct

import numpy as np
import pytest

import b200_sensor as bs

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def fx():
    return json.load(open(os.path.join(GOLDEN, "kafka_v2_batches.json")))


def _varint(buf, p):
    u, shift = 0, 0
    while True:
        b = buf[p]
        p += 1
        u |= (b & 0x7F) << shift
        shift += 7
        if b < 0x80:
            return (u >> 1) ^ -(u & 1), p


def walk_segment(seg: bytes):
    """[(attributes, [(key, value, n_headers), ...]), ...] straight from the protocol layout (no CRC check)."""
    out, pos = [], 0
    while pos + 61 <= len(seg):
        base_offset, blen, _epoch, magic, _crc, attrs, last_delta = struct.unpack_from(">qiibIhi", seg, pos)
        assert magic == 2
        (count,) = struct.unpack_from(">i", seg, pos + 57)
        assert last_delta == count - 1
        p, recs = pos + 61, []
        for _ in range(count):
            ln, p = _varint(seg, p)
            end = p + ln
            p += 1
            _, p = _varint(seg, p)
            _, p = _varint(seg, p)
            kl, p = _varint(seg, p)
            key = None if kl < 0 else seg[p:p + kl]
            p += max(kl, 0)
            vl, p = _varint(seg, p)
            val = None if vl < 0 else seg[p:p + vl]
            p += max(vl, 0)
            nh, p = _varint(seg, p)
            for _ in range(nh):
                hk, p = _varint(seg, p)
                p += hk
                hv, p = _varint(seg, p)
                p += max(hv, 0)
            assert p == end
            recs.append((key, val, nh))
        assert p == pos + 12 + blen
        out.append((attrs, recs))
        pos += 12 + blen
    assert pos == len(seg)
    return out


def test_rfc3720_crc32c_answers():
    c = bs.kafka.crc32c
    assert c(bytes(32)) == 0x8A9136AA and c(b"\xff" * 32) == 0x62A8AB43
    assert c(bytes(range(32))) == 0x46DD794E and c(bytes(range(31, -1, -1))) == 0x113FDB5C and c(b"123456789") == 0xE3069283


def test_golden_segment_is_what_the_protocol_says(fx):
    seg = bytes.fromhex(fx["segment_hex"])
    batches = walk_segment(seg)
    assert [len(r) for _, r in batches] == fx["record_counts"]
    assert [a for a, _ in batches] == [0, 0x30, 0]
    vals = [v for _, recs in batches for _, v, _ in recs]
    want = [None if h is None else bytes.fromhex(h) for h in fx["values_hex"]]
    want[6] = vals[6]  # the control record's value is not a consumer-visible message
    assert vals == want
    pos, cnt = bs.kafka.scan_batches(seg)
    assert cnt.tolist() == fx["record_counts"] and pos[0] == 0
    for p, nxt in zip(pos.tolist(), pos.tolist()[1:] + [len(seg)]):
        stored = struct.unpack_from(">I", seg, p + 17)[0]
        assert bs.kafka.crc32c(seg[p + 21:nxt]) == stored
    assert struct.unpack_from(">I", seg, 17)[0] == fx["batch0_crc"]


def test_write_segment_reads_back_with_the_independent_walker():
    g = np.random.default_rng(5)
    vals = [bytes(g.integers(0, 256, int(g.integers(0, 200)), dtype=np.uint8)) for _ in range(300)]
    vals[3], vals[10] = None, b""
    keys = [None if i % 4 else b"key%d" % i for i in range(300)]
    seg = bs.kafka.write_segment(vals, records_per_batch=64, keys=keys, headers=True, base_offset=77)
    batches = walk_segment(seg)
    assert [len(r) for _, r in batches] == [64, 64, 64, 64, 44]
    assert [v for _, recs in batches for _, v, _ in recs] == vals
    assert [k for _, recs in batches for k, _, _ in recs] == keys
    assert struct.unpack_from(">q", seg, 0)[0] == 77


def test_scan_batches_and_uniform_writer_cpu():
    """The C chain walk (b2_kafka_scan_batches) against a struct.unpack walk, incl. a trailing partial batch and an empty buffer;
    the vectorised segment writer against the record-by-record one."""
    import struct

    import b200_sensor as bs

    rng = np.random.default_rng(5)
    vals = [bytes(rng.integers(0, 256, int(rng.integers(0, 300)), dtype=np.uint8)) for _ in range(700)]
    seg = bs.kafka.write_segment(vals, records_per_batch=97, headers=True)

    def walk(buf):
        pos, out = 0, []
        while pos + 61 <= len(buf):
            (blen,) = struct.unpack_from(">i", buf, pos + 8)
            if blen < 49 or pos + 12 + blen > len(buf):
                break
            out.append((pos, max(struct.unpack_from(">i", buf, pos + 57)[0], 0)))
            pos += 12 + blen
        return out

    for cut in (0, 1, 30, 61, 500):
        buf = seg[:len(seg) - cut]
        bp, rc = bs.kafka.scan_batches(buf)
        assert list(zip(bp.tolist(), rc.tolist())) == walk(buf), cut
    bp, rc = bs.kafka.scan_batches(b"")
    assert bp.size == 0 and rc.size == 0
    bad = bytearray(seg)
    bad[8:12] = struct.pack(">i", -5)  # a negative batch length ends the walk at once
    assert bs.kafka.scan_batches(bytes(bad))[0].size == 0

    fixed = rng.integers(0, 256, (512, 63), dtype=np.uint8)
    a = bs.kafka.write_segment_uniform(fixed, records_per_batch=128, base_offset=11)
    b = bs.kafka.write_segment([bytes(v) for v in fixed], records_per_batch=128, base_offset=11)
    assert bytes(a) == b
