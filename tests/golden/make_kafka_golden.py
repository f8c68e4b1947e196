"""Hand-assembled Kafka record batches (message format v2) for tests/test_kafka_gpu.py and tests/test_kafka_golden_cpu.py.

Independent of b200_sensor.kafka.write_segment on purpose: every field below is packed explicitly, in the order and with the
widths of the Kafka protocol guide's "RecordBatch" / "Record" / "Header" layout (kafka.apache.org/documentation/#recordbatch),
and the checksum comes from a bit-by-bit CRC-32C (Castagnoli, reflected polynomial 0x82F63B78) whose RFC 3720 §B.4 answers are
asserted first.  Run:  python tests/golden/make_kafka_golden.py  → tests/golden/kafka_v2_batches.json
"""
import json
import os
import struct


def crc32c_bitwise(data: bytes) -> int:
    c = 0xFFFFFFFF
    for b in data:
        c ^= b
        for _ in range(8):
            c = (c >> 1) ^ (0x82F63B78 if c & 1 else 0)
    return c ^ 0xFFFFFFFF


# RFC 3720 B.4 "CRC Examples"
assert crc32c_bitwise(bytes(32)) == 0x8A9136AA
assert crc32c_bitwise(b"\xff" * 32) == 0x62A8AB43
assert crc32c_bitwise(bytes(range(32))) == 0x46DD794E
assert crc32c_bitwise(bytes(range(31, -1, -1))) == 0x113FDB5C
assert crc32c_bitwise(b"123456789") == 0xE3069283


def zigzag_varint(v: int) -> bytes:
    """Protocol Buffers zig-zag varint, as Kafka's record fields use: 0→00, -1→01, 1→02, -2→03, 150→ac 02."""
    u = (v << 1) ^ (v >> 63)
    out = bytearray()
    while True:
        if u < 0x80:
            out.append(u)
            return bytes(out)
        out.append((u & 0x7F) | 0x80)
        u >>= 7


assert zigzag_varint(0) == b"\x00" and zigzag_varint(-1) == b"\x01" and zigzag_varint(1) == b"\x02"
assert zigzag_varint(-2) == b"\x03" and zigzag_varint(150) == b"\xac\x02" and zigzag_varint(-64) == b"\x7f" and zigzag_varint(64) == b"\x80\x01"


def record(offset_delta, ts_delta, key, value, headers=()):
    body = bytearray()
    body += b"\x00"                                   # attributes: int8, unused
    body += zigzag_varint(ts_delta)                   # timestampDelta: varlong
    body += zigzag_varint(offset_delta)               # offsetDelta: varint
    body += zigzag_varint(-1) if key is None else zigzag_varint(len(key)) + key        # keyLength + key
    body += zigzag_varint(-1) if value is None else zigzag_varint(len(value)) + value  # valueLen + value
    body += zigzag_varint(len(headers))               # headers array count: varint
    for hk, hv in headers:
        body += zigzag_varint(len(hk)) + hk           # headerKeyLength + headerKey
        body += zigzag_varint(-1) if hv is None else zigzag_varint(len(hv)) + hv       # headerValueLength + value
    return zigzag_varint(len(body)) + bytes(body)     # length: varint


def batch(base_offset, records_bytes, n_records, attributes=0, first_ts=1567606196000, max_ts=1567606196009,
          producer_id=-1, producer_epoch=-1, base_sequence=-1):
    crc_part = (struct.pack(">h", attributes)         # attributes: int16 (bits 0-2 codec, 3 timestampType, 4 transactional, 5 control)
                + struct.pack(">i", n_records - 1)    # lastOffsetDelta: int32
                + struct.pack(">q", first_ts)         # baseTimestamp: int64 (named firstTimestamp in older guides)
                + struct.pack(">q", max_ts)           # maxTimestamp: int64
                + struct.pack(">q", producer_id)      # producerId: int64
                + struct.pack(">h", producer_epoch)   # producerEpoch: int16
                + struct.pack(">i", base_sequence)    # baseSequence: int32
                + struct.pack(">i", n_records)        # records: array count int32
                + records_bytes)
    crc = crc32c_bitwise(crc_part)                    # covers attributes .. end of batch
    tail = (struct.pack(">i", 7)                      # partitionLeaderEpoch: int32
            + struct.pack(">b", 2)                    # magic: int8 = 2
            + struct.pack(">I", crc)                  # crc: uint32
            + crc_part)
    return struct.pack(">q", base_offset) + struct.pack(">i", len(tail)) + tail   # baseOffset int64, batchLength int32


def main():
    avro_a = bytes.fromhex("00000000" "01") + bytes(range(58))  # Confluent-framed value stand-in (magic 0, schema id 1, 58 body bytes)
    values = [b"hello", None, b"", avro_a, b"1567606196,car1,0.5", b"x" * 200]
    keys = [b"car1", None, b"", b"k" * 130, None, b"car7"]
    hdrs = [(), ((b"trace", b"abc"),), ((b"a", None), (b"bb", b"")), (), (), ((b"h" * 70, b"v" * 300),)]
    recs = b"".join(record(i, 3 * i, keys[i], values[i], hdrs[i]) for i in range(len(values)))
    b0 = batch(1000, recs, len(values))
    # a control batch (attributes bit 5): one record with a 4-byte key (version, type) and a 6-byte value, as an abort marker has
    ctl = batch(1006, record(0, 0, struct.pack(">hh", 0, 0), struct.pack(">hi", 0, 0)), 1, attributes=0x20 | 0x10,
                producer_id=4000, producer_epoch=1)
    b2 = batch(1007, record(0, 0, None, b"after-the-marker") + record(1, 1, b"k", b"last"), 2)
    # hostile lengths (checksum valid, so only the length checks stand between these and an out-of-bounds read):
    huge = (1 << 62) + 12345
    evil_key = batch(1009, zigzag_varint(30) + b"\x00" + zigzag_varint(0) + zigzag_varint(0) + zigzag_varint(huge) + bytes(16), 1)
    evil_val = batch(1010, zigzag_varint(30) + b"\x00" + zigzag_varint(0) + zigzag_varint(0) + zigzag_varint(-1) + zigzag_varint(huge) + bytes(15), 1)
    evil_neg = batch(1011, zigzag_varint(8) + b"\x00" + zigzag_varint(0) + zigzag_varint(0) + zigzag_varint(-5) + zigzag_varint(1) + b"v" + b"\x00", 1)
    # a good record followed by one whose length runs past the batch: the first record's range must be withdrawn
    trunc = batch(1012, record(0, 0, None, b"fine") + zigzag_varint(90) + bytes(10), 2)
    out = {
        "comment": "made by tests/golden/make_kafka_golden.py (explicit field packing per the Kafka protocol guide, bitwise CRC-32C)",
        "segment_hex": (b0 + ctl + b2).hex(),
        "values_hex": [None if v is None else v.hex() for v in values] + [None] + ["after-the-marker".encode().hex(), b"last".hex()],
        "batch_status": [0, 5, 0],
        "record_counts": [len(values), 1, 2],
        "batch0_crc": struct.unpack_from(">I", b0, 17)[0],
        "hostile_hex": {"key_length_2^62": evil_key.hex(), "value_length_2^62": evil_val.hex(), "key_length_-5": evil_neg.hex(),
                        "second_record_truncated": trunc.hex()},
    }
    p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kafka_v2_batches.json")
    with open(p, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(p, len(b0 + ctl + b2), "bytes")


if __name__ == "__main__":
    main()
