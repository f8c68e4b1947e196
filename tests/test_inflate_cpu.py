"""The decompressors behind compressed Kafka record batches (csrc/inflate.cuh: gzip / Snappy / LZ4 frame, host+device code) built
for the host and checked against independent compressors: zlib / gzip from the Python standard library and pyarrow's Snappy and LZ4
codecs — round trips on text-like, random and highly repetitive data, the size-only pass, truncated and corrupted input."""
import ctypes as C
import gzip
import os
import struct
import subprocess
import zlib

import numpy as np
import pyarrow as pa
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GZIP, SNAPPY, LZ4, ZSTD = 1, 2, 3, 4


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("inflate") / "libinflate_host.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tests", "helpers", "inflate_host.cpp")], check=True)
    l = C.CDLL(so)
    l.inflate_codec.restype = C.c_longlong
    l.inflate_codec.argtypes = [C.c_int, C.c_char_p, C.c_longlong, C.c_void_p, C.c_longlong]
    return l


def _inflate(lib, codec, z: bytes, cap=None):
    n = lib.inflate_codec(codec, z, len(z), None, 0)      # size pass
    if n < 0:
        return n, None
    buf = C.create_string_buffer(max(n if cap is None else cap, 1))
    m = lib.inflate_codec(codec, z, len(z), buf, n if cap is None else cap)
    return m, buf.raw[:max(m, 0)]


def _xerial(blocks):
    out = b"\x82SNAPPY\x00" + struct.pack(">ii", 1, 1)
    for b in blocks:
        z = pa.compress(b, codec="snappy", asbytes=True)
        out += struct.pack(">i", len(z)) + z
    return out


def _zstd_stream(raw, piece=50_000):
    """The streaming API's frames (window descriptor, no content size) — what the Java client's ZstdOutputStream writes."""
    sink = pa.BufferOutputStream()
    with pa.CompressedOutputStream(sink, "zstd") as f:
        for i in range(0, len(raw), piece):
            f.write(raw[i:i + piece])
    return sink.getvalue().to_pybytes()


def _samples():
    g = np.random.default_rng(7)
    csv = b"".join(b"1567606196,car%d,%.7f,%.5f,%d,1000\n" % (i % 100, g.random(), g.random() * 40, int(g.integers(20, 35))) for i in range(3000))
    return {"empty": b"", "one": b"x", "short": b"hello hello hello hello", "csv": csv, "random": g.bytes(70_000),
            "big": b"".join(b"%d,%f\n" % (i, g.random()) for i in range(40_000)),   # several 128 KB zstd blocks: tables and offsets carried over
            "text": (b"the quick brown fox jumps over the lazy dog. " * 50 + bytes(g.integers(97, 123, size=5000, dtype=np.uint8))) * 8,
            "zeros": bytes(200_000), "runs": (b"ab" * 5000 + b"c" * 70_000 + bytes(range(256)) * 40),
            "avro": b"".join(b"\x00\x00\x00\x00\x01" + g.bytes(52) + b"\x3c\x3c\x3c\x3c\xd0\x0f" for _ in range(2000))}


@pytest.mark.parametrize("name", list(_samples()))
def test_round_trips_against_independent_compressors(lib, name):
    raw = _samples()[name]
    cases = [(GZIP, gzip.compress(raw, 6)), (GZIP, gzip.compress(raw, 1)), (GZIP, gzip.compress(raw, 0)),   # dynamic / fixed / stored blocks
             (SNAPPY, pa.compress(raw, codec="snappy", asbytes=True)), (SNAPPY, _xerial([raw[i:i + 32768] for i in range(0, len(raw), 32768)])),
             (LZ4, pa.compress(raw, codec="lz4", asbytes=True))]
    for level in (-5, 1, 3, 9, 19):   # raw / RLE / Huffman literals (1 and 4 streams, treeless), predefined / RLE / described / repeated FSE tables
        cases.append((ZSTD, pa.Codec("zstd", compression_level=level).compress(raw, asbytes=True)))
    cases.append((ZSTD, _zstd_stream(raw)))
    co = zlib.compressobj(9, zlib.DEFLATED, 31, 9, zlib.Z_FIXED)   # gzip container, fixed Huffman codes only
    cases.append((GZIP, co.compress(raw) + co.flush()))
    for codec, z in cases:
        n, out = _inflate(lib, codec, z)
        assert n == len(raw) and out == raw, (name, codec, n)


def test_truncated_corrupt_and_overflowing_input_is_refused(lib):
    raw = _samples()["csv"]
    for codec, z in ((GZIP, gzip.compress(raw)), (SNAPPY, pa.compress(raw, codec="snappy", asbytes=True)), (LZ4, pa.compress(raw, codec="lz4", asbytes=True)),
                     (ZSTD, pa.compress(raw, codec="zstd", asbytes=True)), (ZSTD, _zstd_stream(raw, 20_000))):
        for cut in (1, 5, len(z) // 2, len(z) - 1):
            assert lib.inflate_codec(codec, z[:cut], cut, None, 0) < 0, (codec, cut)
        n, _ = _inflate(lib, codec, z, cap=len(raw) - 1)       # output buffer one byte short
        assert n == -3
        g = np.random.default_rng(codec)
        for _ in range(300):                                     # random damage: any answer but a crash / an over-long write
            b = bytearray(z)
            for _ in range(int(g.integers(1, 4))):
                b[int(g.integers(0, len(b)))] = int(g.integers(0, 256))
            buf = C.create_string_buffer(len(raw) + 64)
            m = lib.inflate_codec(codec, bytes(b), len(b), buf, len(raw))
            assert m <= len(raw) and buf.raw[len(raw):] == bytes(64)
    assert lib.inflate_codec(5, b"\x28\xb5\x2f\xfd", 4, None, 0) == -4   # no codec 5
    # zstd: concatenated frames with a skippable frame in between; a dictionary id is refused as unsupported
    z1, z2 = pa.compress(raw[:1000], codec="zstd", asbytes=True), pa.Codec("zstd", compression_level=19).compress(raw[1000:], asbytes=True)
    n, out = _inflate(lib, ZSTD, z1 + struct.pack("<II", 0x184D2A53, 5) + b"abcde" + z2)
    assert n == len(raw) and out == raw
    with_dict = bytearray(z1)
    with_dict[4] |= 1            # Dictionary_ID_flag = 1 → one more header byte
    with_dict.insert(5 + (0 if with_dict[4] & 0x20 else 1), 7)
    assert lib.inflate_codec(ZSTD, bytes(with_dict), len(with_dict), None, 0) == -4
    # a snappy copy that reaches before the start of the output, an LZ4 offset of zero
    assert lib.inflate_codec(SNAPPY, bytes([4, 0x01 | (0 << 2), 9]), 3, None, 0) < 0
    assert lib.inflate_codec(LZ4, bytes.fromhex("04224d186040820500000010610000" "00000000"), 19, None, 0) < 0
