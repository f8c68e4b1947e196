"""JSON wire format: the record walk the CUDA kernel runs (csrc/json_walk.cuh, host+device code) built for the
host and compared with Python's json module + the float32 cast — on the reference's own fixture (the 10 simulator messages of
LSTM-…/cardata-v1.json), random objects, mutated (malformed) objects and grammar corner cases."""
import os
import struct
import subprocess

import numpy as np
import pytest

import b200_sensor as bs
from oracle import cpu_oracle as orc
from json_objects import rnd_objects as _rnd_objects

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = list(bs.schemas.FEATURE_NAMES)
LABEL = "failure_occurred"


@pytest.fixture(scope="module")
def walk(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("json") / "json_walk_host")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "helpers", "json_walk_host.cpp")], check=True)
    tmp = tmp_path_factory.mktemp("json_io")

    def run(recs, keys=KEYS, label=LABEL):
        names = list(keys) + ([label] if label else [])
        fs, fr, fo = (str(tmp / n) for n in ("spec.bin", "recs.bin", "out.bin"))
        with open(fs, "wb") as f:
            f.write(struct.pack("ii", len(names), len(keys)))
            for j, k in enumerate(names):
                kb = k.encode("latin-1")
                f.write(struct.pack("bB", j if j < len(keys) else -2, len(kb)))
                f.write(kb.ljust(32, b"\0"))
        offs = np.zeros(len(recs) + 1, np.int32)
        np.cumsum([len(r) for r in recs], out=offs[1:])
        data = b"".join(recs)
        with open(fr, "wb") as f:
            f.write(struct.pack("i", len(recs)))
            f.write(offs.tobytes())
            f.write(data)
        subprocess.run([exe, fs, fr, fo], check=True)
        b = open(fo, "rb").read()
        n, d = len(recs), len(keys)
        out = np.frombuffer(b[:4 * n * d], np.float32).reshape(n, d)
        st = np.frombuffer(b[4 * n * d:4 * n * d + 4 * n], np.int32)
        lab = np.frombuffer(b[4 * n * d + 4 * n:], np.uint8)
        ro, rl, rs = orc.decode_json(np.frombuffer(data, np.uint8), offs, keys, label)
        np.testing.assert_array_equal(st, rs)
        ok = rs != orc.REC_BAD_JSON
        np.testing.assert_array_equal(out.view(np.uint32)[ok], ro.view(np.uint32)[ok])
        np.testing.assert_array_equal(lab[ok], rl[ok])
        return out, lab, st

    return run


def test_reference_fixture(walk, golden_dir):
    recs = [l.strip().encode() for l in open(os.path.join(golden_dir, "cardata_v1_sample.json")) if l.strip()]
    assert len(recs) == 10
    out, lab, st = walk(recs)
    assert not st.any() and (lab == 2).all()
    # first message of the fixture, as printed there
    np.testing.assert_array_equal(out[0, :4], np.array([262.26807, 18.780598, 15.997148, 35.11709], np.float32))
    assert out[0, 9] == 33 and out[0, 17] == 2000


def test_random_and_mutated_objects(walk):
    recs, g = _rnd_objects(7, 6000)
    _, _, st = walk(recs)
    assert (st == 0).sum() > 300 and (st == orc.REC_BAD_NUMBER).sum() > 1000  # random values often put a string where a number is wanted

    def mutate(r):
        r = bytearray(r)
        if not r:
            return bytes(r)
        t, p = g.integers(0, 5), int(g.integers(0, len(r)))
        if t == 0:
            del r[p]
        elif t == 1:
            r[p] = int(g.integers(32, 127))
        elif t == 2:
            r.insert(p, int(g.integers(32, 127)))
        elif t == 3:
            r = r[:p]
        else:
            r[p:p] = b",,"
        return bytes(r)

    _, _, st = walk([mutate(r) for r in recs])
    assert (st == orc.REC_BAD_JSON).sum() > 1000


def test_grammar_corner_cases(walk):
    walk([b"", b"{}", b" { } ", b"[]", b"3", b'{"speed":1}x', b'{"speed":01}', b'{"speed":1.}', b'{"speed":.5}', b'{"speed":+1}',
          b'{"speed":1e5}', b'{"speed":1E+5}', b'{"speed":NaN}', b'{"speed":Infinity}', b'{"speed":"1"}', b'{"speed":true}',
          b'{"speed":null}', b'{"speed":1,"speed":2}', b'{"speed":"x","speed":2}', b'{"speed":1,}', b'{,"speed":1}', b'{"speed" 1}',
          b'{"speed":[1,2]}', b'{"a":[1,,2]}', b'{"a":{"b":[{"c":1}]},"speed":3}', b'{"a":"\\x"}', b'{"a":"\\u12G4"}',
          b'{"a":"tab\there"}', b'{"failure_occurred":"true"}', b'{"failure_occurred":false}', b'{"failure_occurred":"maybe"}',
          b'{"failure_occurred":null}', b'{"failure_occurred":""}', b'{"failure_occurred":7}', b'{"speed":1e999}',
          b'{"speed":-1e-999}', b'{"speed":123456789012345678901234567890}', b'{"speed":0.1234567890123456789012345}',
          b'{"sp\\u0065ed":5}', b'{"speed":1}\n', b'\t{"speed":\r\n1}', b'{"speed":1} {"speed":2}',
          b'{"coolant_temp":1,"coolant_temq":2}', b'{"coolant_temp_":1}', b'{"' + b"k" * 40 + b'":1,"speed":2}',
          b'{"a":' + b"[" * 60 + b"]" * 60 + b"}"])  # (containers nested deeper than 64 levels are refused: documented limit)


def test_synthetic_batches_round_trip(walk):
    rows = bs.synth.sensor_rows(2000, seed=5)
    lab = (np.arange(2000) % 7 == 0).astype(np.int64)
    for dup in (False, True):
        d, o = bs.synth.json_batch(rows, labels=lab, camel_duplicates=dup)
        recs = [bytes(d[o[i]:o[i + 1]]) for i in range(len(o) - 1)]
        out, l2, st = walk(recs)
        assert not st.any()
        want = rows.copy()
        want[:, [9, 10, 11, 12, 17]] = np.rint(want[:, [9, 10, 11, 12, 17]])
        np.testing.assert_array_equal(out, want)  # Float.toString text → double → float32 gives the float32 back
        np.testing.assert_array_equal(l2, lab)


def test_mutated_in_order_records_fast_prefix_hand_over(walk):
    """Records in spec order take the fast prefix; a mutation anywhere must hand over to the careful walk with the same verdict
    and values as Python's json (also: numbers in forms the SWAR parser declines — '+' exponents, long mantissas, blanks)."""
    g = np.random.default_rng(23)
    rows = bs.synth.sensor_rows(3000, seed=6)
    rows[::7, 3] *= 1e-6   # E-notation values
    rows[::11, 5] *= 1e9
    d, o = bs.synth.json_batch(rows, labels=(np.arange(3000) % 3 == 0).astype(np.int64))
    recs = [bytearray(d[o[i]:o[i + 1]].tobytes()) for i in range(3000)]
    out = []
    for r in recs:
        t, p = g.integers(0, 8), int(g.integers(0, len(r)))
        if t == 0:
            del r[p]
        elif t == 1:
            r[p] = int(g.integers(32, 127))
        elif t == 2:
            r.insert(p, int(g.integers(32, 127)))
        elif t == 3:
            r = r[:p]
        elif t == 4:
            r = bytearray(bytes(r).replace(b":", b": ", 3))
        elif t == 5:
            r = bytearray(bytes(r).replace(b'"speed":', b'"speed":1234567890.12345678901234567890e+2,"x":', 1))
        elif t == 6:
            r = bytearray(bytes(r).replace(b'"failure_occurred":"', b'"failure_occurred":"un', 1))
        out.append(bytes(r))
    walk(out)
