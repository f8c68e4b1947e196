"""Result sink: the row formatter the CUDA kernel runs (csrc/float_format.cuh, host+device code) built for
the host and compared with np.array2string — the call the reference's OutputCallback makes per row (cardata-v1.py:222) — on
the committed golden rows (text recorded from NumPy 2.3) and on fresh random rows against the installed NumPy."""
import json
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle import cpu_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NPY = tuple(int(v) for v in np.__version__.split(".")[:2])


@pytest.fixture(scope="module")
def host_fmt(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("fmt") / "format_rows_host")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "helpers", "format_rows_host.cpp")], check=True)
    tmp = tmp_path_factory.mktemp("fmt_io")

    def run(x, compat):
        x = np.ascontiguousarray(x, np.float32)
        fi, fo = str(tmp / "in.bin"), str(tmp / "out.bin")
        with open(fi, "wb") as f:
            f.write(struct.pack("ii", *x.shape))
            f.write(x.tobytes())
        subprocess.run([exe, fi, fo, str(compat)], check=True)
        b = open(fo, "rb").read()
        out, p = [], 0
        for _ in range(x.shape[0]):
            (n,) = struct.unpack_from("i", b, p)
            out.append(b[p + 4:p + 4 + n].decode("ascii"))
            p += 4 + n
        return out

    return run


def test_golden_rows_all_numpy_generations(host_fmt, golden_dir):
    doc = json.load(open(os.path.join(golden_dir, "array2string_rows.json")))
    assert len(doc["rows"]) >= 200
    for compat, key in ((0, "np23"), (1, "np121"), (2, "np114")):
        for i, r in enumerate(doc["rows"]):
            x = np.array(r["bits"], dtype=np.uint32).view(np.float32)[None, :]
            assert host_fmt(x, compat)[0] == r[key], f"row {i} compat {compat}"


def _random_rows(seed):
    g = np.random.default_rng(seed)
    yield np.maximum(g.normal(0.3, 0.5, (1500, 18)), 0).astype(np.float32)
    yield g.uniform(-1, 1, (1500, 18)).astype(np.float32)
    yield (g.uniform(-1, 1, (1500, 18)) * 10.0 ** g.integers(-6, 9, (1500, 18))).astype(np.float32)
    yield (g.uniform(-1, 1, (1500, 5)) * 10.0 ** g.integers(-3, 3, (1500, 1))).astype(np.float32)
    yield g.integers(0, 2 ** 32, (1500, 7), dtype=np.uint64).astype(np.uint32).view(np.float32)
    yield g.integers(2 ** 24, 10 ** 8, (800, 6)).astype(np.float32)
    yield np.ldexp(1.0, g.integers(-149, 128, (800, 6))).astype(np.float32)
    x = g.uniform(0, 2, (500, 30)).astype(np.float32)
    x[g.random(x.shape) < 0.1] = np.nan
    x[g.random(x.shape) < 0.05] = -np.inf
    yield x
    yield g.uniform(0, 1, (100, 64)).astype(np.float32)
    yield g.uniform(0, 1, (300, 1)).astype(np.float32)


@pytest.mark.parametrize("compat", [0, 1, 2])
def test_random_rows_against_installed_numpy(host_fmt, compat):
    if compat < 2 and NPY < (2, 3):
        pytest.skip("the library oracle for this generation needs NumPy >= 2.3 (printoptions legacy='2.2')")
    for x in _random_rows(1000 + compat):
        got, ref = host_fmt(x, compat), orc.array2string_rows(x, compat)
        bad = [i for i, (a, b) in enumerate(zip(got, ref)) if a != b]
        assert not bad, f"compat {compat}: row {bad[0]}\n got {got[bad[0]]!r}\n ref {ref[bad[0]]!r}"


def test_every_length_stays_inside_the_bound(host_fmt):
    # widest elements: negative, 8 fraction digits, scientific → 15 characters; the harness itself exits 3 beyond the bound
    x = np.full((4, 64), -1.2345678e-30, np.float32)
    x[1] = -12345678.0
    x[2] = np.float32(-0.12345678)
    x[3, ::2] = np.nan
    for c in (0, 1, 2):
        assert all(len(s) <= 64 * 20 + 2 for s in host_fmt(x, c))
