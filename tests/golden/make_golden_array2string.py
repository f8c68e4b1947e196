"""Regenerate tests/golden/array2string_rows.json: float32 rows (as uint32 bit patterns) and the text NumPy's
np.array2string gives for each — the library call the reference's OutputCallback makes (cardata-v1.py:222).

    python tests/golden/make_golden_array2string.py        # needs NumPy >= 2.3 (records its version in the file)

"np23" is plain np.array2string of the installed NumPy, "np121" the same under printoptions(legacy='2.2') (NumPy 1.21-2.2
text), "np114" the restatement of NumPy 1.14-1.20 in oracle/cpu_oracle.py (the generation of the reference's TF 2.0 images).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import cpu_oracle as orc  # noqa: E402


def rows():
    g = np.random.default_rng(314)
    out = []
    out += list(np.maximum(g.normal(0.3, 0.5, (60, 18)), 0).astype(np.float32))          # relu-like reconstruction rows
    out += list(np.maximum(g.normal(0.0, 0.8, (30, 30)), 0).astype(np.float32))          # 30-wide model
    out += list((g.uniform(-1, 1, (40, 18)) * 10.0 ** g.integers(-6, 9, (40, 18))).astype(np.float32))
    out += list(g.integers(0, 2 ** 32, (40, 7), dtype=np.uint64).astype(np.uint32).view(np.float32))  # any bit pattern
    out += list(g.integers(2 ** 24, 10 ** 8, (10, 6)).astype(np.float32))
    out += list(g.integers(-3000, 3000, (10, 18)).astype(np.float32))
    x = g.uniform(0, 2, (20, 18)).astype(np.float32)
    x[g.random((20, 18)) < 0.1] = np.nan
    x[g.random((20, 18)) < 0.05] = -np.inf
    x[g.random((20, 18)) < 0.05] = np.inf
    out += list(x)
    out += [np.zeros(18, np.float32), np.array([0.0, -0.0, 1.0], np.float32), np.array([np.nan], np.float32),
            np.array([np.inf, -np.inf], np.float32), np.array([0.5], np.float32)]
    for pair in ([0.0001, 0.1], [9.9999997e-05, 0.05], [1e8, 1], [99999992, 1e5], [1, 1000], [1, 1000.0001], [1e-45, 1e-45],
                 [3.4028235e38, 1], [2e6, 1e4], [999999.94, 1000.0], [1e6, 1e3], [16777216, 16777218], [0.99999994, 0.5],
                 [9.9999994e-05, 9.9999994e-05], [123456.79, 123.0]):
        out.append(np.array(pair, np.float32))
    out += list(g.uniform(0, 1, (5, 64)).astype(np.float32))
    return out


def main():
    assert tuple(int(v) for v in np.__version__.split(".")[:2]) >= (2, 3), "run with NumPy >= 2.3"
    rs = rows()
    doc = {"numpy_version": np.__version__, "rows": []}
    for r in rs:
        r2 = r[None, :]
        doc["rows"].append({"bits": r.view(np.uint32).tolist(), "np23": orc.array2string_rows(r2, 0)[0],
                            "np121": orc.array2string_rows(r2, 1)[0], "np114": orc.array2string_rows(r2, 2)[0]})
    with open(os.path.join(HERE, "array2string_rows.json"), "w") as f:
        json.dump(doc, f, indent=0)
    print(len(rs), "rows written")


if __name__ == "__main__":
    main()
