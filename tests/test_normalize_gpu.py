"""normalize_fn (cardata-v1.py:40-131) on the GPU: the stand-alone operator b2_normalize — which runs the very device function
the fused score / decode / fit kernels apply — against the reference's float32 arithmetic evaluated by NumPy on the CPU
(true IEEE division), bit for bit; and the device-side exhaustive comparison of that function with IEEE division over ALL 2^32
float32 inputs for every feature of the reference's table."""
import ctypes as C

import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import _ffi
from oracle import cpu_oracle as orc

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _special_rows(d):
    v = np.array([0.0, -0.0, 1.0, -1.0, 15.0, 40.0, 20.0, 35.0, 7.0, 7500.0, 1000.0, 2000.0, 3.4028235e38, -3.4028235e38, 1e-45, -1e-45,
                  1.1754944e-38, np.inf, -np.inf, 2.5e38, 1e30, 16777216.0, 16777217.0, 0.1, 1 / 3, 27.5, 27.500002, 99.99999], dtype=np.float32)
    return np.repeat(v[:, None], d, axis=1)


def test_normalize_is_bit_exact_with_the_reference_order(cuda_device):
    kind, lo, rng = bs.schemas.norm_arrays()
    g = np.random.default_rng(6)
    x = np.concatenate([bs.synth.sensor_rows(1 << 20, seed=6), _special_rows(18),
                        g.integers(0, 1 << 32, size=(1 << 18, 18), dtype=np.uint64).astype(np.uint32).view(np.float32)])  # any bit pattern
    got = bs.keras.normalize_fn(x, kind, lo, rng)
    with np.errstate(all="ignore"):
        want = orc.normalize(x, kind, lo, rng, dtype=np.float32)
    nan = np.isnan(want)
    assert (np.isnan(got) == nan).all()
    np.testing.assert_array_equal(_bits(got)[~nan], _bits(want)[~nan])
    assert (got[:, [0, 2, 4, 5]] == 0).all()   # the four features forced to 0.0 (:72,78,84,87)


def test_normalize_odd_width_and_identity(cuda_device):
    kind = np.array([2, 0, 1, 2, 2], dtype=np.int32)
    lo = np.array([1.5, 0, 0, -3.0, 0.25], dtype=np.float32)
    rng = np.array([0.3, 1, 1, 7.0, 1e-3], dtype=np.float32)
    x = np.random.default_rng(1).normal(size=(1001, 5)).astype(np.float32) * 50
    got = bs.keras.normalize_fn(x, kind, lo, rng)
    np.testing.assert_array_equal(_bits(got), _bits(orc.normalize(x, kind, lo, rng, dtype=np.float32)))


def test_division_free_form_equals_ieee_division_for_every_float32(cuda_device):
    """All 2^32 bit patterns x all 18 features of the reference's table, plus a table of awkward ranges (tiny, huge, non
    power-of-two, reciprocal inexact; all inside the documented domain 2^-64 <= |range| <= 2^64): zero
    differences between norm_pair_fast and sub -> IEEE div -> fma on the device."""
    l, ctx = _ffi.lib(), bs.keras.get_ctx(0)
    tables = [bs.schemas.norm_arrays(),
              (np.array([2, 2, 2, 2, 2, 2, 2, 2, 0, 1], np.int32), np.array([0, -1e18, 3, 1e-18, 0.1, -7, 123456.7, 0, 5, 5], np.float32),
               np.array([3.0, 1e18, 1e-18, 3e-18, 0.7, 1.0000001, 3.3333333e7, 1.8e19, 9, 9], np.float32))]  # + identity, zero kinds
    for kind, lo, rng in tables:
        nm = _ffi.make_norm(kind, lo, rng)
        out = torch.zeros(2, dtype=torch.int64, device="cuda")
        for first in range(0, 1 << 32, 1 << 30):
            _ffi.check(l.b2_selftest_normalize(ctx, C.byref(nm), len(kind), first, 1 << 30, out.data_ptr(), None), "b2_selftest_normalize")
        torch.cuda.synchronize()
        bad, where = (int(v) for v in out.cpu())
        assert bad == 0, f"{bad} differences, e.g. feature {where >> 32} at bits {where & 0xFFFFFFFF:#010x}"
