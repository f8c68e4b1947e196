"""JSON records on the GPU (b2_decode_json / b2_decode_score_json through the C ABI) against Python's json + float32 cast."""
import os

import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import decode as dec, synth
from oracle import cpu_oracle as orc

pytestmark = pytest.mark.gpu
KEYS = list(bs.schemas.FEATURE_NAMES)
LABEL = "failure_occurred"


def _check(recs, keys=KEYS, label=LABEL):
    rb = dec.RecordBatch.from_messages(recs)
    out, lab, st = dec.decode_json_matrix(rb, keys, label, raise_on_error=False)
    d, o = rb.data.cpu().numpy(), rb.offsets.cpu().numpy()
    ro, rl, rs = orc.decode_json(d, o, keys, label)
    st = st.cpu().numpy()
    np.testing.assert_array_equal(st, rs)
    ok = rs != orc.REC_BAD_JSON
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32)[ok & (rs == 0)], ro.view(np.uint32)[ok & (rs == 0)])
    np.testing.assert_array_equal(lab.cpu().numpy()[rs == 0], rl[rs == 0])
    return out, lab, st


def test_reference_fixture(cuda_device, golden_dir):
    recs = [l.strip().encode() for l in open(os.path.join(golden_dir, "cardata_v1_sample.json")) if l.strip()]
    out, lab, st = _check(recs)
    assert not st.any()
    assert out[0, 0].item() == np.float32(262.26807) and out[0, 17].item() == 2000.0


def test_random_mutated_and_corner_cases(cuda_device):
    from json_objects import rnd_objects as _rnd_objects

    recs, g = _rnd_objects(11, 5000)
    _check(recs)
    cut = [r[:int(g.integers(0, len(r) + 1))] for r in recs[:2000]]
    _, _, st = _check(cut)
    assert (st == orc.REC_BAD_JSON).sum() > 1000
    _check([b"", b"{}", b"[]", b'{"speed":01}', b'{"speed":1e5}', b'{"speed":"1"}', b'{"speed":null}', b'{"speed":"x","speed":2}',
            b'{"a":[1,,2]}', b'{"a":{"b":[{"c":1}]},"speed":3}', b'{"sp\\u0065ed":5}', b'{"speed":1} {"speed":2}',
            b'{"speed":123456789012345678901234567890}', b'{"failure_occurred":"true"}', b'{"failure_occurred":false}'])


@pytest.mark.parametrize("dup", [False, True])
def test_large_synthetic_batch_and_fused_score(cuda_device, dup):
    rows = synth.sensor_rows(100_003, seed=9)
    labels = (np.arange(rows.shape[0]) % 50 == 0).astype(np.int64)
    d, o = synth.json_batch(rows, labels=labels, camel_duplicates=dup)
    rb = dec.RecordBatch(d, o)
    out, lab, st = dec.decode_json_matrix(rb, KEYS, LABEL)
    want = rows.copy()
    want[:, [9, 10, 11, 12, 17]] = np.rint(want[:, [9, 10, 11, 12, 17]])
    np.testing.assert_array_equal(out.cpu().numpy(), want)
    np.testing.assert_array_equal(lab.cpu().numpy(), labels)
    # fused decode + normalise + score == decode, then score (bit for bit), and the counters add up
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    h = k.layers.Dense(14, activation="tanh", activity_regularizer=k.regularizers.l1(1e-7))(inp)
    h = k.layers.Dense(7, activation="relu")(h)
    h = k.layers.Dense(7, activation="tanh")(h)
    m = k.models.Model(inputs=inp, outputs=k.layers.Dense(18, activation="relu")(h))
    m.set_normalization(*bs.schemas.norm_arrays())
    cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
    mse, flag, lab2, st2 = dec.decode_score(m, rb, json_keys=KEYS, json_label=LABEL, threshold=5.0, counters=cnt)
    mse2, flag2 = m.score_device(out)
    assert torch.equal(mse, mse2) and torch.equal(flag, flag2) and torch.equal(lab2, lab) and not st2.any()
    c = cnt.cpu().numpy()
    assert c[0] == rows.shape[0] and c[1] == int(flag.sum()) and c[2] == 0 and c[3] == int(o[-1])
