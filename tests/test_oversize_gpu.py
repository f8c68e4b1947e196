"""A record far longer than the batch average (the size hint sizes each warp's stage) is still decoded: the main launch hands
it to the retry pass with the 24 KB stage; only records beyond 24 KB are refused (B2_REC_TOO_LONG = 7)."""
import json

import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import decode as dec, schemas, synth

pytestmark = pytest.mark.gpu


def _csv_lines(n, seed):
    rows = synth.sensor_rows(n, seed=seed)
    d, o = synth.csv_batch(rows, seed=seed)
    raw = d.tobytes()
    return [raw[o[i]:o[i + 1]] for i in range(n)], rows


def _model18():
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    h = k.layers.Dense(14, activation="tanh")(inp)
    h = k.layers.Dense(7, activation="relu")(h)
    h = k.layers.Dense(7, activation="tanh")(h)
    m = k.models.Model(inputs=inp, outputs=k.layers.Dense(18, activation="relu")(h))
    m.set_normalization(*schemas.norm_arrays())
    return m


@pytest.mark.parametrize("where", [0, 5000, 9999])
def test_csv_one_8kb_record_among_100_byte_records(cuda_device, where):
    lines, rows = _csv_lines(10_000, seed=3)
    # same numbers, but the (skipped) car-name field is 8 KB long, quoted, with a delimiter and an escaped quote inside
    f = lines[where].split(b",")
    big = b",".join([f[0], b'"' + b"x" * 4000 + b',""' + b"y" * 4150 + b'"'] + f[2:])
    huge = b",".join([f[0], b"z" * 30_000] + f[2:])          # beyond the 24 KB stage: refused
    lines2 = list(lines)
    lines2[where] = big
    other = (where + 77) % 10_000
    lines2[other] = huge
    rb = dec.RecordBatch.from_messages(lines2)
    kinds = [0, 0] + [1] * 18
    out, _, status = dec.decode_csv_matrix(rb, kinds, raise_on_error=False)
    ref = dec.decode_csv_matrix(dec.RecordBatch.from_messages(lines), kinds)[0]
    st = status.cpu().numpy()
    assert st[where] == 0 and st[other] == 7 and (np.delete(st, other) == 0).all()
    keep = np.arange(10_000) != other
    assert torch.equal(out[torch.from_numpy(keep).cuda()], ref[torch.from_numpy(keep).cuda()])
    assert torch.isnan(out[other]).all()
    # fused path: same scores as without the long field, counters add up (9 999 ok, 1 refused)
    m = _model18()
    cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
    mse, flag, _, st2 = dec.decode_score(m, rb, csv_kinds=kinds, counters=cnt)
    mse_ref = dec.decode_score(m, dec.RecordBatch.from_messages(lines), csv_kinds=kinds)[0]
    k = torch.from_numpy(keep).cuda()
    assert torch.equal(mse[k], mse_ref[k]) and bool(torch.isnan(mse[other]))
    assert st2.cpu().numpy()[other] == 7 and cnt.cpu().tolist()[:3] == [9_999, int(flag.sum()), 1]
    # and again on the same stream: the retry list was left empty
    out2, _, status2 = dec.decode_csv_matrix(rb, kinds, raise_on_error=False)
    assert torch.equal(status2, status) and torch.equal(torch.nan_to_num(out2), torch.nan_to_num(out))


def test_json_long_string_member(cuda_device):
    rows = synth.sensor_rows(4096, seed=9)
    d, o = synth.json_batch(rows)
    raw = d.tobytes()
    msgs = [raw[o[i]:o[i + 1]] for i in range(4096)]
    obj = json.loads(msgs[100])
    obj["comment"] = "n" * 12_000
    msgs2 = list(msgs)
    msgs2[100] = json.dumps(obj, separators=(",", ":")).encode()
    keys = list(schemas.FEATURE_NAMES)
    a = dec.decode_json_matrix(dec.RecordBatch.from_messages(msgs2), keys, raise_on_error=False)
    b = dec.decode_json_matrix(dec.RecordBatch.from_messages(msgs), keys)
    assert not a[2].any() and torch.equal(a[0], b[0])


def test_many_oversize_records_and_every_stream_gets_its_own_list(cuda_device):
    lines, _ = _csv_lines(3000, seed=4)
    g = np.random.default_rng(4)
    idx = sorted(g.choice(3000, size=300, replace=False).tolist())
    lines2 = list(lines)
    for i in idx:
        f = lines2[i].split(b",")
        lines2[i] = b",".join([f[0], b"w" * int(g.integers(6000, 20_000))] + f[2:])
    kinds = [0, 0] + [1] * 18
    ref = dec.decode_csv_matrix(dec.RecordBatch.from_messages(lines), kinds)[0]
    rb = dec.RecordBatch.from_messages(lines2)
    rb.avg_record_bytes = 170  # what a consumer would announce from its running average; the long ones are the exception
    outs = []
    streams = [torch.cuda.Stream() for _ in range(3)]
    torch.cuda.synchronize()
    for s in streams:
        with torch.cuda.stream(s):
            outs.append(dec.decode_csv_matrix(rb, kinds, raise_on_error=False))
    torch.cuda.synchronize()
    for out, _, status in outs:
        assert not status.any() and torch.equal(out, ref)
