"""K2 parity: CSV / Avro decode and the fused decode→score kernel against the CPU oracle, through the C ABI.

Decoded floats must be BIT-exact (glibc strtof / IEEE casts are the oracle); statuses must match the oracle's per record;
fused decode+score must equal decode followed by score bit for bit, and the oracle's MSE within 1e-5 relative.
"""
import os

import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import avro_schema as av, decode as dec, schemas, synth
from oracle import c_oracle as co, cpu_oracle as orc

pytestmark = pytest.mark.gpu

ACTS = [orc.ACT_TANH, orc.ACT_RELU, orc.ACT_TANH, orc.ACT_RELU]
SENSOR_KINDS = [0, 0] + [1] * 18


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _model18(seed=314):
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    e = k.layers.Dense(14, activation="tanh", activity_regularizer=k.regularizers.l1(1e-7))(inp)
    e = k.layers.Dense(7, activation="relu")(e)
    d = k.layers.Dense(7, activation="tanh")(e)
    d = k.layers.Dense(18, activation="relu")(d)
    m = k.models.Model(inputs=inp, outputs=d, seed=seed)
    m.set_normalization(*schemas.norm_arrays())
    return m


def test_csv_reference_sample_bit_exact(cuda_device, golden_dir):
    lines = open(os.path.join(golden_dir, "car_sensor_sample.csv")).read().strip().split("\n")[1:]
    rb = dec.RecordBatch.from_messages(lines)
    out, _, status = dec.decode_csv_matrix(rb, SENSOR_KINDS)
    ref = np.array([[np.float32(float(t)) for t in ln.split(",")[2:]] for ln in lines], dtype=np.float32)
    assert int(status.abs().sum()) == 0
    np.testing.assert_array_equal(_bits(out.cpu().numpy()), _bits(ref))


def test_csv_every_reference_token_bit_exact(cuda_device, golden_dir):
    d = np.load(os.path.join(golden_dir, "csv_tokens.npz"))
    toks, bits = d["tokens"].tolist(), d["bits"]
    rb = dec.RecordBatch.from_messages(toks)
    out, _, status = dec.decode_csv_matrix(rb, [1])
    assert int(status.abs().sum()) == 0
    np.testing.assert_array_equal(_bits(out.cpu().numpy()).ravel(), bits)


@pytest.mark.parametrize("n", [1, 2, 255, 256, 257, 1000, 100_003])
def test_csv_synthetic_roundtrip_and_oracle(cuda_device, n):
    rows = synth.sensor_rows(n, seed=40 + n)
    data, offs = synth.csv_batch(rows, seed=n)
    out, _, status = dec.decode_csv_matrix(dec.RecordBatch(data, offs), SENSOR_KINDS)
    ref, _, st = co.decode_csv(data, offs, SENSOR_KINDS)
    assert st.sum() == 0 and int(status.abs().sum()) == 0
    np.testing.assert_array_equal(_bits(out.cpu().numpy()), _bits(ref))
    exp = rows.copy()
    exp[:, [9, 10, 11, 12, 17]] = np.rint(exp[:, [9, 10, 11, 12, 17]])
    np.testing.assert_array_equal(_bits(out.cpu().numpy()), _bits(exp))  # text → float32 round trip is lossless


def test_csv_number_grammar_fuzz_vs_strtof(cuda_device):
    g = np.random.default_rng(99)
    toks = []
    f32 = g.integers(0, 2 ** 32, size=60000, dtype=np.uint64).astype(np.uint32).view(np.float32)
    f32 = f32[np.isfinite(f32)]
    toks += ["%.9g" % v for v in f32[:20000]]
    toks += [repr(float(v)) for v in f32[20000:30000]]
    toks += ["%.17g" % ((float(a) + float(np.nextafter(a, np.float32(np.inf)))) / 2) for a in f32[30000:40000]]  # midpoints
    toks += ["%.6E" % v for v in f32[40000:45000]]
    toks += [" %s " % s for s in ["1.5", "-2", "+3e2", ".5", "5.", "1E-4", "inf", "-inf", "nan", "Infinity", "0", "-0.0"]]
    toks += ["%d.%dE%d" % (g.integers(0, 10 ** 9), g.integers(0, 10 ** 9), g.integers(-60, 40)) for _ in range(5000)]
    toks += ["1e39", "1e-46", "1.4e-45", "7.0064923216240854e-46", "7.0064923216240853e-46", "3.4028235677973366e38",
             "123456789012345678901234567890", "0.000000000000000000000000000000000000000000001"]
    toks = [t for t in toks if "," not in t]
    rb = dec.RecordBatch.from_messages(toks)
    out, _, status = dec.decode_csv_matrix(rb, [1], raise_on_error=False)
    got = _bits(out.cpu().numpy()).ravel()
    st = status.cpu().numpy()
    ref = np.array([np.float32(orc.strtof(t.encode())[0]) for t in toks], dtype=np.float32)
    ok = st == 0
    assert ok.mean() > 0.999  # only >19-digit ambiguous tokens may be refused (status 7), never wrong
    assert set(np.unique(st[~ok]).tolist()) <= {7}
    nan = np.isnan(ref)
    np.testing.assert_array_equal(got[ok & ~nan], _bits(ref)[ok & ~nan])
    assert np.isnan(out.cpu().numpy().ravel()[ok & nan]).all()


def test_csv_error_semantics_match_oracle(cuda_device):
    recs = [b"1,2,3", b"1,,3", b"1,x,3", b'1,"2",3', b'1,"2,3', b"1,2", b" 1 ,2.5e0,3", b'1,2"x,3', b"1,2,3,4", b"1,2,", b"",
            b",,", b'"1","2","3"', b'1,"",3', b"1,2\n,3", b"0x10,1,2", b"1e,2,3", b"1,2,3 "]
    data, offs = synth.lines_to_batch(recs)
    ref, _, st_ref = orc.decode_csv(data, offs, [1, 1, 1], defaults=[None, 7.0, None])
    out, _, st = dec.decode_csv_matrix(dec.RecordBatch(data, offs), [1, 1, 1], defaults=[None, 7.0, None], raise_on_error=False)
    np.testing.assert_array_equal(st.cpu().numpy(), st_ref)
    ok = st_ref == 0
    np.testing.assert_array_equal(_bits(out.cpu().numpy())[ok], _bits(ref)[ok])
    assert np.isnan(out.cpu().numpy()[~ok]).all()
    with pytest.raises(dec.DecodeError):
        dec.decode_csv(dec.RecordBatch(data, offs), [[0.0], [7.0], [0.0]])


def test_csv_credit_card_form_subtiles_and_strings(cuda_device):
    """31 fields (30 full-precision floats + quoted class label, ≈550 B/record → the 48 KB stage holds a sub-tile only)."""
    x = synth.credit_rows(5000, seed=3).astype(np.float64) * 1.000000123
    labels = (np.arange(5000) % 7 == 0).astype(np.int64)
    data, offs = synth.credit_csv_batch(x, labels)
    assert (offs[1:] - offs[:-1]).mean() > 400
    cols = dec.decode_csv(dec.RecordBatch(data, offs), [[0.0]] * 30 + [[""]])
    assert len(cols) == 31 and cols[30].dtype == object
    ref, _, st = co.decode_csv(data, offs, [1] * 30 + [2])
    assert st.sum() == 0
    np.testing.assert_array_equal(_bits(np.stack(cols[:30], 1)), _bits(ref))
    assert [int(c) for c in cols[30]] == labels.tolist()
    # tf.stack(entry[0:30], 1) + to_number(entry[30]) — process_x_y of the reference script
    np.testing.assert_array_equal(dec.to_number(cols[30], np.int32), labels.astype(np.int32))
    with pytest.raises(dec.DecodeError):
        dec.to_number([b"1", b"x"], np.int32)


def test_csv_malformed_injection(cuda_device):
    rows = synth.sensor_rows(50_000, seed=8)
    data, offs = synth.csv_batch(rows, seed=8)
    data, offs, bad = synth.corrupt_csv(data, offs, frac=2e-3, seed=8)
    out, _, status = dec.decode_csv_matrix(dec.RecordBatch(data, offs), SENSOR_KINDS, raise_on_error=False)
    ref, _, st_ref = co.decode_csv(data, offs, SENSOR_KINDS)
    np.testing.assert_array_equal(status.cpu().numpy(), st_ref)
    assert set(np.nonzero(st_ref)[0].tolist()) == set(bad.tolist())
    ok = st_ref == 0
    np.testing.assert_array_equal(_bits(out.cpu().numpy())[ok], _bits(ref)[ok])


def test_avro_both_schemas_match_oracle(cuda_device):
    x = synth.sensor_rows(70_001, seed=21)
    for which in ("a", "b"):
        if which == "a":
            data, offs = synth.avro_a_batch(x)
            labs = None
            sch = schemas.schema_a_json()
        else:
            data, offs, labs = synth.avro_b_batch(x, null_frac=0.03, seed=5)
            sch = schemas.schema_b_json()
        prog = av.compile_schema(sch)
        rb = dec.substr(dec.RecordBatch(data, offs), 5, -1)
        out, label, status, _ = dec.decode_avro_matrix(rb, prog)
        ref, lab_ref, st_ref = co.decode_avro(data, offs, [(f.type, f.nullable, f.out) for f in prog.fields], prog.n_out, skip_prefix=5)
        assert st_ref.sum() == 0 and int(status.abs().sum()) == 0
        np.testing.assert_array_equal(_bits(out.cpu().numpy()), _bits(ref))
        np.testing.assert_array_equal(label.cpu().numpy(), lab_ref)
        if labs is not None:
            np.testing.assert_array_equal(label.cpu().numpy(), labs)
    # the drop-in list-of-columns form with the dtype list of cardata-v3.py:56-74
    cols = dec.decode_avro(rb, sch, dtype=["float64"] * 9 + ["int32"] * 4 + ["float64"] * 4 + ["int32", "string"])
    assert len(cols) == 19 and cols[9].dtype == np.int32 and cols[0].dtype == np.float64 and cols[18][0] in (b"false", b"true", b"")


def test_avro_truncated_and_bad_branch(cuda_device):
    x = synth.sensor_rows(600, seed=2)
    data, offs, _ = synth.avro_b_batch(x, seed=2)
    prog = av.compile_schema(schemas.schema_b_json())
    fields = [(f.type, f.nullable, f.out) for f in prog.fields]
    data = data.copy()
    data[offs[7] + 5] = 0x06  # union branch 3
    offs2 = offs.copy()
    cut = offs2[-1] - 9
    data2, offs2[-1] = data[:cut], cut
    out, label, status, _ = dec.decode_avro_matrix(dec.RecordBatch(data2, offs2, skip_prefix=5), prog, raise_on_error=False)
    ref, lab_ref, st_ref = co.decode_avro(data2, offs2, fields, prog.n_out, skip_prefix=5)
    np.testing.assert_array_equal(status.cpu().numpy(), st_ref)
    assert st_ref[7] == orc.REC_BAD_BRANCH and st_ref[-1] == orc.REC_TRUNCATED
    ok = st_ref == 0
    np.testing.assert_array_equal(_bits(out.cpu().numpy())[ok], _bits(ref)[ok])


@pytest.mark.parametrize("fmt", ["csv", "avro_a", "avro_b"])
def test_fused_decode_score_equals_two_step_and_oracle(cuda_device, fmt):
    m = _model18()
    x = synth.sensor_rows(120_007, seed=77)
    if fmt == "csv":
        data, offs = synth.csv_batch(x, seed=1)
        data, offs, bad = synth.corrupt_csv(data, offs, frac=1e-3, seed=1)
        rb = dec.RecordBatch(data, offs)
        kw = dict(csv_kinds=SENSOR_KINDS)
        feats, _, st0 = dec.decode_csv_matrix(rb, SENSOR_KINDS, raise_on_error=False)
    else:
        if fmt == "avro_a":
            data, offs = synth.avro_a_batch(x)
            sch = schemas.schema_a_json()
        else:
            data, offs, _ = synth.avro_b_batch(x, null_frac=0.01, seed=4)
            sch = schemas.schema_b_json()
        rb = dec.RecordBatch(data, offs, skip_prefix=5)
        kw = dict(avro_schema=sch)
        feats, _, st0, _ = dec.decode_avro_matrix(rb, sch, raise_on_error=False)
    cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
    mse, flag, label, status = dec.decode_score(m, rb, threshold=5.0, counters=cnt, **kw)
    torch.testing.assert_close(status, st0, rtol=0, atol=0)
    ok = (st0 == 0)
    mse2, flag2 = m.score_device(feats[ok].contiguous())
    assert torch.equal(mse[ok], mse2) and torch.equal(flag[ok], flag2)  # fused == decode then score, bit for bit
    assert torch.isnan(mse[~ok]).all() and int(flag[~ok].sum()) == 0
    c = cnt.cpu().numpy()
    assert c[0] == int(ok.sum()) and c[1] == int(flag.sum()) and c[2] == int((~ok).sum()) and c[3] == int(offs[-1])
    # oracle: the whole CPU streaming loop
    w = m.get_flat_weights()
    nm = schemas.norm_arrays()
    if fmt == "csv":
        mr, fr, sr, _ = co.decode_score(data, offs, w, [18, 14, 7, 7, 18], ACTS, norm=nm, csv=(SENSOR_KINDS, None, ",", True))
    else:
        prog = av.compile_schema(sch)
        mr, fr, sr, _ = co.decode_score(data, offs, w, [18, 14, 7, 7, 18], ACTS, norm=nm,
                                        avro=[(f.type, f.nullable, f.out) for f in prog.fields], skip_prefix=5)
    okh = ok.cpu().numpy()
    np.testing.assert_array_equal(sr, st0.cpu().numpy())
    got = mse.cpu().numpy()[okh].astype(np.float64)
    ref = mr[okh].astype(np.float64)
    err = np.abs(got - ref)
    assert (err <= np.maximum(2e-5 * ref, 2e-7)).all()  # oracle.c is float32 itself: twice the 1e-5 budget
    safe = np.abs(ref - 5.0) > 1e-4
    np.testing.assert_array_equal(flag.cpu().numpy()[okh][safe], fr[okh][safe])


def test_empty_and_oversized_records(cuda_device):
    out, _, st = dec.decode_csv_matrix(dec.RecordBatch(np.zeros(0, np.uint8), np.zeros(1, np.int32)), [1, 1])
    assert out.shape == (0, 2)
    big = b"1," + b"9" * 60000  # one record larger than the 48 KB stage → refused with status 7, neighbours unaffected
    recs = [b"1,2", big, b"3,4"]
    out, _, st = dec.decode_csv_matrix(dec.RecordBatch.from_messages(recs), [1, 1], raise_on_error=False)
    assert st.cpu().numpy().tolist() == [0, 7, 0]
    assert out.cpu().numpy()[[0, 2]].tolist() == [[1.0, 2.0], [3.0, 4.0]]


@pytest.mark.parametrize("fmt", ["csv", "avro_a"])
def test_host_end_to_end_path_matches_device_path(cuda_device, fmt):
    """decode_score_host (pinned host buffers → chunked H2D overlapped with decode+score → D2H) == the device-resident call."""
    m = _model18()
    x = synth.sensor_rows(50_003, seed=123)
    if fmt == "csv":
        data, offs = synth.csv_batch(x, seed=3)
        kw, skip = dict(csv_kinds=SENSOR_KINDS), 0
    else:
        data, offs = synth.avro_a_batch(x)
        kw, skip = dict(avro_schema=schemas.schema_a_json()), 5
    mse_d, flag_d, _, _ = dec.decode_score(m, dec.RecordBatch(data, offs, skip_prefix=skip), **kw)
    hd, ho = torch.from_numpy(data).pin_memory(), torch.from_numpy(offs).pin_memory()
    for reuse, overlap in ((False, True), (True, True), (True, True), (True, False), (False, False)):
        # results copied back chunk by chunk on a third stream (default), or in one copy at the end
        mse_h, flag_h = dec.decode_score_host(m, hd, ho, skip_prefix=skip, reuse_output=reuse, overlap_d2h=overlap, **kw)
        np.testing.assert_array_equal(mse_h, mse_d.cpu().numpy())
        np.testing.assert_array_equal(flag_h, flag_d.cpu().numpy())
    mse_h, flag_h = dec.decode_score_host(m, data[: offs[7]], offs[:8], skip_prefix=skip, chunks=16, **kw)  # more chunks than records/4
    np.testing.assert_array_equal(mse_h, mse_d.cpu().numpy()[:7])


@pytest.mark.parametrize("fmt", ["avro_a", "avro_b"])
def test_reference_schema_walks_equal_the_generic_walk(cuda_device, fmt):
    """The straight-line walks for the reference's two schemas (avro_walk_schema_a / _b) against the run-interpreting generic
    walk (context option avro_generic = 1): bit-identical values, labels and statuses, including nulls, wide varints and damaged records."""
    x = synth.sensor_rows(50_003, seed=21)
    x[::97, 9] = 20000.0      # three-byte varint → the record leaves the straight-line walk
    x[::89, 17] = -3.0        # negative firmware
    if fmt == "avro_a":
        data, offs = synth.avro_a_batch(x)
        sch = schemas.schema_a_json()
    else:
        data, offs, _ = synth.avro_b_batch(x, null_frac=0.02, seed=5, label_probs=(0.9, 0.05, 0.05))
        sch = schemas.schema_b_json()
    data = data.copy()
    g = np.random.default_rng(3)
    for i in g.integers(0, len(offs) - 1, 200):       # damage some records: flip a byte / cut the record short
        b, e = int(offs[i]), int(offs[i + 1])
        data[b + 5 + int(g.integers(0, e - b - 5))] ^= 0xFF
    offs2 = offs.copy()
    offs2[-1] -= 7
    rb = dec.RecordBatch(data[:offs2[-1]], offs2, skip_prefix=5)
    res = {}
    try:
        for mode in ("0", "1"):
            bs.keras.set_option("avro_generic", int(mode))
            out, label, status, _ = dec.decode_avro_matrix(rb, sch, raise_on_error=False)
            res[mode] = (out.cpu().numpy().copy(), label.cpu().numpy().copy(), status.cpu().numpy().copy())
    finally:
        bs.keras.set_option("avro_generic", 0)
    np.testing.assert_array_equal(res["0"][2], res["1"][2])
    ok = res["1"][2] == 0
    assert ok.sum() > 40_000 and (~ok).sum() > 5
    np.testing.assert_array_equal(_bits(res["0"][0])[ok], _bits(res["1"][0])[ok])
    np.testing.assert_array_equal(res["0"][1][ok], res["1"][1][ok])


@pytest.mark.parametrize("fmt", ["csv", "avro_a", "json"])
def test_full_size_decode_score_properties(cuda_device, fmt):
    """At the benchmark's batch size (2^20 records + a ragged tail) parity is checked through size-independent properties: the
    batch scored in one launch, as two unequal pieces (different chunking, different stage alignment) and with a different
    record-size hint (different stage size, items split differently) gives bit-identical per-record results; counters add up;
    a sample is compared with the CPU oracle."""
    import ctypes as C

    m = _model18()
    n = (1 << 20) + 37
    x = synth.sensor_rows(n, seed=31)
    if fmt == "csv":
        data, offs = synth.csv_batch(x, seed=2)
        kw, skip = dict(csv_kinds=SENSOR_KINDS), 0
    elif fmt == "json":
        data, offs = synth.json_batch(x, labels=(np.arange(n) % 9 == 0))
        kw, skip = dict(json_keys=list(schemas.FEATURE_NAMES), json_label="failure_occurred"), 0
    else:
        data, offs = synth.avro_a_batch(x)
        kw, skip = dict(avro_schema=schemas.schema_a_json()), 5
    rb = dec.RecordBatch(data, offs, skip_prefix=skip)
    cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
    mse, flag, _, st = dec.decode_score(m, rb, counters=cnt, **kw)
    assert not st.any()
    c = cnt.cpu().numpy()
    assert c[0] == n and c[1] == int(flag.sum()) and c[2] == 0 and c[3] == int(offs[-1])
    # two unequal pieces
    h = n // 3 + 5
    d_dev, o_dev = rb.data, rb.offsets
    parts = []
    for lo, hi in ((0, h), (h, n)):
        b0 = int(offs[lo])
        sub = dec.RecordBatch(d_dev[b0:int(offs[hi])], (o_dev[lo:hi + 1] - b0).contiguous(), skip_prefix=skip)
        parts.append(dec.decode_score(m, sub, **kw)[:2])
    assert torch.equal(mse, torch.cat([p[0] for p in parts])) and torch.equal(flag, torch.cat([p[1] for p in parts]))
    # a much smaller stage (hint = a third of the real record size): every chunk is taken in several items
    rb2 = dec.RecordBatch(d_dev, o_dev, skip_prefix=skip)
    rb2.avg_record_bytes = max(8, int(offs[-1]) // n // 3)
    mse2, flag2, _, st2 = dec.decode_score(m, rb2, **kw)
    assert torch.equal(mse, mse2) and torch.equal(flag, flag2) and not st2.any()
    # sample against the oracle's scoring of the source rows
    idx = np.arange(0, n, 3001)
    xs = x[idx].copy()
    xs[:, [9, 10, 11, 12, 17]] = np.rint(xs[:, [9, 10, 11, 12, 17]])
    ref = orc.score(xs, m.get_flat_weights(), [18, 14, 7, 7, 18], ACTS, norm=schemas.norm_arrays(), dtype=np.float64)[0]
    got = mse.cpu().numpy()[idx].astype(np.float64)
    assert (np.abs(got - ref) <= np.maximum(1e-5 * ref, 1e-7)).all()


@pytest.mark.parametrize("fmt", ["avro_a", "avro_b"])
def test_two_records_per_lane_variant_is_bit_identical(cuda_device, fmt):
    """The row-packed two-records-per-lane form of the fused Avro kernel (context option decode_r2 = 2 / 3; not the default: it is
    slower on B200) must give the same bits as one record per lane — values, flags, labels, statuses, counters — on ragged
    batches with damaged records."""
    x = synth.sensor_rows(70_001, seed=33)
    x[::97, 9] = 20000.0
    if fmt == "avro_a":
        data, offs = synth.avro_a_batch(x)
        sch = schemas.schema_a_json()
    else:
        data, offs, _ = synth.avro_b_batch(x, null_frac=0.02, seed=5, label_probs=(0.9, 0.05, 0.05))
        sch = schemas.schema_b_json()
    data, offs, bad = synth.corrupt_batch(data, offs, fmt, 1e-3, seed=9)
    m = _model18()
    res = {}
    try:
        for mode in (0, 2, 3):
            bs.keras.set_option("decode_r2", mode)
            for n in (70_001, 63, 64, 65, 1):
                rb = dec.RecordBatch(data[: offs[n]], offs[: n + 1], skip_prefix=5)
                cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
                mse, flag, label, status = dec.decode_score(m, rb, avro_schema=sch, counters=cnt)
                res[(mode, n)] = (_bits(mse.cpu().numpy()), flag.cpu().numpy(), label.cpu().numpy(), status.cpu().numpy(), cnt.cpu().tolist())
    finally:
        bs.keras.set_option("decode_r2", 0)
    assert (res[(0, 70_001)][3] != 0).sum() == len(bad)
    for mode in (2, 3):
        for n in (70_001, 63, 64, 65, 1):
            for a, b in zip(res[(0, n)][:4], res[(mode, n)][:4]):
                np.testing.assert_array_equal(a, b)
            assert res[(0, n)][4] == res[(mode, n)][4]


@pytest.mark.parametrize("fmt", ["avro_a", "csv"])
def test_tensor_core_score_phase_variant_agrees(cuda_device, fmt):
    """decode_kernel_tc (context option decode_tc = 1; not the default: slower than the FFMA2 score phase on B200): the same
    statuses, labels and counters, scores within the parity tolerance of the default kernel's, flags equal away from the threshold."""
    x = synth.sensor_rows(50_003, seed=41)
    if fmt == "avro_a":
        data, offs = synth.avro_a_batch(x)
        kw, skip = dict(avro_schema=schemas.schema_a_json()), 5
    else:
        data, offs = synth.csv_batch(x, seed=41)
        kw, skip = dict(csv_kinds=SENSOR_KINDS), 0
    data, offs, bad = synth.corrupt_batch(data, offs, fmt, 1e-3, seed=2)
    m = _model18()
    res = {}
    try:
        for tc in (0, 1):
            bs.keras.set_option("decode_tc", tc)
            for n in (50_003, 127, 129, 1):
                rb = dec.RecordBatch(data[: offs[n]], offs[: n + 1], skip_prefix=skip)
                cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
                mse, flag, label, status = dec.decode_score(m, rb, counters=cnt, **kw)
                res[(tc, n)] = (mse.cpu().numpy().astype(np.float64), flag.cpu().numpy(), status.cpu().numpy(), cnt.cpu().tolist())
    finally:
        bs.keras.set_option("decode_tc", 0)
    for n in (50_003, 127, 129, 1):
        a, b = res[(0, n)], res[(1, n)]
        np.testing.assert_array_equal(a[2], b[2])
        ok = a[2] == 0
        assert np.array_equal(np.isnan(a[0]), np.isnan(b[0])) and (np.abs(a[0][ok] - b[0][ok]) <= np.maximum(1e-5 * a[0][ok], 1e-7)).all()
        safe = ok & (np.abs(a[0] - 5.0) > 1e-4)
        np.testing.assert_array_equal(a[1][safe], b[1][safe])
        assert a[3][0] == b[3][0] and a[3][2] == b[3][2] and a[3][3] == b[3][3]


@pytest.mark.parametrize("fmt", ["avro_a", "avro_b", "csv"])
def test_zero_set_instantiation_is_bit_identical(cuda_device, fmt):
    """The fused kernel compiled for the zero set of the reference's normalisation table (features 0, 2, 4, 5 are forced to 0.0 by
    normalize_fn, cardata-v1.py:72-87: their layer-1 products are left out) must give the same bits as the generic instantiation
    (context option zero_skip = 0) — also when the zeroed columns carry NaN / inf / huge values, and with damaged records."""
    x = synth.sensor_rows(60_001, seed=77)
    rng = np.random.default_rng(5)
    for col in (0, 2, 4, 5):
        rows = rng.choice(len(x), 300, replace=False)
        x[rows[:100], col] = np.nan if fmt != "csv" else 1e30
        x[rows[100:200], col] = np.inf if fmt != "csv" else -3.4e38
        x[rows[200:], col] = -1e30
    if fmt == "avro_a":
        data, offs = synth.avro_a_batch(x)
        kw, skip = dict(avro_schema=schemas.schema_a_json()), 5
    elif fmt == "avro_b":
        data, offs, _ = synth.avro_b_batch(x, null_frac=0.02, seed=5, label_probs=(0.9, 0.05, 0.05))
        kw, skip = dict(avro_schema=schemas.schema_b_json()), 5
    else:
        data, offs = synth.csv_batch(x, seed=7)
        kw, skip = dict(csv_kinds=SENSOR_KINDS), 0
    data, offs, bad = synth.corrupt_batch(data, offs, fmt, 1e-3, seed=11)
    m = _model18()
    res = {}
    try:
        for zs in (1, 0):
            bs.keras.set_option("zero_skip", zs)
            for n in (60_001, 33, 1):
                rb = dec.RecordBatch(data[: offs[n]], offs[: n + 1], skip_prefix=skip)
                cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
                mse, flag, label, status = dec.decode_score(m, rb, counters=cnt, **kw)
                res[(zs, n)] = (_bits(mse.cpu().numpy()), flag.cpu().numpy(), status.cpu().numpy(), cnt.cpu().tolist())
    finally:
        bs.keras.set_option("zero_skip", 1)
    assert (res[(1, 60_001)][2] != 0).sum() == len(bad)
    assert np.isfinite(res[(1, 60_001)][0].view(np.float32)[res[(1, 60_001)][2] == 0]).all()  # the poisoned columns never reach the score
    for n in (60_001, 33, 1):
        for a, b in zip(res[(1, n)][:3], res[(0, n)][:3]):
            np.testing.assert_array_equal(a, b)
        assert res[(1, n)][3] == res[(0, n)][3]
    # a table with another zero set, or a non-finite weight in a skipped row, runs the generic instantiation: same answer as the oracle path
    kind, lo, rg = schemas.norm_arrays()
    kind2 = kind.copy()
    kind2[1] = schemas.NORM_ZERO
    m2 = _model18()
    m2.set_normalization(kind2, lo, rg)
    rb = dec.RecordBatch(data, offs, skip_prefix=skip)
    a = dec.decode_score(m2, rb, **kw)[0].cpu().numpy()
    bs.keras.set_option("zero_skip", 0)
    try:
        b = dec.decode_score(m2, rb, **kw)[0].cpu().numpy()
    finally:
        bs.keras.set_option("zero_skip", 1)
    np.testing.assert_array_equal(_bits(a), _bits(b))


def test_csv_zeroed_columns_are_still_validated(cuda_device):
    """A CSV float field whose column the normalisation forces to zero is not converted, but a token tf.io.decode_csv would refuse
    there must still mark the record: same statuses (and scores) with zero_skip on and off."""
    x = synth.sensor_rows(200, seed=5)
    data, offs = synth.csv_batch(x, seed=5)
    lines = [bytes(data[offs[i]:offs[i + 1]]).decode() for i in range(len(x))]
    bad_tokens = ["abc", "", "1e", "--1", "1.2.3", "0x10", " 1", "1e400", "-", "+", "1,5"]
    want_bad = set()
    for k, tok in enumerate(bad_tokens * 4):
        row, col = 3 * k + 1, (0, 2, 4, 5)[k % 4]
        f = lines[row].split(",")
        f[2 + col] = tok
        lines[row] = ",".join(f)
        if tok in ("abc", "1e", "--1", "1.2.3", "-", "+"):  # not a number under any reading
            want_bad.add(row)
    rb = dec.RecordBatch.from_messages(lines)
    m = _model18()
    res = {}
    try:
        for zs in (1, 0):
            bs.keras.set_option("zero_skip", zs)
            mse, flag, _, status = dec.decode_score(m, rb, csv_kinds=SENSOR_KINDS)
            res[zs] = (_bits(mse.cpu().numpy()), flag.cpu().numpy(), status.cpu().numpy())
    finally:
        bs.keras.set_option("zero_skip", 1)
    for a, b in zip(res[1], res[0]):
        np.testing.assert_array_equal(a, b)
    got_bad = set(np.nonzero(res[1][2])[0].tolist())
    assert want_bad <= got_bad, [lines[i] for i in sorted(want_bad - got_bad)][:4]
    assert all(i % 3 == 1 for i in got_bad)  # only rows that were touched


def test_float64_normalisation_order_of_cardata_v3(cuda_device):
    """cardata-v3.py decodes the KSQL Avro records as float64, runs normalize_fn on float64 and only then lets Keras round the
    model input to float32 (:53-168).  With set_normalization(..., order="f64") the fused kernel follows that order: its scores
    carry the same bits as scoring rows that NumPy decoded and normalised in float64 and rounded to float32 last — and they differ
    from the float32 order (cardata-v1.py) in a few last bits.  Calls that have no float64 values refuse the option."""
    from oracle import cpu_oracle as orc

    rng = np.random.default_rng(21)
    x = synth.sensor_rows(3000, seed=21).astype(np.float64)
    x *= 1.0 + 1e-9 * rng.standard_normal(x.shape)           # not representable in float32: the two orders can differ
    x[5, 1], x[6, 3], x[7, 6] = 1e300, -1e-300, np.nan
    data, offs, _ = synth.avro_b_batch(x, null_frac=0.03, seed=4, label_probs=(0.9, 0.05, 0.05))
    data, offs, bad = synth.corrupt_batch(data, offs, "avro_b", 2e-3, seed=3)
    sch = schemas.schema_b_json()
    prog = bs.avro_schema.compile_schema(sch)
    kind, lo, rg = schemas.norm_arrays()
    m64 = _model18()
    m64.set_normalization(kind, lo, rg, order="f64")
    rb = dec.RecordBatch(data, offs, skip_prefix=5)
    cnt = torch.zeros(4, dtype=torch.int64, device="cuda")
    mse64, flag64, label64, status64 = dec.decode_score(m64, rb, avro_schema=sch, counters=cnt)
    # the reference order on the CPU: float64 decode, float64 normalize_fn, float32 cast, then the model (here: the same kernel math
    # on pre-normalised rows, which is bit-identical to the fused kernel's score phase)
    fields = [(f.type, f.nullable, f.out) for f in prog.fields]
    v64, lab_ref, st_ref = orc.decode_avro(data, offs, fields, 18, skip_prefix=5, dtype=np.float64)
    np.testing.assert_array_equal(status64.cpu().numpy(), st_ref)
    ok = st_ref == 0
    assert (~ok).sum() == len(bad)
    with np.errstate(all="ignore"):
        xn = orc.normalize(v64[ok], kind, lo.astype(np.float64), rg.astype(np.float64), dtype=np.float64).astype(np.float32)
    plain = _model18()
    plain.set_normalization(None, None, None)
    ref_mse, ref_flag = plain.score_device(torch.from_numpy(xn).cuda())
    np.testing.assert_array_equal(_bits(mse64.cpu().numpy()[ok]), _bits(ref_mse.cpu().numpy()))
    np.testing.assert_array_equal(flag64.cpu().numpy()[ok], ref_flag.cpu().numpy())
    np.testing.assert_array_equal(label64.cpu().numpy()[ok], lab_ref[ok])
    assert np.isnan(mse64.cpu().numpy()[~ok]).all() and cnt.cpu().tolist()[2] == len(bad)
    # the float32 order on the same records: close, not identical
    m32 = _model18()
    mse32 = dec.decode_score(m32, rb, avro_schema=sch)[0].cpu().numpy()
    a, b = mse64.cpu().numpy()[ok], mse32[ok]
    fin = np.isfinite(a) & np.isfinite(b)
    assert (np.abs(a[fin] - b[fin]) <= 2e-6 * np.abs(b[fin]) + 1e-9).all() and (_bits(a[fin]) != _bits(b[fin])).any()
    # no float64 values there: refused
    for call in (lambda: dec.decode_score(m64, dec.RecordBatch(*synth.csv_batch(x[:10].astype(np.float32), seed=1)), csv_kinds=SENSOR_KINDS),
                 lambda: m64.score(x[:10].astype(np.float32))):
        with pytest.raises(Exception, match="float64"):
            call()


def test_host_score_pipeline_matches_one_batch_at_a_time(cuda_device):
    """HostScorePipeline (submit / result, two batches in flight, per-slot buffers) returns what decode_score_host returns, batch by
    batch — ragged sizes, an empty batch, results collected late, out of order and after the slot was about to be reused."""
    m = _model18()
    sch = schemas.schema_a_json()
    batches = []
    for i, n in enumerate((50_003, 7, 0, 1, 33_333, 64, 20_001)):
        if n == 0:
            batches.append((np.zeros(0, np.uint8), np.zeros(1, np.int32)))
        else:
            batches.append(synth.avro_a_batch(synth.sensor_rows(n, seed=100 + i)))
    want = [dec.decode_score_host(m, d, o, avro_schema=sch, skip_prefix=5) for d, o in batches]
    pipe = dec.HostScorePipeline(m, depth=2, avro_schema=sch, skip_prefix=5)
    pinned = [(torch.from_numpy(d).pin_memory(), torch.from_numpy(o).pin_memory()) for d, o in batches]
    got = {}
    prev = None
    for k, (d, o) in enumerate(pinned):       # the consumer loop: submit k, then collect k - 1
        t = pipe.submit(d, o)
        if prev is not None:
            mse, flag = pipe.result(prev)
            got[prev] = (mse.copy(), flag.copy())
        prev = t
    for t, mse, flag in pipe.drain():
        got[t] = (mse.copy(), flag.copy())
    assert sorted(got) == list(range(len(batches)))
    for k, (mse, flag) in enumerate(want):
        np.testing.assert_array_equal(_bits(got[k][0]), _bits(mse))
        np.testing.assert_array_equal(got[k][1], flag)
    # a ticket that was never collected is waited for when its slot comes round again; collecting it afterwards is an error
    t0 = pipe.submit(*pinned[0])
    t1 = pipe.submit(*pinned[4])
    t2 = pipe.submit(*pinned[6])               # reuses t0's slot
    with pytest.raises(ValueError):
        pipe.result(t0)
    m2, f2 = pipe.result(t2)
    m1, f1 = pipe.result(t1)
    np.testing.assert_array_equal(_bits(m2), _bits(want[6][0]))
    np.testing.assert_array_equal(_bits(m1), _bits(want[4][0]))
    # CSV through the same pipeline, depth 3
    xs = synth.sensor_rows(9_000, seed=5)
    dcsv, ocsv = synth.csv_batch(xs, seed=2)
    p3 = dec.HostScorePipeline(m, depth=3, csv_kinds=SENSOR_KINDS)
    ts = [p3.submit(dcsv, ocsv) for _ in range(3)]
    ref = dec.decode_score_host(m, dcsv, ocsv, csv_kinds=SENSOR_KINDS)
    for t in ts:
        mse, flag = p3.result(t)
        np.testing.assert_array_equal(_bits(mse), _bits(ref[0]))
        np.testing.assert_array_equal(flag, ref[1])
