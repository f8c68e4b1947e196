"""Pin the CPU oracle against everything the reference checkout offers for this path: the derived
App. G known-answer vectors on the shipped weights, float64 shadows, torch autograd, and glibc/pyarrow parses."""
import json
import os

import numpy as np
import pytest

from oracle import cpu_oracle as orc

W30 = [30, 14, 7, 7, 30]
ACTS = [orc.ACT_TANH, orc.ACT_RELU, orc.ACT_TANH, orc.ACT_RELU]

# SURVEY.md App. G (float32 NumPy restatement on the shipped weights; tolerance 1e-6 absolute)
APPG = {"ae30_5epochs": {"zeros": 0.00108109054, "ones": 1.38512659, "ramp": 0.901749313},
        "ae30_100epochs": {"zeros": 0.0013377479, "ones": 0.868697047, "ramp": 0.722347617}}
APPG_Y = {"ae30_5epochs": {"zeros": [0, 0, 0, 0.176839, 0, 0], "ones": [0, 0, 0.31200135, 0, 0, 0.7072669]},
          "ae30_100epochs": {"zeros": [0.20026195, 0, 0, 0.005252615, 0, 0], "ones": [0.2021935, 1.5314744, 0, 0, 0.14146084, 0]}}


def _inputs():
    return {"zeros": np.zeros((1, 30), np.float32), "ones": np.ones((1, 30), np.float32),
            "ramp": ((np.arange(30, dtype=np.float32) - 15) / 10).reshape(1, 30)}


@pytest.mark.parametrize("key", ["ae30_5epochs", "ae30_100epochs"])
def test_appendix_g_vectors(golden_dir, key):
    w = np.load(os.path.join(golden_dir, key + ".npz"))["weights"]
    for nm, x in _inputs().items():
        m, fl, y, _ = orc.score(x, w, W30, ACTS, dtype=np.float32)
        assert abs(float(m[0]) - APPG[key][nm]) < 1e-6
        assert fl[0] == 0
        if nm in APPG_Y[key]:
            np.testing.assert_allclose(y[0, :6], APPG_Y[key][nm], atol=1e-6)
        m64 = orc.score(x, w, W30, ACTS, dtype=np.float64)[0]
        assert abs(float(m64[0]) - float(m[0])) < 1e-6


def test_committed_appg_json_is_current(golden_dir):
    a = json.load(open(os.path.join(golden_dir, "appg_vectors.json")))
    for key in APPG:
        for nm in APPG[key]:
            assert abs(a[key][nm]["mse"] - APPG[key][nm]) < 1e-6


def test_normalize_matches_formula_in_f64():
    from b200_sensor import schemas, synth

    kind, lo, rng = schemas.norm_arrays()
    x = synth.sensor_rows(2000, seed=1)
    a = orc.normalize(x, kind, lo, rng, dtype=np.float32)
    b = orc.normalize(x.astype(np.float64), kind, lo, rng, dtype=np.float64)
    assert np.all(a[:, [0, 2, 4, 5]] == 0)
    np.testing.assert_allclose(a, b, rtol=0, atol=2e-6)
    # firmware 1000/2000 → exactly -1 / +1; tyre pressure 20..35 → [-1, 1]
    assert set(np.unique(a[:, 17]).tolist()) <= {-1.0, 1.0}
    assert a[:, 9:13].min() >= -1.0 and a[:, 9:13].max() <= 1.0


def test_gradients_against_torch_autograd(ae30):
    import torch

    rng = np.random.default_rng(3)
    x = rng.standard_normal((64, 30))
    w = ae30["weights"].astype(np.float64)
    l1 = 1e-7
    loss, g, correct = orc.ae_loss_and_grads(x, w, W30, ACTS, l1, dtype=np.float64)
    wt = torch.tensor(w, dtype=torch.float64, requires_grad=True)
    xt = torch.tensor(x, dtype=torch.float64)
    h, p = xt, 0
    h1 = None
    for i, (a, b) in enumerate(zip(W30[:-1], W30[1:])):
        k = wt[p:p + a * b].reshape(a, b)
        p += a * b
        bias = wt[p:p + b]
        p += b
        z = h @ k + bias
        h = torch.tanh(z) if ACTS[i] == orc.ACT_TANH else torch.relu(z)
        if i == 0:
            h1 = h
    lt = ((h - xt) ** 2).mean(dim=1).mean() + l1 * h1.abs().sum() / x.shape[0]
    lt.backward()
    assert abs(float(lt.detach()) - float(loss)) < 1e-12
    np.testing.assert_allclose(g, wt.grad.numpy(), rtol=1e-9, atol=1e-12)


def test_adam_step_against_torch(ae30):
    import torch

    rng = np.random.default_rng(5)
    w = rng.standard_normal(50)
    wt = torch.tensor(w, dtype=torch.float64, requires_grad=True)
    opt = torch.optim.Adam([wt], lr=1e-3, betas=(0.9, 0.999), eps=1e-7)
    m = np.zeros(50)
    v = np.zeros(50)
    for t in range(1, 6):
        g = rng.standard_normal(50)
        wt.grad = torch.tensor(g)
        opt.step()
        w, m, v = orc.adam_step(w, m, v, g, t)
    # torch applies eps to sqrt(v_hat), Keras/TF to sqrt(v) with the bias correction folded into alpha_t: the two
    # differ by O(lr * eps / sqrt(v)) for the smallest |g| (measured 4.5e-7 here) and agree otherwise
    np.testing.assert_allclose(w, wt.detach().numpy(), rtol=0, atol=1e-6)


def test_lstm_against_float64_and_torch_cell():
    import torch

    w = None
    from b200_sensor import keras_compat as kc

    w = kc.init_lstm_weights(18, (32, 16, 16, 32), seed=314)
    assert w.size == 18642
    rng = np.random.default_rng(0)
    x = rng.uniform(-1, 1, size=(5, 4, 18)).astype(np.float32)
    y32 = orc.lstm_forward(x, w, dtype=np.float32)
    y64 = orc.lstm_forward(x, w, dtype=np.float64)
    assert y32.shape == (5, 4, 18)
    np.testing.assert_allclose(y32, y64, atol=2e-5)
    # independent check of ONE layer against a hand-written torch cell (gate order i,f,c,o; relu activation)
    ws = orc.lstm_unflatten(w.astype(np.float64), 18, (32, 16, 16, 32))
    W, U, b = (torch.tensor(a) for a in ws[0:3])
    xt = torch.tensor(x.astype(np.float64))
    h = torch.zeros(5, 32, dtype=torch.float64)
    c = torch.zeros(5, 32, dtype=torch.float64)
    outs = []
    for t in range(4):
        z = xt[:, t] @ W + h @ U + b
        i, f, g, o = z.split(32, dim=1)
        c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.relu(g)
        h = torch.sigmoid(o) * torch.relu(c)
        outs.append(h)
    ref = torch.stack(outs, 1).numpy()
    got = orc.lstm_layer(x.astype(np.float64), *ws[0:3], return_sequences=True, dtype=np.float64)
    np.testing.assert_allclose(got, ref, atol=1e-12)


def test_strtof_agrees_with_numpy_and_pyarrow_on_csv_tokens(golden_dir):
    import io

    import pyarrow as pa
    import pyarrow.csv as pacsv

    d = np.load(os.path.join(golden_dir, "csv_tokens.npz"))
    toks, bits = d["tokens"], d["bits"]
    assert len(toks) > 70000
    sub = toks[:: max(1, len(toks) // 20000)]
    sbits = bits[:: max(1, len(toks) // 20000)]
    mine = np.array([np.float32(orc.strtof(t.encode())[0]).view(np.uint32) for t in sub], dtype=np.uint32)
    np.testing.assert_array_equal(mine, sbits)
    via_np = np.array([np.float32(float(t)) for t in sub], dtype=np.float32).view(np.uint32)
    np.testing.assert_array_equal(via_np, sbits)
    tbl = pacsv.read_csv(io.BytesIO(("v\n" + "\n".join(sub.tolist()) + "\n").encode()),
                         convert_options=pacsv.ConvertOptions(column_types={"v": pa.float32()}))
    np.testing.assert_array_equal(tbl["v"].to_numpy().view(np.uint32), sbits)


def test_decode_csv_oracle_on_reference_sample(golden_dir):
    from b200_sensor import synth

    lines = open(os.path.join(golden_dir, "car_sensor_sample.csv")).read().strip().split("\n")[1:]
    data, offs = synth.lines_to_batch(lines)
    kinds = [0, 0] + [1] * 18
    out, spans, st = orc.decode_csv(data, offs, kinds)
    assert st.sum() == 0 and out.shape == (512, 18)
    ref = np.array([[np.float32(float(t)) for t in ln.split(",")[2:]] for ln in lines], dtype=np.float32)
    np.testing.assert_array_equal(out.view(np.uint32), ref.view(np.uint32))


def test_decode_csv_oracle_error_semantics():
    from b200_sensor import synth

    recs = [b"1,2,3", b"1,,3", b"1,x,3", b'1,"2",3', b'1,"2,3', b"1,2", b" 1 ,2.5e0,3", b'1,2"x,3', b"1,2,3,4", b"1,2,"]
    data, offs = synth.lines_to_batch(recs)
    out, _, st = orc.decode_csv(data, offs, [1, 1, 1], defaults=[None, 7.0, None])
    assert st.tolist() == [0, 0, orc.REC_BAD_NUMBER, 0, orc.REC_BAD_QUOTE, orc.REC_FIELD_COUNT, 0, orc.REC_BAD_QUOTE,
                           orc.REC_FIELD_COUNT, orc.REC_EMPTY_FIELD]
    assert out[1].tolist() == [1.0, 7.0, 3.0] and out[3].tolist() == [1.0, 2.0, 3.0] and out[6].tolist() == [1.0, 2.5, 3.0]


def test_decode_avro_oracle_roundtrip_both_schemas():
    from b200_sensor import avro_schema as av, schemas, synth

    x = synth.sensor_rows(300, seed=9)
    d, o = synth.avro_a_batch(x)
    assert np.all(o[1:] - o[:-1] == 63)  # 5 framing + 58 body
    pa_ = av.compile_schema(schemas.schema_a_json())
    out, lab, st = orc.decode_avro(d, o, [(f.type, f.nullable, f.out) for f in pa_.fields], pa_.n_out)
    exp = x.copy()
    exp[:, [9, 10, 11, 12, 17]] = np.rint(exp[:, [9, 10, 11, 12, 17]])
    assert st.sum() == 0
    np.testing.assert_array_equal(out, exp)
    d, o, labs = synth.avro_b_batch(x, null_frac=0.05)
    pb = av.compile_schema(schemas.schema_b_json())
    out, lab, st = orc.decode_avro(d, o, [(f.type, f.nullable, f.out) for f in pb.fields], pb.n_out)
    assert st.sum() == 0
    np.testing.assert_array_equal(lab, labs)
    nz = out != 0
    np.testing.assert_array_equal(out[nz], exp[nz])  # nulls decode to 0, everything else survives double→float exactly
    assert 0.02 < 1 - nz[:, [1, 3, 6]].mean() < 0.09
    # truncation is reported, not fatal
    d2, o2 = d[: o[-1] - 3], o.copy()
    o2[-1] -= 3
    _, _, st2 = orc.decode_avro(d2, o2, [(f.type, f.nullable, f.out) for f in pb.fields], pb.n_out)
    assert st2[-1] == orc.REC_TRUNCATED and st2[:-1].sum() == 0


def test_fit_oracle_decreases_loss():
    from b200_sensor import keras_compat as kc

    rng = np.random.default_rng(1)
    x = rng.uniform(-1, 1, size=(512, 18))
    w0 = kc.init_dense_weights([18, 14, 7, 7, 18], seed=314)
    w, m, v, t, losses, accs = orc.fit(x, w0, [18, 14, 7, 7, 18], ACTS, 1e-7, batch=32, epochs=3)
    assert t == 48 and losses[-16:].mean() < losses[:16].mean()


def test_lstm_bptt_gradients_match_torch_autograd():
    """oracle.lstm_loss_and_grads (the float64 BPTT the CUDA training kernel is compared with) against torch autograd."""
    import torch

    F, units = 18, (32, 16, 16, 32)
    g = np.random.default_rng(0)
    w = orc.init_lstm_weights(F, units, seed=3).astype(np.float64) + g.normal(size=18642) * 0.05
    for B, T in ((1, 1), (3, 4), (5, 2)):
        x, y = g.uniform(-1, 1, (B, T, F)), g.uniform(-1, 1, (B, T, F))
        loss, grads, acc, out = orc.lstm_loss_and_grads(x, y, w)
        tw = torch.tensor(w, dtype=torch.float64, requires_grad=True)
        ws, p = [], 0
        for shp in orc.lstm_param_shapes(F, units):
            n = int(np.prod(shp))
            ws.append(tw[p:p + n].reshape(shp))
            p += n

        def layer(xx, W, U, b, seq):
            u = U.shape[0]
            h = torch.zeros(xx.shape[0], u, dtype=torch.float64)
            c = torch.zeros_like(h)
            hs = []
            for t in range(xx.shape[1]):
                z = xx[:, t] @ W + h @ U + b
                i, f, gg, o = torch.sigmoid(z[:, :u]), torch.sigmoid(z[:, u:2 * u]), torch.relu(z[:, 2 * u:3 * u]), torch.sigmoid(z[:, 3 * u:])
                c = f * c + i * gg
                h = o * torch.relu(c)
                hs.append(h)
            return torch.stack(hs, 1) if seq else h

        h = layer(torch.tensor(x), *ws[0:3], True)
        h = layer(h, *ws[3:6], False)
        h = layer(h[:, None, :].repeat(1, T, 1), *ws[6:9], True)
        h = layer(h, *ws[9:12], True)
        o = h @ ws[12] + ws[13]
        L = ((o - torch.tensor(y)) ** 2).mean()
        L.backward()
        assert abs(loss - L.item()) < 1e-14 and np.abs(grads - tw.grad.numpy()).max() < 1e-14
        np.testing.assert_allclose(out, orc.lstm_forward(x, w, dtype=np.float64), rtol=0, atol=1e-15)
