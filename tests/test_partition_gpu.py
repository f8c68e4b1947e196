"""K5: the per-GPU partition driver through the C ABI — several partition batches over two CUDA streams give the same
per-record results and counters as scoring each batch alone."""
import numpy as np
import pytest
import torch

import b200_sensor as bs
from b200_sensor import decode as dec, partition as part, schemas, synth

pytestmark = pytest.mark.gpu


def _model18():
    k = bs.keras
    inp = k.layers.Input(shape=(18,))
    e = k.layers.Dense(14, activation="tanh", activity_regularizer=k.regularizers.l1(1e-7))(inp)
    e = k.layers.Dense(7, activation="relu")(e)
    d = k.layers.Dense(7, activation="tanh")(e)
    d = k.layers.Dense(18, activation="relu")(d)
    m = k.models.Model(inputs=inp, outputs=d, seed=314)
    m.set_normalization(*schemas.norm_arrays())
    return m


@pytest.mark.parametrize("fmt", ["csv", "avro_a"])
def test_partition_run_matches_single_calls(cuda_device, fmt):
    m = _model18()
    sizes = [30_000, 1, 77_777, 0, 4096]
    batches, singles = [], []
    for p, n in enumerate(sizes):
        rows = synth.sensor_rows(max(n, 1), seed=314 + p)[:n]
        if fmt == "csv":
            d, o = synth.csv_batch(rows, seed=p) if n else (np.zeros(0, np.uint8), np.zeros(1, np.int32))
            rb = dec.RecordBatch(d, o)
            kw = dict(csv_kinds=[0, 0] + [1] * 18)
        else:
            d, o = synth.avro_a_batch(rows) if n else (np.zeros(0, np.uint8), np.zeros(1, np.int32))
            rb = dec.RecordBatch(d, o, skip_prefix=5)
            kw = dict(avro_schema=schemas.schema_a_json())
        batches.append(rb)
        singles.append(dec.decode_score(m, rb, **kw)[:2] if n else None)
    drv = part.PartitionDriver(m, n_streams=2, **kw)
    st = drv.run(batches)
    assert st.records == sum(sizes) and st.decode_errors == 0 and st.elapsed_ms > 0
    assert st.bytes == sum(int(b.offsets[-1]) for b in batches)
    flagged = 0
    for b, s in zip(batches, singles):
        if s is None:
            continue
        assert torch.equal(b.mse, s[0]) and torch.equal(b.flag, s[1])
        flagged += int(s[1].sum())
    assert st.flagged == flagged
    assert part.reduce_counters([st.records, st.flagged, st.decode_errors, st.bytes]) == [st.records, st.flagged, 0, st.bytes]
