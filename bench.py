#!/usr/bin/env python
"""Benchmark of the sensor-record scoring hot path (BASELINE.json metric: records/s scored, % of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic records that is already resident in HBM
(`value`), or that starts in pinned HOST memory and whose results end in host memory (`e2e`).  Multi-GPU runs are
launched by torchrun (one rank per GPU); ranks own disjoint Kafka-partition-like shards (weak scaling, no data-path
collective; one NCCL all-reduce of the 4 counters per step).  `--impl reference` times the CPU restatement of the
reference's path (oracle/oracle.c, OpenMP over every host core) on the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "sensor records/sec scored (decode+AE+MSE)"
UNIT = "records/s"
W30 = [30, 14, 7, 7, 30]
W18 = [18, 14, 7, 7, 18]
ACTS = [1, 2, 1, 2]
THRESHOLD = 5.0


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(kernel_key):
    """DRAM bytes/record of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        return json.load(open(p)).get(kernel_key)
    return None


class ClockSampler:
    """nvidia-smi sampling DURING the timed region (B200_PROFILING.md clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------------------------
def golden_weights():
    return np.load(os.path.join(ROOT, "tests", "golden", "ae30_5epochs.npz"))["weights"]


class ScoreWorkload:
    """BASELINE.json configs[1]: synthetic pre-decoded float sensor rows, fused AE forward + MSE + flag."""

    def __init__(self, d, rows_log2, e2e_rows_log2, repeat=1):
        self.d = d
        self.launch_rows = 1 << rows_log2
        self.repeat = repeat
        self.rows = self.launch_rows * repeat
        self.e2e_rows = 1 << e2e_rows_log2
        self.name = f"score_predecoded_d{d}"
        self.bytes_per_record = 4 * d + 5
        self.flop_per_record = 2 * (14 * d + 98 + 49 + 7 * d)
        self.launches_per_step = repeat
        self.decode_launches_per_step = repeat
        self.kernel_key = f"ae_score_d{d}"
        self.kernel_ms = 0.0

    def describe(self):
        return {"workload": self.name, "features": self.d, "model": "Dense AE %d-14-7-7-%d" % (self.d, self.d),
                "weights": "autoencoder_sensor_anomaly_detection.h5 (shipped)" if self.d == 30 else "Glorot-uniform seed 314 (no 18-wide weights are shipped)",
                "normalise": self.d == 18, "rows_per_launch": self.launch_rows, "launches_per_step": self.repeat,
                "records_per_step_per_gpu": self.rows, "threshold": THRESHOLD,
                "l2_policy": "input per launch (%.1f GB) is far larger than the 126 MB L2; offsets rotate" % (self.launch_rows * 4 * self.d / 1e9),
                "e2e_rows_per_step_per_gpu": self.e2e_rows}

    def weights(self):
        import b200_sensor as bs

        return golden_weights() if self.d == 30 else bs.keras_compat.init_dense_weights(W18, seed=314)

    def widths(self):
        return W30 if self.d == 30 else W18

    def norm(self):
        import b200_sensor as bs

        return bs.schemas.norm_arrays() if self.d == 18 else None

    def model(self):
        import b200_sensor as bs

        k = bs.keras
        inp = k.layers.Input(shape=(self.d,))
        h = k.layers.Dense(14, activation="tanh", activity_regularizer=k.regularizers.l1(1e-7))(inp)
        h = k.layers.Dense(7, activation="relu")(h)
        h = k.layers.Dense(7, activation="tanh")(h)
        h = k.layers.Dense(self.d, activation="relu")(h)
        m = k.models.Model(inputs=inp, outputs=h)
        m.set_flat_weights(self.weights())
        if self.d == 18:
            m.set_normalization(*self.norm())
        return m

    def host_rows(self, n, seed):
        import b200_sensor as bs

        return bs.synth.credit_rows(n, seed=seed) if self.d == 30 else bs.synth.sensor_rows(n, seed=seed)

    def setup(self, dev, rank):
        import torch

        self.m = self.model()
        base = torch.from_numpy(self.host_rows(1 << 20, 314 + rank)).to(dev)
        # tile the seeded 2^20-row block with per-copy jitter so that all rows differ, entirely on the device
        reps = self.launch_rows // base.shape[0] + 1  # one extra block so that launches can start at rotating offsets
        self.x = torch.empty((reps * base.shape[0], self.d), dtype=torch.float32, device=dev)
        g = torch.Generator(device=dev).manual_seed(314 + rank)
        for r in range(reps):
            blk = self.x[r * base.shape[0]:(r + 1) * base.shape[0]]
            blk.copy_(base)
            if self.d == 30:
                blk.mul_(1.0 + 0.01 * torch.rand((), device=dev, generator=g))
        self.mse = torch.empty(self.launch_rows, dtype=torch.float32, device=dev)
        self.flag = torch.empty(self.launch_rows, dtype=torch.uint8, device=dev)
        self.counters = torch.zeros(4, dtype=torch.int64, device=dev)
        self.max_off = self.x.shape[0] - self.launch_rows
        self.e0, self.e1, self.pending = self._ev(), self._ev(), []

    def step(self, i):
        self.e0.record()
        for r in range(self.repeat):
            off = (((i * self.repeat + r) * 104729 * 2) % max(self.max_off, 1)) & ~1  # even row offset: 16-byte aligned bulk path
            self.m.score_device(self.x[off:off + self.launch_rows], THRESHOLD, mse=self.mse, flag=self.flag, counters=self.counters)
        self.e1.record()
        self.pending.append((self.e0, self.e1))
        self.e0, self.e1 = self._ev(), self._ev()
        return self.rows

    def _ev(self):
        import torch

        return torch.cuda.Event(enable_timing=True)

    def finish(self):
        """Sum of the per-launch CUDA-event times (events on torch's current stream = the stream the kernel is launched on)."""
        import torch

        torch.cuda.synchronize()
        self.kernel_ms += sum(a.elapsed_time(b) for a, b in self.pending)
        self.pending = []

    def setup_e2e(self, dev, rank):
        import torch

        self.hx = torch.from_numpy(self.host_rows(self.e2e_rows, 1000 + rank)).pin_memory()
        self.h2d = self.hx.numel() * 4
        self.d2h = self.e2e_rows * 5

    def step_e2e(self, i):
        mse, flag = self.m.score(self.hx, THRESHOLD, reuse_output=True)  # host in → host out through the public API
        return self.e2e_rows

    def host_stats(self):
        return self.bytes_per_record

    def copy_only(self, dev, steps):
        import torch

        dx = torch.empty_like(self.hx, device=dev)
        res = torch.empty(self.e2e_rows * 5, dtype=torch.uint8, device=dev)
        hres = torch.empty(self.e2e_rows * 5, dtype=torch.uint8).pin_memory()
        back = torch.cuda.Stream(device=dev)

        def copies():
            dx.copy_(self.hx, non_blocking=True)
            with torch.cuda.stream(back):
                hres.copy_(res, non_blocking=True)
            torch.cuda.synchronize()

        for _ in range(2):
            copies()
        t0 = time.perf_counter()
        for _ in range(steps):
            copies()
        return (time.perf_counter() - t0) * 1e3 / steps

    def cpu_prepare(self, threads=None):
        from oracle import c_oracle as co

        co.set_threads(threads or os.cpu_count())  # torchrun exports OMP_NUM_THREADS=1: ask for every host thread explicitly
        x = self.host_rows(1 << 22, 7)
        w, wd, nm = self.weights(), self.widths(), self.norm()

        def run():
            co.score(x, w, wd, ACTS, norm=nm, threshold=THRESHOLD)

        return run, x.shape[0], (co.max_threads() if not threads else threads)


class DecodeScoreWorkload:
    """BASELINE.json configs[2] / SURVEY §8d C3: synthetic Kafka-shaped record batches (CSV lines / Confluent-framed Avro / JSON),
    1e-4 of the records malformed, fused decode → normalize_fn → Dense AE 18-14-7-7-18 → MSE → flag.  One step = the rank's
    partition batches pushed `repeat` times through b2_partition_run_repeat (the per-GPU partition driver, K5): one kernel launch
    per partition batch, no Python between launches."""

    MALFORMED = 1e-4

    def __init__(self, fmt, batch_log2=20, n_batches=4, repeat=32):
        self.fmt = fmt
        self.batch = 1 << batch_log2
        self.nb = n_batches
        self.repeat = repeat
        self.name = f"decode_score_{fmt}"
        # CSV / JSON launches are followed by their (normally empty) oversize-retry launch
        self.launches_per_step = n_batches * repeat * (2 if fmt in ("csv", "json", "avro_b") else 1)
        self.decode_launches_per_step = n_batches * repeat
        self.kernel_key = f"decode_score_{fmt}"
        self.flop_per_record = 2 * (14 * 18 + 98 + 49 + 7 * 18)
        self.bytes_per_record = None  # measured from the generated batches (host_stats / setup)
        self.rows = self.batch * self.nb * self.repeat
        self.e2e_rows = self.batch
        self.kernel_ms = 0.0
        self.n_streams = 2

    def describe(self):
        return {"workload": self.name, "format": {"csv": "CSV line per message (car-sensor-data.csv layout, 20 fields)",
                                                  "avro_a": "Confluent-framed Avro, testdata/cardata-v1.avsc (63 B/record)",
                                                  "avro_b": "Confluent-framed Avro, KSQL cardata-v1.avsc (nullable unions, ~139 B/record)",
                                                  "json": "flat JSON object per message, the simulator's sensor-data topic (LSTM-…/cardata-v1.json member names, ~520 B/record)"}[self.fmt],
                "features": 18, "model": "Dense AE 18-14-7-7-18", "weights": "Glorot-uniform seed 314 (no 18-wide weights are shipped)",
                "normalise": True, "records_per_batch": self.batch, "partition_batches_per_gpu": self.nb,
                "passes_per_step": self.repeat, "records_per_step_per_gpu": self.rows, "malformed_fraction": self.MALFORMED,
                "threshold": THRESHOLD,
                "partitioning": "rank r owns partition batches p with seed 314 + (r * partition_batches_per_gpu + p); no data-path "
                                "collective; driver = b2_partition_run_repeat on 2 CUDA streams",
                "l2_policy": "every pass streams %d distinct batches back to back (%s in total, the L2 holds 126 MB), so no launch "
                             "finds its input in L2" % (self.nb, "> 250 MB"),
                "bytes_per_record_payload_plus_offset_plus_out": self.bytes_per_record,
                "e2e_rows_per_step_per_gpu": self.e2e_rows}

    def weights(self):
        import b200_sensor as bs

        return bs.keras_compat.init_dense_weights(W18, seed=314)

    def _host_batch(self, seed, n=None):
        import b200_sensor as bs

        rows = bs.synth.sensor_rows(n or self.batch, seed=seed)
        if self.fmt == "csv":
            d, o = bs.synth.csv_batch(rows, seed=seed)
        elif self.fmt == "avro_a":
            d, o = bs.synth.avro_a_batch(rows)
        elif self.fmt == "json":
            d, o = bs.synth.json_batch(rows, labels=(np.arange(rows.shape[0]) % 64 == 0))
        else:
            d, o, _ = bs.synth.avro_b_batch(rows, seed=seed)
        d, o, _ = bs.synth.corrupt_batch(d, o, self.fmt, self.MALFORMED, seed=seed)
        return d, o

    def host_stats(self):
        """bytes/record of this workload from one generated batch (same generator and seed in both arms)."""
        if self.bytes_per_record is None:
            d, o = self._host_batch(314, n=1 << 16)
            self.bytes_per_record = round(float(o[-1]) / (len(o) - 1) + 4 + 5, 1)
        return self.bytes_per_record

    def _kw(self):
        import b200_sensor as bs

        if self.fmt == "csv":
            return dict(csv_kinds=[0, 0] + [1] * 18)
        if self.fmt == "json":
            return dict(json_keys=list(bs.schemas.FEATURE_NAMES), json_label="failure_occurred")
        return dict(avro_schema=bs.schemas.schema_a_json() if self.fmt == "avro_a" else bs.schemas.schema_b_json())

    def setup(self, dev, rank):
        import torch

        import b200_sensor as bs

        self.host_stats()
        self.m = ScoreWorkload(18, 20, 20).model()
        self.kw = self._kw()
        self.batches = []
        for p in range(self.nb):
            d, o = self._host_batch(314 + rank * self.nb + p)
            rb = bs.decode.RecordBatch(d, o, dev, skip_prefix=0 if self.fmt in ("csv", "json") else 5)
            rb.mse = torch.empty(rb.n, dtype=torch.float32, device=dev)
            rb.flag = torch.empty(rb.n, dtype=torch.uint8, device=dev)
            rb.status = torch.empty(rb.n, dtype=torch.int32, device=dev)
            self.batches.append(rb)
        self.counters = torch.zeros(4, dtype=torch.int64, device=dev)
        self.driver = None
        if self.fmt != "json":  # the partition driver takes CSV / Avro specs; JSON batches go through the Python loop
            self.driver = bs.partition.PartitionDriver(self.m, threshold=THRESHOLD, n_streams=self.n_streams, **self.kw)
        self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def step(self, i):
        import b200_sensor as bs

        if self.driver is not None:
            st = self.driver.run(self.batches, counters=self.counters, repeat=self.repeat)
            self.kernel_ms += st.elapsed_ms   # CUDA events recorded on the driver's own streams around the launches
            return self.rows
        self.e0.record()
        for _ in range(self.repeat):
            for rb in self.batches:
                bs.decode.decode_score(self.m, rb, threshold=THRESHOLD, mse=rb.mse, flag=rb.flag, status=rb.status,
                                       counters=self.counters, **self.kw)
        self.e1.record()
        self.e1.synchronize()
        self.kernel_ms += self.e0.elapsed_time(self.e1)
        return self.rows

    def setup_e2e(self, dev, rank):
        import torch

        d, o = self._host_batch(2000 + rank)
        self.hd = torch.from_numpy(d).pin_memory()
        self.ho = torch.from_numpy(o).pin_memory()
        self.h2d = self.hd.numel() + self.ho.numel() * 4
        self.d2h = self.e2e_rows * 5
        import b200_sensor as bs

        # the consumer-loop form of the public API: batch i + 1 is submitted (H2D, kernels, D2H enqueued) before the results of batch i
        # are collected; every batch still pays its own H2D and D2H inside the timed region
        self.pipe = bs.decode.HostScorePipeline(self.m, depth=2, threshold=THRESHOLD, skip_prefix=0 if self.fmt in ("csv", "json") else 5, **self.kw)
        self.pending = None

    def step_e2e(self, i):
        t = self.pipe.submit(self.hd, self.ho)
        if self.pending is not None:
            mse, flag = self.pipe.result(self.pending)   # host arrays of the previous batch
        self.pending = t
        return self.e2e_rows

    def copy_only(self, dev, steps):
        """The e2e step's transfers alone (same pinned buffers: payload + offsets H2D, 5 B/record D2H, no kernel)."""
        import torch

        dd = torch.empty(self.hd.numel(), dtype=torch.uint8, device=dev)
        do = torch.empty(self.ho.numel(), dtype=torch.int32, device=dev)
        res = torch.empty(self.e2e_rows * 5, dtype=torch.uint8, device=dev)
        hres = torch.empty(self.e2e_rows * 5, dtype=torch.uint8).pin_memory()
        back = torch.cuda.Stream(device=dev)  # the results go back while the payload comes in (PCIe is full duplex), as in the e2e step

        def copies():
            dd.copy_(self.hd, non_blocking=True), do.copy_(self.ho, non_blocking=True)
            with torch.cuda.stream(back):
                hres.copy_(res, non_blocking=True)

        for _ in range(2):
            copies()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):   # back to back, like the pipelined e2e step: one wait at the end
            copies()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3 / steps

    def cpu_prepare(self, threads=None):
        """→ (run, n_records, cores): `run()` pushes one generated batch through the CPU restatement (oracle/oracle.c, OpenMP)."""
        import b200_sensor as bs
        from oracle import c_oracle as co

        if self.fmt == "json":  # no C restatement of the JSON walk: Python's json + the NumPy scoring restatement, one core
            from oracle import cpu_oracle as orc

            d, o = self._host_batch(7, n=1 << 14)
            w, nm = self.weights(), bs.schemas.norm_arrays()

            def run_json():
                x, _, _ = orc.decode_json(d, o, list(bs.schemas.FEATURE_NAMES), "failure_occurred")
                orc.score(np.nan_to_num(x), w, W18, ACTS, norm=nm, threshold=THRESHOLD)

            return run_json, len(o) - 1, 1
        co.set_threads(threads or os.cpu_count())  # torchrun exports OMP_NUM_THREADS=1: ask for every host thread explicitly
        d, o = self._host_batch(7, n=1 << 20)
        w, nm = self.weights(), bs.schemas.norm_arrays()
        if self.fmt == "csv":
            kw = dict(csv=([0, 0] + [1] * 18, None, ",", True))
        else:
            prog = bs.avro_schema.compile_schema(self._kw()["avro_schema"])
            kw = dict(avro=[(f.type, f.nullable, f.out) for f in prog.fields], skip_prefix=5)

        def run():
            co.decode_score(d, o, w, W18, ACTS, norm=nm, threshold=THRESHOLD, **kw)

        return run, len(o) - 1, (co.max_threads() if not threads else threads)


def aux_fit(dev):
    """BASELINE.json configs[3]: autoencoder training, 227 468 x 30 rows, batch 128 (the notebook's run, nb2.0:653-853)."""
    import torch

    import b200_sensor as bs

    m = ScoreWorkload(30, 20, 20).model()
    m.set_flat_weights(bs.keras_compat.init_dense_weights(W30, seed=314))
    m.compile(optimizer="adam", loss="mean_squared_error", metrics=["accuracy"])
    x = torch.from_numpy(bs.synth.credit_rows(227468, seed=314, outlier_frac=0.0)).to(dev)
    m.fit(x, epochs=1, batch_size=128, shuffle=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    h = m.fit(x, epochs=3, batch_size=128, shuffle=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 3
    return {"value": 227468 / dt, "unit": "samples/s", "us_per_sample": dt / 227468 * 1e6, "s_per_epoch": dt, "batch": 128,
            "loss_after_4_epochs": h.history["loss"][-1], "kernel": "ae_fit_kernel (one persistent CTA per epoch)",
            "reference_published": {"us_per_sample": 16.0, "source": "nb2.0:653-853 (MacBook CPU, TF 2.0.0-alpha0)"}}


def aux_format_rows(dev):
    """np.array2string(row) for 2^20 reconstruction-like rows (relu outputs, 18 wide), device resident;
    the reference does this per row in Python (cardata-v1.py:218-224) — NumPy timed on a bounded sample beside it."""
    import numpy as np
    import torch

    import b200_sensor as bs

    n, d = 1 << 20, 18
    g = torch.Generator(device=dev).manual_seed(314)
    x = torch.relu(torch.randn((n, d), device=dev, generator=g) * 0.5 + 0.3)
    lib = bs._ffi.lib()
    bound = int(lib.b2_format_rows_bound(d))
    text = torch.empty(n * bound, dtype=torch.uint8, device=dev)
    offs = torch.empty(n, dtype=torch.int64, device=dev)
    lens = torch.empty(n, dtype=torch.int32, device=dev)
    cur = torch.zeros(2, dtype=torch.int64, device=dev)
    ctx, st = bs.keras_compat.get_ctx(dev.index), torch.cuda.current_stream().cuda_stream
    compat = bs.sink.installed_numpy_compat()

    def launch():
        cur.zero_()
        bs._ffi.check(lib.b2_format_rows(ctx, x.data_ptr(), n, d, compat, text.data_ptr(), text.numel(), offs.data_ptr(),
                                         lens.data_ptr(), cur.data_ptr(), st))

    for _ in range(3):
        launch()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        launch()
    e1.record()
    torch.cuda.synchronize()
    dt = e0.elapsed_time(e1) * 1e-3 / 10
    used = int(cur[0].item())
    xs = x[:20000].cpu().numpy()
    t0 = time.perf_counter()
    ref = [np.array2string(r) for r in xs]
    t_np = time.perf_counter() - t0
    return {"value": n / dt, "unit": "rows/s", "text_bytes_per_row": used / n, "gb_per_s_out": used / dt / 1e9,
            "kernel": "format_rows_kernel", "numpy_array2string_rows_per_s_1core": len(ref) / t_np, "numpy": np.__version__}


def aux_metrics(dev):
    """the threshold sweep behind roc_curve / precision_recall_curve (nb2.0 cells 23-24) on 2^22 device-resident
    scores; scikit-learn's roc_curve on the same data on one host core beside it."""
    import torch

    import b200_sensor as bs

    n = 1 << 22
    g = torch.Generator(device=dev).manual_seed(314)
    y = (torch.rand(n, device=dev, generator=g) < 0.002).to(torch.uint8)
    s = torch.randn(n, device=dev, generator=g) ** 2 + y * 6.0
    for _ in range(2):
        bs.metrics.threshold_sweep(y, s)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        fps, tps, thr = bs.metrics.threshold_sweep(y, s)   # includes the small D2H of the counters
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    out = {"value": n / dt, "unit": "scores/s", "n": n, "thresholds": int(thr.numel()), "kernel": "mt_bitonic_* + mt_sweep"}
    try:
        from sklearn.metrics import roc_curve

        yh, sh = y.cpu().numpy(), s.cpu().numpy()
        t0 = time.perf_counter()
        roc_curve(yh, sh)
        out["sklearn_roc_curve_scores_per_s_1core"] = n / (time.perf_counter() - t0)
    except Exception as e:  # scikit-learn is optional for the bench
        out["sklearn"] = "unavailable: %s" % type(e).__name__
    return out


def aux_lstm(dev, T):
    """BASELINE.json configs[4]: LSTM stack forward on synthetic sequences (B = 2^20 / T)."""
    import torch

    import b200_sensor as bs

    k = bs.keras
    m = k.models.Sequential()
    m.add(k.layers.LSTM(32, activation="relu", input_shape=(T, 18), return_sequences=True))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=False))
    m.add(k.layers.RepeatVector(T))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=True))
    m.add(k.layers.LSTM(32, activation="relu", return_sequences=True))
    m.add(k.layers.TimeDistributed(k.layers.Dense(18)))
    B = (1 << 20) // T
    g = torch.Generator(device=dev).manual_seed(314)
    x = torch.rand((B, T, 18), device=dev, generator=g) * 2 - 1
    y = torch.empty_like(x)
    lib = bs._ffi.lib()
    h = m._handle()
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        bs._ffi.check(lib.b2_lstm_forward(h, x.data_ptr(), B, T, y.data_ptr(), st))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        bs._ffi.check(lib.b2_lstm_forward(h, x.data_ptr(), B, T, y.data_ptr(), st))
    e1.record()
    torch.cuda.synchronize()
    dt = e0.elapsed_time(e1) * 1e-4
    return {"value": B * T / dt, "unit": "record-steps/s", "batch": B, "timesteps": T, "fp32_tflops": 36480 * B * T / dt / 1e12,
            "fp32_frac_of_74.4": 36480 * B * T / dt / 74.4e12}


# records per step are sized so that a step lasts ~0.1 s and the default 20 timed steps ~2 s (clocks and power reach their
# sustained state inside the timed region); the batches of one pass always exceed the L2
def aux_lstm_fit(dev):
    """SURVEY §8 f.4 / LSTM-TensorFlow-IO-Kafka/cardata-v1.py:191-194: fit(zip(window(1), skip(1)).batch(1).take(1000), epochs=5)
    — 5 000 dependent Adam steps of the 18 642-parameter stack; one persistent-CTA launch per epoch."""
    import torch

    import b200_sensor as bs

    k = bs.keras
    m = k.models.Sequential()
    m.add(k.layers.LSTM(32, activation="relu", input_shape=(1, 18), return_sequences=True))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=False))
    m.add(k.layers.RepeatVector(1))
    m.add(k.layers.LSTM(16, activation="relu", return_sequences=True))
    m.add(k.layers.LSTM(32, activation="relu", return_sequences=True))
    m.add(k.layers.TimeDistributed(k.layers.Dense(18)))
    m.compile(metrics=["accuracy"], loss="mean_squared_error", optimizer="adam")
    rows = k.normalize_fn(torch.from_numpy(bs.synth.sensor_rows(1001, seed=314, outlier_frac=0.0)).to(dev), *bs.schemas.norm_arrays())
    win = k.Windows(rows, 1, 1)
    m.fit(win, epochs=1, batch_size=1, shuffle=False)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    h = m.fit(win, epochs=5, batch_size=1, shuffle=False)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"value": 5000 / dt, "unit": "steps/s", "us_per_step": dt / 5000 * 1e6, "batch": 1, "look_back": 1, "epochs": 5,
            "steps_per_epoch": 1000, "loss_first_last": [h.history["loss"][0], h.history["loss"][-1]],
            "kernel": "lstm_fit_kernel (one persistent CTA per epoch: BPTT + Adam)"}


def aux_kafka_ingest(dev):
    """what KafkaDataset / librdkafka does in front of cardata-v1.py:10 — record-batch (v2) bytes of one
    partition segment → CRC-32C check + record walk (one warp per batch) → fused decode + score on the value ranges inside the
    segment (no packed copy).  2^20 Avro-A records in 4 096 uncompressed batches of 256; CUDA-event timed."""
    import torch

    import b200_sensor as bs
    from b200_sensor import _ffi, decode as dec
    from b200_sensor.keras_compat import _stream_ptr, get_ctx

    n, per = 1 << 20, 256
    data, offs = bs.synth.avro_a_batch(bs.synth.sensor_rows(n, seed=314 + 99))
    L = int(offs[1] - offs[0])
    seg = bs.kafka.write_segment_uniform(data.reshape(n, L), records_per_batch=per)
    d = torch.from_numpy(seg).to(dev)
    pos, cnt = bs.kafka.scan_batches(seg)
    nb = len(pos)
    base = np.zeros(nb, dtype=np.int64)
    np.cumsum(cnt[:-1], out=base[1:])
    pos_d, base_d = torch.from_numpy(pos).to(dev), torch.from_numpy(base).to(dev)
    begin, end = torch.zeros(n, dtype=torch.int32, device=dev), torch.zeros(n, dtype=torch.int32, device=dev)
    status = torch.zeros(nb, dtype=torch.int32, device=dev)
    lib, ctx = _ffi.lib(), get_ctx(dev.index)
    m = ScoreWorkload(18, 20, 20).model()
    spans = dec.RecordSpans(d, begin, end, n, skip_prefix=5)
    spans.avg_record_bytes = (int(d.numel()) + n - 1) // n
    out = dict(mse=torch.empty(n, dtype=torch.float32, device=dev), flag=torch.empty(n, dtype=torch.uint8, device=dev),
               label=torch.empty(n, dtype=torch.uint8, device=dev), status=torch.empty(n, dtype=torch.int32, device=dev))

    def index():
        _ffi.check(lib.b2_kafka_index_records(ctx, _ffi.ptr(d), int(d.numel()), _ffi.ptr(pos_d), _ffi.ptr(base_d), nb, 1, _ffi.ptr(begin),
                                              _ffi.ptr(end), _ffi.ptr(status), _stream_ptr()), "b2_kafka_index_records")

    def chain():
        index()
        dec.decode_score(m, spans, avro_schema=bs.schemas.schema_a_json(), threshold=THRESHOLD, **out)

    def timed(fn, iters=30):
        for _ in range(3):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters * 1e-3

    t_idx, t_chain = timed(index), timed(chain)
    assert not bool(status.any()) and int((out["status"] != 0).sum()) == 0
    return {"value": n / t_chain, "unit": UNIT, "records": n, "batches": nb, "segment_bytes": int(d.numel()),
            "index_us": t_idx * 1e6, "index_gbs": d.numel() / t_idx / 1e9, "segment_to_scores_us": t_chain * 1e6,
            "kernels": "kafka_index_kernel (one warp per batch) + decode_kernel<18,1,1,1> on ranges inside the segment"}


WORKLOADS = {
    "score30": lambda: ScoreWorkload(30, 27, 22, repeat=24),
    "score18": lambda: ScoreWorkload(18, 27, 22, repeat=24),
    "csv": lambda: DecodeScoreWorkload("csv", n_batches=4, repeat=128),
    "avro_a": lambda: DecodeScoreWorkload("avro_a", n_batches=8, repeat=320),
    "avro_b": lambda: DecodeScoreWorkload("avro_b", n_batches=4, repeat=384),
    "json": lambda: DecodeScoreWorkload("json", batch_log2=19, n_batches=4, repeat=96),
}
DEFAULT_WORKLOAD = "avro_a"   # decode+score on the schema BASELINE.json names (testdata/cardata-v1.avsc)


# ------------------------------------------------------------------------------------------------------------
def cpu_steady(wl, seconds, threads=None):
    """Steady-state rate of the CPU restatement: one batch generated up front, one untimed pass to wake the OpenMP threads and
    warm the caches, then whole passes over it until `seconds` have gone by.  → (records, seconds, cores)"""
    run, nrec, cores = wl.cpu_prepare(threads)
    run()
    done, t0 = 0, time.perf_counter()
    while True:
        run()
        done += nrec
        dt = time.perf_counter() - t0
        if dt >= seconds:
            return done, dt, cores


CPU_SAMPLE_TEXT = ("oracle/oracle.c (OpenMP, gcc -O3 -march=native) on every host thread; TensorFlow/Keras is not installable offline, "
                   "so the reference's predict path is timed as its C restatement")


def run_reference(args, wl, rank, world):
    """`--impl reference`: the reference's CPU path for the same workload, timed on rank 0's host cores.  One step = a bounded
    sample: whole passes over one generated batch for ~cpu_seconds / steps of wall time, in steady state (same loop as the
    cpu_baseline leg of the b200 arm)."""
    if rank != 0:
        return
    wl.host_stats()
    run, nrec, cores = wl.cpu_prepare()
    per_step_s = max(args.cpu_seconds * 2.0 / max(args.steps, 1), 0.25)
    for _ in range(max(args.warmup, 1)):
        run()
    t_tot, n_tot = 0.0, 0
    for _ in range(args.steps):
        t0 = time.perf_counter()
        while True:
            run()
            n_tot += nrec
            if time.perf_counter() - t0 >= per_step_s:
                break
        t_tot += time.perf_counter() - t0
    v = n_tot / t_tot
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": wl.describe(),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{n_tot} records of the same workload in {t_tot:.1f} s ({nrec} per pass, whole passes for "
                                       f">= {per_step_s:.2f} s per step); " + CPU_SAMPLE_TEXT},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), file=_JSON_OUT, flush=True)


_JSON_OUT = sys.stdout


def _quiet_stdout():
    """Keep stdout for the ONE JSON line: libraries that print to file descriptor 1 (NCCL's version banner) go to stderr."""
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def load_instr(kernel_key):
    """Executed warp-instructions per record of the dominant kernel from the committed ncu --set full capture
    (smsp__inst_executed.sum / records of that launch; profiles/instr_per_record.json)."""
    p = os.path.join(ROOT, "profiles", "instr_per_record.json")
    if os.path.exists(p):
        return json.load(open(p)).get(kernel_key)
    return None


def timed_steps(wl, steps, first_step, torch):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wl.kernel_ms = 0.0
    t0 = time.perf_counter()
    e0.record()
    recs = 0
    for i in range(steps):
        recs += wl.step(first_step + i)
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    if hasattr(wl, "finish"):
        wl.finish()
    # the partition driver synchronises its own streams inside each step, so the host clock brackets all device work; the
    # event pair on torch's stream covers launches made there.  Whichever is longer is the step time.
    return recs, max(e0.elapsed_time(e1), wall)


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("B2_BENCH_WORKLOAD", DEFAULT_WORKLOAD), choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target wall time of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--also", default=None, help="comma list of further workloads to measure (device-resident value only); "
                                                 "default: all others at N=1, none under torchrun")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]()

    if args.impl == "reference":
        run_reference(args, wl, rank, world)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device visible — the b200 arm has no CPU fallback (use --impl reference for the CPU arm)")
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"bench.py: --gpus {args.gpus} needs torchrun (one rank per GPU); WORLD_SIZE is 1")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    wl.setup(dev, rank)
    for i in range(args.warmup):
        wl.step(i)
    if hasattr(wl, "finish"):
        wl.finish()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    wl.counters.zero_()
    barrier()
    recs, ms_local = timed_steps(wl, args.steps, args.warmup, torch)
    kernel_ms = wl.kernel_ms
    if world > 1:  # the path's only exchange: the final count reduction over NVLink (cardata-style per-partition counters)
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        dist.all_reduce(wl.counters, op=dist.ReduceOp.SUM)
        c1.record()
        torch.cuda.synchronize()
        ms_local += c0.elapsed_time(c1)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(ms_local)
    total_records = recs * world
    value = total_records / (ms * 1e-3)
    counters = wl.counters.cpu().numpy().tolist()

    # end-to-end: pinned host → device → kernel → host, through the public API; and the same transfers with no kernel
    e2e = None
    if not args.no_e2e:
        wl.setup_e2e(dev, rank)
        for i in range(max(args.warmup, 3) * 8):   # the link and the host side have idled through the device-resident region: 24+ batches (≈ 30 ms) of warm-up
            wl.step_e2e(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e2e_steps = max(args.steps, 8) * 8   # an e2e step is one 2^20-record batch (~1.5 ms): many of them for a stable figure
        t0 = time.perf_counter()
        e0.record()
        n_e2e = 0
        for i in range(e2e_steps):
            n_e2e += wl.step_e2e(i)
        e1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        ms_e2e = max_over_ranks(max(e0.elapsed_time(e1), wall))
        barrier()
        ms_copy = max_over_ranks(wl.copy_only(dev, 32))
        e2e = {"value": n_e2e * world / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h,
               "ms_per_step": ms_e2e / e2e_steps, "steps": e2e_steps,
               "copy_only": {"ms_per_step": ms_copy, "gb_per_s_per_gpu": (wl.h2d + wl.d2h) / (ms_copy * 1e-3) / 1e9,
                             "records_per_s_ceiling": wl.e2e_rows * world / (ms_copy * 1e-3),
                             "note": "the same pinned buffers copied H2D + D2H (on two streams: full duplex, steps back to back) with no "
                                     "kernel, max over ranks: what the host / PCIe side alone allows at this N"},
               "api": "Model.score(pinned host tensor, reuse_output=True) -> (mse, flag) numpy; rows in 4 chunks, H2D / kernel / D2H overlapped" if isinstance(wl, ScoreWorkload) else
                      "decode.HostScorePipeline(model).submit(pinned bytes, pinned offsets) / .result(ticket) -> (mse, flag) numpy, two "
                      "batches in flight; per batch 2 chunks, H2D of a chunk under the kernel of the one before, results D2H chunk by "
                      "chunk on a third stream (decode.decode_score_host is the one-batch-at-a-time form of the same steps)"}

    if rank == 0:
        peak, peak_src = load_peaks()
        config = wl.describe()
        n_launch = args.steps * wl.decode_launches_per_step
        per_launch_ms = kernel_ms / n_launch
        rec_per_launch = recs / n_launch
        achieved = wl.bytes_per_record * rec_per_launch / (per_launch_ms * 1e-3) / 1e9
        traffic = load_traffic(wl.kernel_key)
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
        ipr = load_instr(wl.kernel_key)
        issue_peak = 148 * 4 * sm_mhz * 1e6     # warp-instructions/s: 4 schedulers per SM, one issue per clock each
        issue = None
        hbm_ceiling = peak * 1e9 / wl.bytes_per_record
        fp32_ceiling = 148 * 128 * 2 * sm_mhz * 1e6 / wl.flop_per_record
        bound = "hbm"
        if ipr:
            ach = ipr * rec_per_launch / (per_launch_ms * 1e-3)
            issue = {"achieved": ach, "peak": issue_peak, "unit": "warp-instr/s", "frac": ach / issue_peak,
                     "warp_instr_per_record": ipr, "records_per_s_ceiling": issue_peak / ipr,
                     "source": "smsp__inst_executed.sum of the committed ncu capture / records of that launch "
                               "(profiles/instr_per_record.json); peak = 148 SMs x 4 schedulers x SM clock under load"}
            if issue_peak / ipr < min(hbm_ceiling, fp32_ceiling):
                bound = "issue"
        if bound == "hbm" and fp32_ceiling < hbm_ceiling:
            bound = "fp32"
        roofline = {"bound": bound, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": None if traffic is None else traffic * rec_per_launch,
                    "peak_source": peak_src, "algorithmic_bytes_per_record": wl.bytes_per_record,
                    "flop_per_record": wl.flop_per_record,
                    "fp32_tflops_achieved": wl.flop_per_record * rec_per_launch / (per_launch_ms * 1e-3) / 1e12,
                    "kernel_ms_per_launch": per_launch_ms, "records_per_launch": rec_per_launch,
                    "kernel_time_share_of_step": kernel_ms / ms_local,
                    "timing": "CUDA events on the launching streams around the launches of each step (partition driver: "
                              "b2_run_stats.elapsed_ms), divided by the launches; with 2 streams consecutive launches overlap at "
                              "their tails, so this is the effective duration per launch",
                    "ceilings_records_per_s": {"hbm": hbm_ceiling, "fp32": fp32_ceiling,
                                               "issue": None if not ipr else issue_peak / ipr},
                    "issue": issue,
                    "note": "`bound` names the lowest of the three ceilings for this kernel; `frac` is always against the HBM peak",
                    "evidence": "profiles/r2_ncu_full_summary.json, profiles/round2_notes.md"}
        cpu = None
        if not args.no_cpu_baseline and world == 1:  # reported at N=1 only
            n, dt, cores = cpu_steady(wl, args.cpu_seconds)
            cpu = {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{n} records of the same workload in {dt:.1f} s, steady state; " + CPU_SAMPLE_TEXT}
        launches = args.steps * wl.launches_per_step
        also = {}
        names = [w for w in WORKLOADS if w != args.workload] if (args.also is None and world == 1) else \
            [w for w in (args.also or "").split(",") if w in WORKLOADS]   # "--also none" → nothing
        import gc

        wl = None  # free the headline workload's device buffers before measuring the others
        for nm in names:
            gc.collect()
            torch.cuda.empty_cache()
            w2 = WORKLOADS[nm]()
            w2.setup(dev, rank)
            for i in range(3):
                w2.step(i)
            if hasattr(w2, "finish"):
                w2.finish()
            torch.cuda.synchronize()
            r2, t2 = timed_steps(w2, max(args.steps // 4, 3), 3, torch)
            t2 *= 1e-3
            ipr2 = load_instr(w2.kernel_key)
            also[w2.name] = {"value": r2 / t2, "unit": UNIT, "hbm_frac": w2.bytes_per_record * r2 / t2 / 1e9 / peak,
                             "algorithmic_bytes_per_record": w2.bytes_per_record,
                             "fp32_tflops": w2.flop_per_record * r2 / t2 / 1e12,
                             "issue_frac": None if not ipr2 else ipr2 * r2 / t2 / issue_peak}
            w2 = None
        if names:
            gc.collect()
            torch.cuda.empty_cache()
            also["fit_d30_batch128"] = aux_fit(dev)
            also["lstm_fit_batch1_lookback1"] = aux_lstm_fit(dev)
            also["lstm_forward_T1"] = aux_lstm(dev, 1)
            also["lstm_forward_T16"] = aux_lstm(dev, 16)
            also["format_rows_d18"] = aux_format_rows(dev)
            also["metrics_threshold_sweep"] = aux_metrics(dev)
            also["kafka_segment_to_scores_avro_a"] = aux_kafka_ingest(dev)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": launches, "clocks": clocks, "also": also,
                "counters": {"records": counters[0], "flagged": counters[1], "decode_errors": counters[2], "bytes": counters[3]}}
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
