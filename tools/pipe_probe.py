"""HostScorePipeline: throughput against chunks per batch and batches in flight (2^20 Avro-A records per batch, pinned buffers)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, b200_sensor as bs
dev = torch.device("cuda", 0)
wl = bench.DecodeScoreWorkload("avro_a", n_batches=1, repeat=1)
wl.setup(dev, 0)
wl.setup_e2e(dev, 0)
for depth, chunks in ((2, 4), (2, 2), (2, 1), (3, 1), (3, 2), (2, 4)):
    pipe = bs.decode.HostScorePipeline(wl.m, depth=depth, chunks=chunks, threshold=5.0, skip_prefix=5, **wl.kw)
    prev = None
    def step():
        global prev
        t = pipe.submit(wl.hd, wl.ho)
        if prev is not None:
            pipe.result(prev)
        prev = t
    for _ in range(6):
        step()
    torch.cuda.synchronize()
    N = 300
    t0 = time.perf_counter()
    c0 = time.process_time()
    for _ in range(N):
        step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / N
    cpu = (time.process_time() - c0) * 1e3 / N
    pipe.drain(); prev = None
    print(f"depth={depth} chunks={chunks}: {ms:.3f} ms per batch = {wl.e2e_rows / (ms * 1e-3):.4g} records/s (host CPU {cpu:.3f} ms per batch)", flush=True)
print("ok")
