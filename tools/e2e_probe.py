"""End-to-end step (pinned host buffers -> H2D -> fused kernel -> D2H) of decode_score_host: chunk count and result copy-back
variants against the transfers alone (serial and full duplex)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, b200_sensor as bs
dev = torch.device("cuda", 0)
fmt = sys.argv[1] if len(sys.argv) > 1 else "avro_a"
wl = bench.DecodeScoreWorkload(fmt, n_batches=1, repeat=1)
wl.setup(dev, 0)
wl.setup_e2e(dev, 0)
sp = 0 if fmt in ("csv", "json") else 5
ref = None
for chunks, ov in ((4, False), (4, True), (8, True), (16, True), (8, False)):
    def step():
        return bs.decode.decode_score_host(wl.m, wl.hd, wl.ho, threshold=5.0, skip_prefix=sp, reuse_output=True, chunks=chunks, overlap_d2h=ov, **wl.kw)
    for _ in range(5):
        mse, flag = step()
    got = (mse.copy(), flag.copy())
    if ref is None:
        ref = got
    same = np.array_equal(ref[0].view(np.uint32), got[0].view(np.uint32)) and np.array_equal(ref[1], got[1])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    N = 200
    for _ in range(N):
        step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / N
    print(f"{fmt} chunks={chunks} overlap_d2h={ov}: {ms:.3f} ms per 2^20-record batch = {wl.e2e_rows / (ms * 1e-3):.4g} records/s; same results {same}", flush=True)
# transfers alone
dd = torch.empty(wl.hd.numel(), dtype=torch.uint8, device=dev)
do = torch.empty(wl.ho.numel(), dtype=torch.int32, device=dev)
res = torch.empty(wl.e2e_rows * 5, dtype=torch.uint8, device=dev)
hres = torch.empty(wl.e2e_rows * 5, dtype=torch.uint8).pin_memory()
side = torch.cuda.Stream()
for duplex in (False, True):
    def copies():
        dd.copy_(wl.hd, non_blocking=True), do.copy_(wl.ho, non_blocking=True)
        if duplex:
            with torch.cuda.stream(side):
                hres.copy_(res, non_blocking=True)
        else:
            hres.copy_(res, non_blocking=True)
        torch.cuda.synchronize()
    for _ in range(3):
        copies()
    t0 = time.perf_counter()
    for _ in range(100):
        copies()
    ms = (time.perf_counter() - t0) * 1e3 / 100
    print(f"copies only, {'full duplex' if duplex else 'serial'}: {ms:.3f} ms = {(wl.h2d + wl.d2h) / (ms * 1e-3) / 1e9:.1f} GB/s -> ceiling {wl.e2e_rows / (ms * 1e-3):.4g} records/s", flush=True)
print("ok")
