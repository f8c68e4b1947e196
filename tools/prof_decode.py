"""Small single-launch driver for ncu: one fused decode+score launch on a 2^18-record batch (fmt from argv)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

fmt = sys.argv[1] if len(sys.argv) > 1 else "csv"
wl = bench.DecodeScoreWorkload(fmt, batch_log2=int(sys.argv[2]) if len(sys.argv) > 2 else 18, n_batches=1, repeat=1)
wl.setup(torch.device("cuda", 0), 0)
for i in range(3):
    wl.step(i)
torch.cuda.synchronize()
print("ok", fmt, wl.counters.cpu().tolist())
