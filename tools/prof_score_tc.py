"""Single-launch driver for ncu: the tcgen05 score kernel (context option score_tc = 1) on 2^22 rows, D from argv (30 | 18)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, b200_sensor as bs
d = int(sys.argv[1]) if len(sys.argv) > 1 else 30
bs.keras.set_option("score_tc", int(sys.argv[2]) if len(sys.argv) > 2 else 1)
wl = bench.ScoreWorkload(d, 22, 20)
wl.setup(torch.device("cuda", 0), 0)
for i in range(3):
    wl.step(i)
wl.finish()
print("ok", d, wl.kernel_ms / 3, "ms per launch")
