"""Single-kernel driver for ncu: LSTM forward (tcgen05 kernel), B = 2^20 / T sequences."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
T = int(sys.argv[1]) if len(sys.argv) > 1 else 1
print(bench.aux_lstm(torch.device("cuda", 0), T))
