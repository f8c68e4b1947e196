"""LSTM forward throughput: tcgen05 kernel vs FP32 FMA kernel (B2_LSTM_TC=1/0), T in {1, 4, 16}."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, b200_sensor as bs
dev = torch.device("cuda", 0)
for tc in ("2", "1", "0"):
    bs.keras.set_option("lstm_tc", int(tc))
    for T in (1, 4, 16):
        r = bench.aux_lstm(dev, T)
        print(f"B2_LSTM_TC={tc} T={T:2d} B={r['batch']:8d}  {r['value']:.4g} record-steps/s  {r['fp32_tflops']:.1f} TFLOP/s-equivalent")
