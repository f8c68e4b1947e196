"""Data-parallel fit over NCCL: needs ≥ 2 GPUs, spawns one rank per GPU with torchrun (tools/dp_fit_check.py)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_dp_fit_matches_single_gpu(cuda_device):
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run under gpurun --gpus 2)")
    n = 2 if n < 4 else 4
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "tools", "dp_fit_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["identical_on_all_ranks"]
    assert d["max_abs_weight_diff_vs_single_gpu"] < 2e-5  # same gradient sums, different float32 reduction order
    assert abs(d["loss"] - d["loss_single_gpu"]) < 1e-5 * max(1.0, abs(d["loss_single_gpu"]))
    print(d)
