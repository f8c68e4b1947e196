"""Data-parallel training check (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/dp_fit_check.py [global_batch] [steps]
Every rank computes gradient sums on ITS slice of each global batch; the sum over GPUs and the Adam update happen inside the same
kernel (b2_ae_step_fused: peer stores into symmetric memory), or — B2_DP_NCCL=1, or no symmetric memory — as gradient kernel,
NCCL all-reduce of the P+3 floats, Adam kernel.  Checks: weights identical on all ranks, and equal (to float32 reduction order)
to a single-GPU run over the full global batches.  Prints one JSON line on rank 0."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import b200_sensor as bs
import bench


def model():
    m = bench.ScoreWorkload(30, 20, 20).model()
    m.set_flat_weights(bs.keras_compat.init_dense_weights(bench.W30, seed=314))
    m.compile(optimizer="adam", loss="mean_squared_error", metrics=["accuracy"])
    return m


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    global_batch = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    x = bs.synth.credit_rows(global_batch * steps, seed=314, outlier_frac=0.0)
    per = global_batch // world
    # rank r owns rows [r*per, (r+1)*per) of every global batch
    xs = x.reshape(steps, world, per, 30)[:, rank].reshape(steps * per, 30)
    m = model()
    xs_d = torch.from_numpy(np.ascontiguousarray(xs)).cuda()
    warm = model()
    warm.fit(xs_d[: 2 * per], epochs=1, batch_size=per, shuffle=False, data_parallel=True)  # symmetric-memory set-up, first launches
    m._exch = getattr(warm, "_exch", None)  # same exchange buffers (sequence numbers continue)
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    h = m.fit(xs_d, epochs=1, batch_size=per, shuffle=False, data_parallel=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    w = torch.from_numpy(m.get_flat_weights()).cuda()
    ws = [torch.empty_like(w) for _ in range(world)]
    dist.all_gather(ws, w)
    same = all(torch.equal(ws[0], t) for t in ws)
    out = {"world": world, "global_batch": global_batch, "steps": steps, "identical_on_all_ranks": bool(same), "path": m.dp_path,
           "symmetric_memory_error": getattr(warm, "_exch_error", None),
           "s_per_step": dt / steps, "samples_per_s": global_batch * steps / dt, "loss": h.history["loss"][-1]}
    if rank == 0:
        ref = model()
        hr = ref.fit(torch.from_numpy(x).cuda(), epochs=1, batch_size=global_batch, shuffle=False)
        wr = ref.get_flat_weights()
        out["max_abs_weight_diff_vs_single_gpu"] = float(np.abs(wr - m.get_flat_weights()).max())
        out["loss_single_gpu"] = hr.history["loss"][-1]
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if not same:
        sys.exit(1)


if __name__ == "__main__":
    main()
