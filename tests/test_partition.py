"""Multi-GPU host logic on CPU: partition → rank map, balanced sub-batching, and the final count reduction over a
world_size-2 gloo group (the NCCL path runs the same code on GPUs)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200_sensor import partition as part


def test_owned_partitions_cover_everything_once():
    for world in (1, 2, 4, 8):
        seen = []
        for r in range(world):
            seen += part.owned_partitions(r, world)
        assert sorted(seen) == list(range(10))
    assert part.owned_partitions(1, 8) == [1, 9]
    with pytest.raises(ValueError):
        part.owned_partitions(8, 8)


def test_balanced_shards_even_out_ten_partitions_on_eight_ranks():
    sizes = [1 << 20] * 10
    shards = part.balanced_shards(sizes, 8, sub_batch=1 << 17)
    loads = [sum(n for _, _, n in s) for s in shards]
    assert sum(loads) == sum(sizes) and max(loads) - min(loads) <= 1 << 17
    flat = sorted(pc for s in shards for pc in s)
    covered = {}
    for p, s, n in flat:
        covered.setdefault(p, []).append((s, n))
    for p, segs in covered.items():
        pos = 0
        for s, n in sorted(segs):
            assert s == pos
            pos += n
        assert pos == sizes[p]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = part.owned_partitions(rank, world)
    local = [1000 * len(mine) + rank, 10 * len(mine), rank, 163 * 1000 * len(mine)]
    total = part.reduce_counters(local)
    q.put((rank, mine, total))
    dist.destroy_process_group()


def test_count_reduction_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == [0, 2, 4, 6, 8] and res[1][1] == [1, 3, 5, 7, 9]
    assert res[0][2] == res[1][2] == [10001, 100, 1, 163 * 10000]


def test_reduce_counters_without_process_group_is_identity():
    assert part.reduce_counters([1, 2, 3, 4]) == [1, 2, 3, 4]
