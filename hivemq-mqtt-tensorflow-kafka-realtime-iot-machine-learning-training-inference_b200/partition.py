"""Per-GPU partition driver (host side of K5) and the multi-GPU sharding of the scoring path.

The reference reads ONE partition per process (``"{}:0:{}".format(topic, offset)``, cardata-v1.py:10) of a 10-partition
topic (infrastructure/confluent/01_installConfluentPlatform.sh:180).  Here rank ``r`` of ``world_size`` owns partitions
``p`` with ``p % world_size == r``; there is no data-path collective, only the final sum of the four counters
(records, flagged, decode errors, payload bytes) over NCCL (or gloo on CPU for the host logic).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np

N_PARTITIONS = 10  # sensor-data topic


def owned_partitions(rank: int, world_size: int, n_partitions: int = N_PARTITIONS) -> List[int]:
    """Round-robin partition → rank map (p mod G)."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return [p for p in range(n_partitions) if p % world_size == rank]


def balanced_shards(sizes: Sequence[int], world_size: int, sub_batch: int) -> List[List[tuple]]:
    """Split partitions into (partition, first_record, n_records) sub-batches of at most ``sub_batch`` records and deal
    them out greedily (largest first, to the least loaded rank) so that 10 partitions load 8 GPUs evenly."""
    pieces = []
    for p, n in enumerate(sizes):
        for s in range(0, n, sub_batch):
            pieces.append((p, s, min(sub_batch, n - s)))
    pieces.sort(key=lambda t: (-t[2], t[0], t[1]))
    load = [0] * world_size
    out: List[List[tuple]] = [[] for _ in range(world_size)]
    for pc in pieces:
        r = min(range(world_size), key=lambda i: (load[i], i))
        out[r].append(pc)
        load[r] += pc[2]
    for r in range(world_size):
        out[r].sort()
    return out


@dataclass
class RunStats:
    records: int
    flagged: int
    decode_errors: int
    bytes: int
    elapsed_ms: float


def reduce_counters(local: Sequence[int], group=None) -> List[int]:
    """Final count reduction: one all-reduce(sum) of int64[4] over the process group (NCCL over NVLink on GPUs)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return [int(v) for v in local]
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    t = torch.tensor([int(v) for v in local], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return [int(v) for v in t.cpu().tolist()]


class PartitionDriver:
    """Runs decode+score over this rank's partition batches with ``b2_partition_run`` (C ABI)."""

    def __init__(self, model, csv_kinds=None, csv_defaults=None, avro_schema=None, threshold: float = 5.0, n_streams: int = 2):
        from . import _ffi, avro_schema as av

        if (csv_kinds is None) == (avro_schema is None):
            raise ValueError("give exactly one of csv_kinds / avro_schema")
        self.model, self.threshold, self.n_streams = model, float(threshold), int(n_streams)
        self._csv = _ffi.make_csv_spec(csv_kinds, csv_defaults) if csv_kinds is not None else None
        self._avro = None
        if avro_schema is not None:
            prog = avro_schema if isinstance(avro_schema, av.AvroProgram) else av.compile_schema(avro_schema)
            self._avro = _ffi.make_avro_prog(prog)

    def run(self, batches, counters=None, repeat: int = 1) -> RunStats:
        """batches: RecordBatch list (device resident).  Per-record outputs land in ``batch.mse`` / ``batch.flag``.
        ``repeat`` > 1 walks the list that many times inside the one C call (a consumer polling the same partitions)."""
        import torch

        from . import _ffi

        l = _ffi.lib()
        dev = batches[0].data.device if batches else torch.device("cuda", torch.cuda.current_device())
        if counters is None:
            counters = torch.zeros(4, dtype=torch.int64, device=dev)
        arr = (_ffi.Partition * max(len(batches), 1))()
        skip = batches[0].skip_prefix if batches else 0
        for i, b in enumerate(batches):
            if b.skip_prefix != skip:
                raise ValueError("all batches of a run must share the same framing prefix")
            if not hasattr(b, "mse") or b.mse is None:
                b.mse = torch.empty(b.n, dtype=torch.float32, device=dev)
                b.flag = torch.empty(b.n, dtype=torch.uint8, device=dev)
            arr[i].bytes_dev, arr[i].offs_dev, arr[i].n = b.data.data_ptr(), b.offsets.data_ptr(), b.n
            arr[i].mse_dev, arr[i].flag_dev = b.mse.data_ptr(), b.flag.data_ptr()
        nm = self.model._norm_struct()
        st = _ffi.RunStats()
        tot_n = sum(b.n for b in batches)
        tot_bytes = sum(int(b.data.numel()) for b in batches)
        from .keras_compat import set_record_bytes_hint
        set_record_bytes_hint(dev.index, (tot_bytes + max(tot_n, 1) - 1) // max(tot_n, 1))
        torch.cuda.current_stream().synchronize()  # inputs were produced on torch's stream; the driver uses its own
        _ffi.check(l.b2_partition_run_repeat(self.model._handle(), arr, len(batches), C.byref(self._csv) if self._csv is not None else None,
                                             C.byref(self._avro) if self._avro is not None else None, skip, C.byref(nm), self.threshold,
                                             self.n_streams, int(repeat), counters.data_ptr(), C.byref(st)), "b2_partition_run_repeat")
        return RunStats(int(st.records), int(st.flagged), int(st.decode_errors), int(st.bytes), float(st.elapsed_ms))
