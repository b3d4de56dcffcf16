"""Multi-GPU plumbing: one process per GPU, the table image replicated, packet
batches sharded, and exactly two collectives, both off the per-batch path:

  * broadcast of the committed table image from rank 0 (NCCL over NVLink on GPUs; gloo over host
    memory for the CPU tests, where offline handles carry the image), and
  * a sum all-reduce of the u64 hit counters.

torch.distributed is the plumbing; the library only exposes device/host pointers to the image and
the counters.  No reference counterpart: the reference drives a single DPU per node.
"""
from __future__ import annotations

import time

import numpy as np


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous split of n packets: rank r gets [r*n/world, (r+1)*n/world)."""
    return (rank * n) // world, ((rank + 1) * n) // world


def broadcast_table(dp, rank: int, world: int, src: int = 0) -> float:
    """Replicate rank `src`'s committed image on every rank.  Returns milliseconds spent."""
    import torch
    import torch.distributed as dist

    on_gpu = dp.device >= 0
    dev = torch.device("cuda", dp.device) if on_gpu else torch.device("cpu")
    t0 = time.perf_counter()
    size = torch.tensor([dp.table_size if rank == src else 0], dtype=torch.int64, device=dev)
    dist.broadcast(size, src=src)
    nbytes = int(size.item())
    buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    if rank == src:
        if on_gpu:
            dp.table_export_device(buf.data_ptr(), nbytes, torch.cuda.current_stream().cuda_stream)
        else:
            buf.copy_(torch.from_numpy(dp.table_export()))
    dist.broadcast(buf, src=src)
    if rank != src:
        if on_gpu:
            dp.table_import_device(buf.data_ptr(), nbytes, torch.cuda.current_stream().cuda_stream)
        else:
            dp.table_import(buf.numpy())
    if on_gpu:
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3


def reduce_counters(dp, world: int) -> float:
    """Sum the hit counters over all ranks, in place on the device.  Returns milliseconds."""
    import torch
    import torch.distributed as dist

    from . import consts

    n = consts.CNT_TOTAL
    # view the library's counter array as a torch tensor (no copy): int64 has the same bits as u64
    iface = {"shape": (n,), "typestr": "<i8", "data": (dp.counters_device_ptr, False), "version": 2}

    class _Wrap:
        __cuda_array_interface__ = iface

    t = torch.as_tensor(_Wrap(), device=torch.device("cuda", dp.device))
    torch.cuda.synchronize()
    # the collective itself, timed on the device (host-side wrapping and the first-use set-up of a communicator are
    # not part of what a reporting interval pays)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def reduce_counters_host(counters: np.ndarray) -> np.ndarray:
    """gloo/CPU twin of reduce_counters for the host-side tests."""
    import torch
    import torch.distributed as dist

    t = torch.from_numpy(counters.view(np.int64).copy())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.numpy().view(np.uint64)
