"""Launcher-side mirror of the reference's operator surface.

`HashJoin(numberOfNodes, nodeId, inner, outer).join()` has the argument meaning of
hpcjoin::operators::HashJoin (operators/HashJoin.h:23-43); the phases run in the
order of operators/HashJoin.cpp:45-220, each phase one call through the C ABI.
With more than one rank (one process per GPU, torch.distributed/NCCL) the only
communication is: one all-gather of the per-rank histograms (replaces
MPI_Allreduce + MPI_Exscan), the library's own peer stores into the other ranks'
receive windows (replace MPI_Put) and, per relation, one barrier once every
rank's stores have landed (operators/HashJoin.cpp:119-121).
"""
from __future__ import annotations

import ctypes as C


class _DeviceArray:
    """Expose a raw device pointer to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, n: int, typestr: str = "<i8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


class Relation:
    """Device-resident relation: `local_size` tuples of this rank out of `global_size`
    (data/Relation.h:18-60).  `data` is a torch tensor (int64) holding Tuple{key,rid}
    pairs (tuple_bytes 16) or packed 8-byte tuples."""

    def __init__(self, data, local_size: int, global_size: int, tuple_bytes: int = 16):
        self.data, self.local_size, self.global_size, self.tuple_bytes = data, local_size, global_size, tuple_bytes

    @staticmethod
    def from_numpy(arr, global_size: int | None = None, device: int = 0) -> "Relation":
        import numpy as np
        import torch
        tb = 16 if arr.dtype.itemsize == 16 else 8
        flat = np.ascontiguousarray(arr).view(np.int64)
        t = torch.from_numpy(flat).to(f"cuda:{device}")
        return Relation(t, len(arr), global_size if global_size is not None else len(arr), tb)

    def getLocalSize(self):
        return self.local_size

    def getGlobalSize(self):
        return self.global_size

    def getData(self):
        return self.data


def exchange_windows(ctx, group=None):
    """Collective: make every rank's receive windows addressable from this rank
    (replaces MPI_Win_create, data/Window.cpp:35-45) via CUDA IPC."""
    import torch.distributed as dist
    from . import WindowHandle
    if ctx.nranks == 1:
        return
    mine = bytes(ctx.window_export())
    handles = [None] * ctx.nranks
    dist.all_gather_object(handles, mine, group=group)
    for r, raw in enumerate(handles):
        if r != ctx.rank:
            ctx.window_import(r, WindowHandle.from_buffer_copy(raw))


_HOST_GROUPS = {}


def _host_barrier_group(group=None):
    """A gloo group over the same ranks, used for the exchange barrier when RHJ_HOST_BARRIER is
    set.  Measured on 8 GPUs it makes no difference to the step time (13.66 vs 13.69 ms), so the
    default is the NCCL all-reduce on the compute stream, which needs no host synchronisation."""
    import os
    import torch.distributed as dist
    if not os.environ.get("RHJ_HOST_BARRIER"):
        return None
    key = id(group)
    if key not in _HOST_GROUPS:
        try:
            ranks = dist.get_process_group_ranks(group) if group is not None else None
            _HOST_GROUPS[key] = dist.new_group(ranks=ranks, backend="gloo")
        except Exception:
            _HOST_GROUPS[key] = None
    return _HOST_GROUPS[key]


def join_multi_rank(ctx, inner: Relation, outer: Relation, group=None, pairs=None, capacity=None):
    """operators/HashJoin.cpp:45-220 for this rank; every rank calls it (collective).  With `pairs`
    (device int64 tensor, 2 values per pair) the build/probe stage also writes this rank's
    (rid_R, rid_S) pairs (rhj_build_probe_pairs)."""
    import torch
    import torch.distributed as dist
    ctx.histogram(inner.data, inner.local_size, outer.data, outer.local_size)
    if ctx.nranks > 1:
        local = torch.as_tensor(_DeviceArray(ctx.local_histogram_ptr(), 2 * ctx.P), device=f"cuda:{ctx.device}")
        gathered = torch.empty(ctx.nranks * 2 * ctx.P, dtype=torch.int64, device=local.device)
        dist.all_gather_into_tensor(gathered, local, group=group)
        ctx.compute_offsets(gathered)
    else:
        ctx.compute_offsets(None)
    ctx.network_partition()
    if ctx.nranks > 1:
        # Per relation: my pushes are complete (exchange_wait), then everybody's are (the
        # all-reduce is the barrier of operators/HashJoin.cpp:119-121).  The inner relation's
        # local pass then runs while the outer relation is still crossing NVLink.
        host_group = _host_barrier_group(group)
        token = torch.zeros(1, dtype=torch.int32, device=f"cuda:{ctx.device}")
        for which in (0, 1):
            if host_group is not None:
                # optional host-side barrier (RHJ_HOST_BARRIER): host waits for its own pushes,
                # then meets the other ranks over gloo instead of NCCL
                ctx.exchange_sync(which)
                dist.barrier(group=host_group)
            else:
                ctx.exchange_wait(which)
                dist.all_reduce(token, group=group)
            ctx.local_partition_rel(which)
    else:
        ctx.local_partition()
    if pairs is not None:
        ctx.build_probe_pairs(pairs, capacity)
    else:
        ctx.build_probe()


class HashJoin:
    """hpcjoin::operators::HashJoin (operators/HashJoin.h:23-43).  RESULT_COUNTER is the
    class-level match count of the last join on this rank, as in the reference
    (operators/HashJoin.cpp:28); RESULT_CHECKSUM is this project's addition."""

    RESULT_COUNTER = 0
    RESULT_CHECKSUM = (0, 0, 0, 0)

    def __init__(self, numberOfNodes: int, nodeId: int, innerRelation: Relation, outerRelation: Relation,
                 config=None, device: int | None = None, group=None, window_caps=(0, 0)):
        from . import Context, make_config
        import torch
        self.numberOfNodes, self.nodeId = numberOfNodes, nodeId
        self.innerRelation, self.outerRelation = innerRelation, outerRelation
        self.group = group
        cfg = config if config is not None else make_config(tuple_bytes=innerRelation.tuple_bytes)
        dev = device if device is not None else torch.cuda.current_device()
        self.ctx = Context(cfg, nodeId, numberOfNodes, dev)
        self.ctx.plan(innerRelation.local_size, outerRelation.local_size, *window_caps)
        exchange_windows(self.ctx, group)

    def join(self, pairs=None, capacity=None):
        """`pairs`: optional device int64 tensor that receives this rank's (rid_R, rid_S) pairs."""
        join_multi_rank(self.ctx, self.innerRelation, self.outerRelation, self.group, pairs, capacity)
        r = self.ctx.result()
        HashJoin.RESULT_COUNTER = r.matches
        HashJoin.RESULT_CHECKSUM = (r.sum_key, r.sum_rid_r, r.sum_rid_s, r.sum_mix)
        return r

    def close(self):
        self.ctx.close()
