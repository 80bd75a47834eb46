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


def exact_window_caps(cfg, nranks: int, rank: int, inner: Relation, outer: Relation, group=None):
    """Receive-window sizes as the reference computes them: after a histogram of the actual relations
    (data/Window.cpp:160-178 sizes every window from assignment + global histogram, so it can never
    overflow, however skewed the keys are).  One extra counting pass over the keys at set-up time,
    done with torch (launcher plumbing, outside every timed join); returns one pair of capacities
    that fits every rank, since all ranks must plan the same capacities."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from . import lib, FLAG_BALANCED_ASSIGNMENT
    P = 1 << cfg.network_fanout_bits
    hists = []
    for rel in (inner, outer):
        n = rel.local_size
        h = torch.zeros(P, dtype=torch.int64, device=rel.data.device)
        step = 1 << 24
        for b in range(0, n, step):
            e = min(n, b + step)
            keys = rel.data[2 * b:2 * e:2] if rel.tuple_bytes == 16 else (rel.data[b:e] >> 32)
            h += torch.bincount(keys & (P - 1), minlength=P)
        g = [torch.zeros_like(h) for _ in range(nranks)]
        if nranks > 1:
            dist.all_gather(g, h, group=group)
        else:
            g = [h]
        hists.append(np.ascontiguousarray(torch.stack(g).cpu().numpy().astype(np.uint64)))
    assign = np.zeros(P, np.uint32)
    if cfg.flags & FLAG_BALANCED_ASSIGNMENT:
        gi = np.ascontiguousarray(hists[0].sum(axis=0, dtype=np.uint64))
        go = np.ascontiguousarray(hists[1].sum(axis=0, dtype=np.uint64))
        lib().rhj_host_assignment_balanced(gi.ctypes.data, go.ctypes.data, nranks, P, assign.ctypes.data)
    else:
        lib().rhj_host_assignment(nranks, P, assign.ctypes.data)
    caps = []
    for k in range(2):
        ws = np.zeros(nranks, np.uint64)
        lib().rhj_host_offsets_assigned(hists[k].ctypes.data, assign.ctypes.data, nranks, rank, P, None, None, None, None,
                                        ws.ctypes.data)
        caps.append(int(ws.max()) + 64)
    return tuple(caps)


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


def _splitmix64(x: int) -> int:
    m = (1 << 64) - 1
    x = (x + 0x9E3779B97F4A7C15) & m
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & m
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & m
    return x ^ (x >> 31)


def perm_positions(n: int, seed: int, device):
    """perm(i) for i in [0, n): the generator's bijection (rhj_fill_perm / oracle orc_perm) evaluated
    with torch integer ops, so that it runs wherever the relation lives (cuda, or cpu under gloo)."""
    import torch
    x = torch.arange(n, dtype=torch.int64, device=device)
    if n <= 1:
        return torch.zeros_like(x)
    b = max(1, (n - 1).bit_length())
    mask, sh = (1 << b) - 1, max(1, b // 2)
    s64 = lambda v: v - (1 << 64) if v >= (1 << 63) else v  # two's complement: int64 products wrap like u64
    ma = [(s64(_splitmix64(seed + 2 * i) | 1), s64(_splitmix64(seed + 2 * i + 1))) for i in range(3)]

    def rounds(v):
        for m_, a_ in ma:
            v = (v * m_ + a_) & mask
            v = v ^ (v >> sh)
        return v

    x = rounds(x)
    while True:  # cycle walking: values that left [0, n) go round again
        out = x >= n
        if not bool(out.any()):
            return x
        x = torch.where(out, rounds(x), x)


def distribute(rel: Relation, nranks: int, rank: int, seed: int, group=None) -> None:
    """Relation::distribute (data/Relation.cpp:99-141) on the device, in place, for relations of the
    harness's counter-based generator: this rank swaps section (rank + i) % n with
    rank i -- tuples cross NVLink through point-to-point NCCL transfers, no host copy of the relation
    exists -- and then permutes its keys over the positions with the generator's bijection where the
    reference shuffles them with rand() (keys move, rids stay, data/Relation.cpp:87-97).  Every rank
    calls it (collective); all ranks hold the same number of 16-byte tuples.  Bit-identical to
    oracle.distribute_perm (orc_distribute_perm)."""
    import torch
    import torch.distributed as dist
    if rel.tuple_bytes != 16:
        raise ValueError("distribute works on 16-byte {key, rid} tuples")
    n = rel.local_size
    t = rel.data[:2 * n].view(n, 2)
    if nranks > 1:
        sec = n // nranks
        last = n - (nranks - 1) * sec
        ops, bufs = [], []
        for i in range(nranks):
            if i == rank:
                continue
            s_ = (rank + i) % nranks
            lo, size = s_ * sec, (last if s_ == nranks - 1 else sec)
            if size == 0:
                continue
            send = t[lo:lo + size].contiguous()
            recv = torch.empty_like(send)
            peer = i if group is None else dist.get_global_rank(group, i)
            ops.append(dist.P2POp(dist.isend, send, peer, group))
            ops.append(dist.P2POp(dist.irecv, recv, peer, group))
            bufs.append((lo, size, recv, send))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        for lo, size, recv, _ in bufs:
            t[lo:lo + size] = recv
    t[:, 0] = t[perm_positions(n, seed + rank, t.device), 0]


_SIDE_STREAMS = {}


def _side_stream(ctx):
    import torch
    key = id(ctx)
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = torch.cuda.Stream(device=ctx.device)
    return _SIDE_STREAMS[key]


def overlap_sms(nranks: int, sm_count: int = 148) -> int:
    """SMs given to the outer relation's network pass while the inner relation's local pass runs
    beside it (0 = no overlap).  With two ranks half of the tuples stay at home and the pass is
    HBM-bound like the local pass, so running them side by side gains nothing; from four ranks on
    the pass waits for NVLink and half of the SMs are enough to keep the links busy."""
    return 0 if nranks < 3 else sm_count // 2


def join_multi_rank(ctx, inner: Relation, outer: Relation, group=None, pairs=None, capacity=None, host=None,
                    overlap=None):
    """operators/HashJoin.cpp:45-220 for this rank; every rank calls it (collective).  With `pairs`
    (device int64 tensor, 2 values per pair) the build/probe stage also writes this rank's
    (rid_R, rid_S) pairs (rhj_build_probe_pairs).  `host` = (inner_host, outer_host): the relations
    are taken from host memory through rhj_stage_host_inputs instead of `inner.data`/`outer.data`.
    `overlap`: SMs of the outer relation's network pass (None = overlap_sms(nranks))."""
    import os
    import torch
    import torch.distributed as dist
    if host is not None:
        d_in, d_out = ctx.stage_host_inputs(host[0], inner.local_size, host[1], outer.local_size)
    else:
        d_in, d_out = inner.data, outer.data
    from . import FLAG_BALANCED_ASSIGNMENT
    # per-relation histogram exchange (inner relation scattered and pushed while the outer one is counted):
    # built and parity-tested, but measured slower on 8 GPUs (12.5 against 11.7 ms per join), so opt-in
    per_relation = (bool(os.environ.get("RHJ_PER_RELATION_EXCHANGE")) and ctx.nranks > 1 and ctx.is_staged()
                    and not (ctx.cfg.flags & FLAG_BALANCED_ASSIGNMENT))
    if not per_relation:
        ctx.histogram(d_in, inner.local_size, d_out, outer.local_size)
    if ctx.nranks == 1:
        ctx.compute_offsets(None)
        ctx.network_partition()
        ctx.local_partition()
    else:
        dev = f"cuda:{ctx.device}"
        if per_relation:
            # Histogram, all-gather and offsets one relation at a time: the inner relation is already being
            # scattered and pushed over NVLink while the outer relation's histogram is computed.
            rels = ((d_in, inner.local_size), (d_out, outer.local_size))
            for which in (0, 1):
                if which == 1:
                    ctx.set_sm_limit(ctx.sm_count() - 32)  # the pushes of the inner relation keep their SMs
                ctx.histogram_rel(which, *rels[which])
                ctx.set_sm_limit(0)
                local = torch.as_tensor(_DeviceArray(ctx.local_histogram_ptr() + which * ctx.P * 8, ctx.P), device=dev)
                gathered = torch.empty(ctx.nranks * ctx.P, dtype=torch.int64, device=dev)
                dist.all_gather_into_tensor(gathered, local, group=group)
                ctx.compute_offsets_rel(which, gathered)
                if which == 0:
                    ctx.network_partition_rel(0)
        else:
            local = torch.as_tensor(_DeviceArray(ctx.local_histogram_ptr(), 2 * ctx.P), device=dev)
            gathered = torch.empty(ctx.nranks * 2 * ctx.P, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(gathered, local, group=group)
            ctx.compute_offsets(gathered)
        # Per relation: my stores into the peers' windows are complete (exchange_wait), then
        # everybody's are (the all-reduce is the barrier of operators/HashJoin.cpp:119-121).
        host_group = _host_barrier_group(group)
        token = torch.zeros(1, dtype=torch.int32, device=dev)

        use_nccl_barrier = not ctx.has_flags() or bool(os.environ.get("RHJ_NCCL_BARRIER"))

        def barrier(which):
            if not use_nccl_barrier and host_group is None:
                ctx.exchange_wait(which)  # waits for this rank's stores and for every peer's arrival flag
            elif host_group is not None:  # optional host-side barrier (RHJ_HOST_BARRIER)
                ctx.exchange_sync(which)
                dist.barrier(group=host_group)
            else:
                ctx.exchange_wait(which)
                dist.all_reduce(token, group=group)

        staged = ctx.is_staged()
        if overlap is None:
            sms = overlap_sms(ctx.nranks)
        else:
            sms = overlap
        if sms <= 0 or host_group is not None:
            if not per_relation:
                ctx.network_partition_rel(0)
            ctx.network_partition_rel(1)
            for which in (0, 1):
                barrier(which)
                ctx.local_partition_rel(which)
        else:
            # The inner relation's wait and local pass run on a second stream beside the outer
            # relation's network pass.  Direct exchange: each gets its share of the SMs (`sms` for the
            # network pass).  Staged exchange: the pushes (library side stream) hold their own SMs, the
            # two passes share the rest as the block scheduler hands them out.
            main, side = torch.cuda.current_stream(), _side_stream(ctx)
            if not per_relation:
                ctx.network_partition_rel(0)
            side.wait_stream(main)
            if not staged:
                ctx.set_sm_limit(sms)
            ctx.network_partition_rel(1)
            with torch.cuda.stream(side):
                barrier(0)
                if not staged:
                    ctx.set_sm_limit(ctx.sm_count() - sms)
                ctx.local_partition_rel(0)
            ctx.set_sm_limit(0)
            barrier(1)
            main.wait_stream(side)
            ctx.local_partition_rel(1)
    if pairs is not None:
        ctx.build_probe_pairs(pairs, capacity)
    else:
        ctx.build_probe()


def host_pieces(outer_host, piece_tuples: int, device):
    """Generator over device-resident pieces of a host relation (numpy array of 16-byte tuples or of packed
    u64 values): two staging tensors and a copy stream, so that piece i+1 crosses PCIe while piece i is
    joined.  A staging tensor is reused two pieces later, after the join stream has passed that point."""
    import numpy as np
    import torch
    words = 2 if outer_host.dtype.itemsize == 16 else 1
    flat = torch.from_numpy(np.ascontiguousarray(outer_host).view(np.int64))
    n = len(outer_host)
    bufs = [torch.empty(piece_tuples * words, dtype=torch.int64, device=device) for _ in range(2)]
    free = [torch.cuda.Event() for _ in range(2)]
    main, copy = torch.cuda.current_stream(), torch.cuda.Stream(device=device)
    for i, b in enumerate(range(0, n, piece_tuples)):
        m = min(piece_tuples, n - b)
        buf = bufs[i & 1]
        if i >= 2:
            copy.wait_event(free[i & 1])  # the join of piece i-2 no longer reads this buffer
        with torch.cuda.stream(copy):
            buf[:m * words].copy_(flat[b * words:(b + m) * words], non_blocking=True)
            arrived = torch.cuda.Event()
            arrived.record(copy)
        main.wait_event(arrived)
        yield buf, m
        free[i & 1].record(main)


def join_probe_pieces(ctx, inner: Relation, pieces, group=None):
    """HashJoin::join() with the probe side supplied piece by piece: the inner relation is
    partitioned once and stays resident, every piece of the outer relation (device tensor, tuples) goes
    through histogram, network pass, local pass and build/probe against it, and the results add up.  Only
    the inner relation and one piece have to fit the context's plan.  Collective with more than one rank
    (every rank supplies the same number of pieces).  Returns the rank's result of the whole join."""
    import torch
    import torch.distributed as dist
    dev = f"cuda:{ctx.device}"
    use_barrier = ctx.nranks > 1 and not ctx.has_flags()
    token = torch.zeros(1, dtype=torch.int32, device=dev) if use_barrier else None

    def phase(which, data, n):
        ctx.histogram_rel(which, data, n)
        if ctx.nranks > 1:
            local = torch.as_tensor(_DeviceArray(ctx.local_histogram_ptr() + which * ctx.P * 8, ctx.P), device=dev)
            gathered = torch.empty(ctx.nranks * ctx.P, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(gathered, local, group=group)
            ctx.compute_offsets_rel(which, gathered)
        else:
            ctx.compute_offsets_rel(which, None)
        ctx.network_partition_rel(which)
        ctx.exchange_wait(which)
        if use_barrier:
            dist.all_reduce(token, group=group)
        ctx.local_partition_rel(which)

    phase(0, inner.data, inner.local_size)
    first = True
    for data, n in pieces:
        phase(1, data, n)
        if first:
            ctx.build_probe()
        else:
            ctx.build_probe_accumulate()
        first = False
    if first:  # no pieces at all: an empty outer relation
        phase(1, None, 0)
        ctx.build_probe()
    return ctx.result()


def hot_partitions(global_outer, nranks: int, inv_share: int = 4):
    """The skew rule, identical on every rank (second half), in integer arithmetic: a network
    partition is HOT when its probe-side tuple count exceeds 1/inv_share of the mean probe-side load of a rank,
    i.e. global_outer[p] * nranks * inv_share > sum(global_outer) -- a single partition that alone is a quarter
    (by default) of what a rank should carry.  Uniform keys never trigger it (a partition is 1/P of the
    relation, and P >= 4 n in every configuration this repo runs).  inv_share <= 0 switches splitting off.
    `global_outer`: int64[P] (torch tensor or numpy array).  Returns a boolean mask over partitions."""
    if inv_share <= 0:
        return global_outer < 0
    return global_outer * (nranks * inv_share) > global_outer.sum()


def split_hot(rel: Relation, hot_mask, P: int):
    """(hot part, cold part) of a relation of 16-byte tuples as two new int64 tensors [m, 2]."""
    t = rel.data[:2 * rel.local_size].view(-1, 2)
    is_hot = hot_mask[t[:, 0] & (P - 1)]
    return t[is_hot].contiguous(), t[~is_hot].contiguous()


def gather_rows(local_rows, nranks: int, group=None):
    """All ranks' rows ([m_r, 2] int64, m_r differs per rank) concatenated in rank order on every rank."""
    import torch
    import torch.distributed as dist
    if nranks == 1:
        return local_rows
    cnt = torch.tensor([local_rows.shape[0]], dtype=torch.int64, device=local_rows.device)
    cnts = [torch.zeros_like(cnt) for _ in range(nranks)]
    dist.all_gather(cnts, cnt, group=group)
    sizes = [int(c.item()) for c in cnts]
    pad = max(max(sizes), 1)
    mine = torch.zeros((pad, 2), dtype=torch.int64, device=local_rows.device)
    mine[:local_rows.shape[0]] = local_rows
    bufs = [torch.empty_like(mine) for _ in range(nranks)]
    dist.all_gather(bufs, mine, group=group)
    return torch.cat([b[:m] for b, m in zip(bufs, sizes)])


class HotSplitJoin:
    """HashJoin with hot-partition splitting (the part a partition -> rank assignment cannot
    do): the probe side of a hot partition is NOT sent to the partition's owner -- every rank keeps its own
    tuples of it -- and the partition's build side is replicated to all ranks instead, so a hot key's matches
    are spread like the probe tuples themselves.  What the reference's un-built kernels do inside one GPU
    (skew_detect / generate_block_mapping / probe_skew, operators/gpu/kernels_optimized.cu:301-364), done across
    ranks.  Built from the library's own stages: the cold partitions go through the ordinary multi-rank join,
    the hot ones through a second, single-rank context per GPU (inner = the replicated build side, outer =
    this rank's probe tuples); the two results add up.  Splitting and replication are launcher plumbing
    (torch indexing, one all-gather of the hot build tuples); `hot_partitions` is the rule.
    16-byte tuples; every rank constructs and calls it alike."""

    def __init__(self, numberOfNodes: int, nodeId: int, innerRelation: Relation, outerRelation: Relation,
                 config=None, device=None, group=None, inv_share: int = 4):
        import copy
        import torch
        import torch.distributed as dist
        from . import Context, make_config
        self.n, self.rank, self.group, self.inv_share = numberOfNodes, nodeId, group, inv_share
        self.inner, self.outer = innerRelation, outerRelation
        cfg = config if config is not None else make_config()
        self.cfg = cfg
        self.P = 1 << cfg.network_fanout_bits
        dev = device if device is not None else torch.cuda.current_device()
        self.device = dev
        # the rule needs the global probe-side histogram: one counting pass at construction, as for the windows
        keys = outerRelation.data[:2 * outerRelation.local_size:2]
        h = torch.bincount(keys & (self.P - 1), minlength=self.P)
        if numberOfNodes > 1:
            dist.all_reduce(h, group=group)
        self.hot = hot_partitions(h, numberOfNodes, inv_share)
        s_hot, s_cold = split_hot(outerRelation, self.hot, self.P)
        r_hot, r_cold = split_hot(innerRelation, self.hot, self.P)
        self.r_hot_all = gather_rows(r_hot, numberOfNodes, group)
        self.s_hot = s_hot
        self.cold_inner = Relation(r_cold.view(-1), r_cold.shape[0], innerRelation.global_size)
        self.cold_outer = Relation(s_cold.view(-1), s_cold.shape[0], outerRelation.global_size)
        # ragged cold sizes per rank: plan the main context for the largest share
        sizes = torch.tensor([r_cold.shape[0], s_cold.shape[0]], dtype=torch.int64, device=r_cold.device)
        if numberOfNodes > 1:
            dist.all_reduce(sizes, op=dist.ReduceOp.MAX, group=group)
        self.ctx = Context(cfg, nodeId, numberOfNodes, dev)
        caps = (0, 0) if numberOfNodes == 1 else exact_window_caps(cfg, numberOfNodes, nodeId, self.cold_inner,
                                                                    self.cold_outer, group)
        self.ctx.plan(max(int(sizes[0].item()), 1), max(int(sizes[1].item()), 1), *caps)
        exchange_windows(self.ctx, group)
        hot_cfg = copy.copy(cfg)
        self.hot_ctx = Context(hot_cfg, 0, 1, dev)
        self.hot_ctx.plan(max(self.r_hot_all.shape[0], 1), max(s_hot.shape[0], 1))

    def join(self):
        from . import Result
        join_multi_rank(self.ctx, self.cold_inner, self.cold_outer, self.group)
        cold = self.ctx.result()
        hot = self.hot_ctx.join(self.r_hot_all.view(-1), self.r_hot_all.shape[0], self.s_hot.view(-1), self.s_hot.shape[0])
        m = (1 << 64) - 1
        out = Result(*[(a + b) & m for a, b in zip(cold.astuple(), hot.astuple())])
        HashJoin.RESULT_COUNTER = out.matches
        HashJoin.RESULT_CHECKSUM = out.astuple()[1:]
        return out

    def close(self):
        self.ctx.close()
        self.hot_ctx.close()


class HashJoin:
    """hpcjoin::operators::HashJoin (operators/HashJoin.h:23-43).  RESULT_COUNTER is the
    class-level match count of the last join on this rank, as in the reference
    (operators/HashJoin.cpp:28); RESULT_CHECKSUM is this project's addition."""

    RESULT_COUNTER = 0
    RESULT_CHECKSUM = (0, 0, 0, 0)

    def __init__(self, numberOfNodes: int, nodeId: int, innerRelation: Relation, outerRelation: Relation,
                 config=None, device: int | None = None, group=None, window_caps=None):
        from . import Context, make_config
        import torch
        self.numberOfNodes, self.nodeId = numberOfNodes, nodeId
        self.innerRelation, self.outerRelation = innerRelation, outerRelation
        self.group = group
        cfg = config if config is not None else make_config(tuple_bytes=innerRelation.tuple_bytes)
        dev = device if device is not None else torch.cuda.current_device()
        self.ctx = Context(cfg, nodeId, numberOfNodes, dev)
        if window_caps is None:
            # the reference sizes each window exactly, after the histogram phase; a static guess
            # (allocation_factor x local size) only holds for near-uniform partition sizes
            window_caps = (0, 0) if numberOfNodes == 1 else \
                exact_window_caps(cfg, numberOfNodes, nodeId, innerRelation, outerRelation, group)
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
