"""Shared helpers for the GPU parity tests: drive n emulated ranks on one device and
compare every stage with the CPU oracle."""
from __future__ import annotations

import numpy as np


def ocfg(O, cfg):
    return O.Config(cfg.network_fanout_bits, cfg.local_fanout_bits, cfg.payload_bits, cfg.enable_two_level,
                    1 if cfg.flags & 4 else 0)  # RHJ_FLAG_BALANCED_ASSIGNMENT


def to_input(arr, tuple_bytes):
    """TUPLE array -> what the library reads (16-byte tuples, or packed key<<32|rid)."""
    if tuple_bytes == 16:
        return np.ascontiguousarray(arr)
    return ((arr["key"] << np.uint64(32)) | arr["rid"]).astype(np.uint64)


class GpuRun:
    """All ranks of one join emulated in this process on cuda:0 (peer windows are plain
    device pointers; kernels of different ranks run back to back, never concurrently
    waiting on each other)."""

    def __init__(self, rhj, cfg, inner, outer, window_caps=None):
        import torch
        self.rhj, self.cfg, self.n = rhj, cfg, len(inner)
        self.torch = torch
        n = self.n
        tb = cfg.tuple_bytes
        self.inner_n = [len(a) for a in inner]
        self.outer_n = [len(a) for a in outer]
        tot_i, tot_o = sum(self.inner_n), sum(self.outer_n)
        caps = window_caps if window_caps is not None else (tot_i + 64, tot_o + 64)
        self.ctx = [rhj.Context(cfg, r, n, 0) for r in range(n)]
        for c in self.ctx:
            c.plan(max(self.inner_n + [1]), max(self.outer_n + [1]), caps[0], caps[1])
        wins = [c.window_get() for c in self.ctx]
        for c in self.ctx:
            for r in range(n):
                c.window_set_peer(r, wins[r][0], wins[r][1])
        self.d_inner = [torch.from_numpy(to_input(a, tb).view(np.int64).copy()).cuda() for a in inner]
        self.d_outer = [torch.from_numpy(to_input(a, tb).view(np.int64).copy()).cuda() for a in outer]

    def histogram(self):
        for r, c in enumerate(self.ctx):
            c.histogram(self.d_inner[r], self.inner_n[r], self.d_outer[r], self.outer_n[r])

    def offsets(self):
        torch = self.torch
        P = self.ctx[0].P
        parts = [torch.as_tensor(self.rhj.hashjoin._DeviceArray(c.local_histogram_ptr(), 2 * P), device="cuda:0")
                 for c in self.ctx]
        self.gathered = torch.cat(parts).contiguous()
        for c in self.ctx:
            c.compute_offsets(self.gathered if self.n > 1 else None)

    def network(self):
        for c in self.ctx:
            c.network_partition()
        self.torch.cuda.synchronize()

    def local(self):
        for c in self.ctx:
            c.local_partition()

    def probe(self):
        for c in self.ctx:
            c.build_probe()
        return [c.result() for c in self.ctx]

    def run_all(self):
        self.histogram()
        self.offsets()
        self.network()
        self.local()
        return self.probe()

    def close(self):
        for c in self.ctx:
            c.close()


def check_against_oracle(rhj, O, cfg, inner, outer, deep=True):
    """Run the CUDA path stage by stage and compare with the oracle.  Returns the per-rank results."""
    oc = ocfg(O, cfg)
    n = len(inner)
    run = GpuRun(rhj, cfg, inner, outer)
    try:
        run.histogram()
        run.offsets()
        plans = []
        assignment = None
        if oc.balanced:
            glob = [np.sum([O.local_histogram(t, oc) for t in rels], axis=0, dtype=np.uint64) for rels in (inner, outer)]
            assignment = O.balanced_assignment(glob[0], glob[1], n)
        for which, rels in ((0, inner), (1, outer)):
            loc = np.stack([O.local_histogram(t, oc) for t in rels])
            pl = O.plan(loc, n, assignment)
            plans.append(pl)
            for r, c in enumerate(run.ctx):
                assert np.array_equal(c.get_local_histogram(which), loc[r]), f"local histogram rel {which} rank {r}"
                assert np.array_equal(c.get_global_histogram(which), pl.global_), f"global histogram rel {which}"
                assert np.array_equal(c.get_base_offsets(which), pl.base), f"base offsets rel {which}"
                assert np.array_equal(c.get_write_offsets(which), pl.abs_[r]), f"write offsets rel {which} rank {r}"
                assert np.array_equal(c.get_assignment(), pl.assignment)
        run.network()
        run.local()
        if deep:
            for which, rels in ((0, inner), (1, outer)):
                pl = plans[which]
                wins = O.network_partition(rels, oc, pl)
                for r, c in enumerate(run.ctx):
                    for p in range(c.P):
                        if pl.assignment[p] != r:
                            continue
                        want = wins[r][int(pl.base[p]):int(pl.base[p] + pl.global_[p])]
                        got = c.get_partition(which, p, False)
                        assert len(got) == len(want)
                        assert np.array_equal(np.sort(got), np.sort(want)), f"window rel {which} rank {r} part {p}"
                        if cfg.enable_two_level:
                            h, pf, out = O.local_partition(want, oc)
                            gh, gpf = c.get_local_partition(which, p)
                            assert np.array_equal(gh, h), f"local hist rel {which} part {p}"
                            assert np.array_equal(gpf, pf), f"local prefix rel {which} part {p}"
                            gout = c.get_partition(which, p, True)
                            assert len(gout) == len(out)
                            for l in range(c.L):
                                a, b = int(pf[l]), int(pf[l] + h[l])
                                assert np.array_equal(np.sort(gout[a:b]), np.sort(out[a:b])), \
                                    f"sub-partition rel {which} part {p} sub {l}"
        res = run.probe()
        want = O.join(inner, outer, oc)
        for r in range(n):
            assert res[r].astuple() == want[r].astuple(), f"rank {r}: {res[r].astuple()} != {want[r].astuple()}"
        return res
    finally:
        run.close()
