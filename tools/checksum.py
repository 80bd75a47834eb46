"""Expected join result of a generated workload, computed on the device independently of the join.

Checker code for bench.py and the full-size GPU tests (sizes the CPU oracle cannot reach): a dense
position table pos[key_R] = rid_R over all ranks' inner tuples, gathered at this rank's outer keys,
then the five result words (count, sum key, sum rid_R, sum rid_S, sum mix) in wrap-around int64
arithmetic == arithmetic modulo 2^64.  The mix term restates rhj_device.cuh mix_triple, which is
bit-identical to oracle/hpcjoin_oracle.c orc_mix (tests/test_oracle_props.py pins that one).
Requires unique inner keys in [0, key_domain) and every outer key present in the inner relation.
"""
from __future__ import annotations

M64 = (1 << 64) - 1


def expected_result(torch, dist, d_inner, n_inner, d_outer, n_outer, key_domain, world, rank, dev):
    """d_inner / d_outer: int64 tensors of 16-byte tuples {key, rid}; returns five Python ints, all-reduced."""
    pos = torch.empty(key_domain, dtype=torch.int64, device=dev)
    for r in range(world):
        blk = d_inner[:2 * n_inner].clone() if r == rank else torch.empty(2 * n_inner, dtype=torch.int64, device=dev)
        if world > 1:
            dist.broadcast(blk, src=r)
        v = blk.view(-1, 2)
        pos[v[:, 0]] = v[:, 1]
        del blk, v
    so = d_outer[:2 * n_outer].view(-1, 2)
    tot = torch.zeros(5, dtype=torch.int64, device=dev)
    step, lo32 = 1 << 24, 0xFFFFFFFF
    for b in range(0, n_outer, step):
        key, rs = so[b:b + step, 0], so[b:b + step, 1]
        rr = pos[key]
        k32 = (key ^ (key >> 32)) & lo32
        r32 = (rr ^ (rr >> 32)) & lo32
        s32 = (rs ^ (rs >> 32)) & lo32
        mix = (r32 ^ 0x9E3779B9) * (s32 ^ 0x85EBCA6B) + (k32 ^ 0xC2B2AE35) * ((r32 + s32 + 0x27D4EB2F) & lo32)
        tot += torch.stack([torch.tensor(key.numel(), device=dev), key.sum(), rr.sum(), rs.sum(), mix.sum()])
    del pos
    if world > 1:
        dist.all_reduce(tot)
    return [int(x) & M64 for x in tot.tolist()]


def reduce_result(torch, dist, res, world, dev):
    """All-reduce one rank's rhj_result into the job's five words."""
    tot = torch.tensor([x - (1 << 64) if x >= (1 << 63) else x for x in res.astuple()], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    return [int(x) & M64 for x in tot.tolist()]
