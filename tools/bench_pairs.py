"""Device time of the build/probe stage, counting (rhj_build_probe) against materialising
(rhj_build_probe_pairs), on the bench workload (256 M x 256 M 16-byte tuples, one GPU).
Prints one JSON line; not part of the bench.py contract."""
import argparse
import importlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2n", type=int, default=28)
    ap.add_argument("--steps", type=int, default=10)
    args = ap.parse_args()
    n = 1 << args.log2n
    cfg = rhj.make_config(8, 8, 24, True)
    dr = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    ds = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    rhj.fill_perm(dr, 16, n, 0, n, seed=1)
    rhj.fill_perm(ds, 16, n, 0, n, seed=2)
    ctx = rhj.Context(cfg)
    ctx.plan(n, n)
    ctx.histogram(dr, n, ds, n)
    ctx.compute_offsets(None)
    ctx.network_partition()
    ctx.local_partition()
    del dr, ds
    pairs = torch.empty(2 * n, dtype=torch.int64, device="cuda")

    def timed(fn):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.steps

    t_count = timed(ctx.build_probe)
    r0 = ctx.result()
    t_nostore = timed(lambda: ctx.build_probe_pairs(pairs, 0))  # capacity 0: everything but the stores
    t_pairs = timed(lambda: ctx.build_probe_pairs(pairs))
    r1 = ctx.result()
    assert r0.astuple() == r1.astuple() and r0.matches == n
    p = pairs.view(-1, 2)
    assert int(p[:, 1].sum()) == n * (n - 1) // 2 and int(p[:, 0].sum()) == n * (n - 1) // 2
    read_b, write_b = 16 * n, 16 * n
    print(json.dumps({
        "workload": f"{n} x {n} 16-byte tuples, 8+8 radix bits, one GPU, stage K4 only",
        "count_only_ms": t_count, "materialise_ms": t_pairs, "materialise_without_stores_ms": t_nostore,
        "count_only_GBps": read_b / t_count / 1e6,
        "materialise_GBps": (read_b + write_b) / t_pairs / 1e6,
        "pairs_written": n, "bytes_per_match_written": 16,
    }))


if __name__ == "__main__":
    main()
