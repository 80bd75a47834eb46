"""Join time with sparse random unique keys (sub-partition sizes scatter around their mean instead
of being exactly equal as with the dense generator), for each build capacity given."""
import importlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")


def main():
    log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 28
    caps = [int(x) for x in sys.argv[2:]] or [4096, 5120]
    n = 1 << log2n
    # unique sparse keys: a permutation of [0, 8n), first n values of it for R, S probes R's keys again
    dr = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    ds = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    rhj.fill_perm(dr, 16, n, 0, 8 * n, seed=7)
    rhj.fill_perm(ds, 16, n, 0, 8 * n, seed=7)
    perm = torch.randperm(n, device="cuda")
    ds.view(-1, 2)[:, 0] = ds.view(-1, 2)[perm, 0]
    del perm
    out = {}
    for cap in caps:
        cfg = rhj.make_config(8, log2n - 20, 24, True, build_capacity=cap)
        ctx = rhj.Context(cfg)
        for _ in range(3):
            r = ctx.join(dr, n, ds, n)
        assert r.matches == n, r.matches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            ctx.join(dr, n, ds, n)
        e1.record()
        torch.cuda.synchronize()
        t = ctx.timings()
        out[str(cap)] = {"ms": e0.elapsed_time(e1) / 10, "build_probe_us": t.bp_us, "local_pass_us": t.lp_us}
        ctx.close()
    print(json.dumps({"workload": f"{n} x {n} sparse unique keys, 8+{log2n - 20} radix bits", "by_build_capacity": out}))


if __name__ == "__main__":
    main()
