"""Join time against input size on one GPU (dense unique keys, 1:1), radix bits chosen so that
sub-partitions hold ~4 K tuples; shows the fixed cost of a join next to its streaming rate."""
import importlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")


def main():
    out = []
    for log2n in (16, 18, 20, 22, 24, 26, 28):
        n = 1 << log2n
        bits = max(2, log2n - 12)
        net = min(8, (bits + 1) // 2)
        loc = bits - net
        cfg = rhj.make_config(net, loc, 32 - net, True)
        dr = torch.empty(2 * n, dtype=torch.int64, device="cuda")
        ds = torch.empty(2 * n, dtype=torch.int64, device="cuda")
        rhj.fill_perm(dr, 16, n, 0, n, seed=1)
        rhj.fill_perm(ds, 16, n, 0, n, seed=2)
        ctx = rhj.Context(cfg)
        for _ in range(3):
            r = ctx.join(dr, n, ds, n)
        assert r.matches == n
        reps = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            ctx.join(dr, n, ds, n)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        t = ctx.timings().asdict()
        out.append({"log2n": log2n, "bits": [net, loc], "ms": round(ms, 4), "Gtuples_per_s": round(2 * n / ms / 1e6, 2),
                    "stages_us": {k: round(v, 1) for k, v in t.items() if v}})
        ctx.close()
        del dr, ds
    print(json.dumps(out))


if __name__ == "__main__":
    main()
