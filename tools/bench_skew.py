"""Stage times of one join on one GPU under skew, the dense uniform workload beside them: a
Zipf(1)-like probe side (P(k) ~ 1/k), one key holding 20 % of the probe side (what the rank owning
the hottest partition sees on 8 GPUs), and the latter sorted on the key."""
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")


def run(dr, ds, n, flags):
    cfg = rhj.make_config(8, 8, 24, True, flags=flags)
    ctx = rhj.Context(cfg)
    ctx.plan(n, n, 2 * n, 2 * n)
    for _ in range(3):
        r = ctx.join(dr, n, ds, n)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        ctx.join(dr, n, ds, n)
    e1.record()
    torch.cuda.synchronize()
    t = ctx.timings().asdict()
    ctx.close()
    return {"ms": round(e0.elapsed_time(e1) / 5, 3), "matches": r.matches,
            "stages_us": {k: round(v) for k, v in t.items() if v}}


def main():
    log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 28
    n = 1 << log2n
    dr = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    ds = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    rhj.fill_perm(dr, 16, n, 0, n, seed=1)
    out = {}
    rhj.fill_perm(ds, 16, n, 0, n, seed=2)
    out["uniform"] = run(dr, ds, n, 0)
    keys = ds.view(-1, 2)[:, 0]
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    step = 1 << 24
    for b in range(0, n, step):
        u = torch.rand(min(step, n - b), device="cuda", dtype=torch.float64, generator=g)
        keys[b:b + step] = torch.clamp(torch.floor(torch.exp(u * float(np.log(n)))).to(torch.int64) - 1, 0, n - 1)
    out["zipf"] = run(dr, ds, n, 0)
    rhj.fill_perm(ds, 16, n, 0, n, seed=2)
    keys[torch.rand(n, device="cuda", generator=g) < 0.2] = 12345
    out["hot20"] = run(dr, ds, n, 0)
    keys.copy_(torch.sort(keys).values)
    out["hot20_sorted"] = run(dr, ds, n, 0)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
