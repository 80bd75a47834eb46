"""The reference's SHIPPED configuration (core/Configuration.h: two-level partitioning off, 5 radix bits)
timed once on the GPU and on the host CPUs, for the record.
With one pass of 32 partitions a partition holds 1/32 of the relation: the reference's build/probe walks
chains of ~32 nodes, and here a partition's build side is joined in shared-memory slices, each of which
probes the partition's whole probe side -- work grows with (partition size / slice size), so this
configuration is timed at BASELINE config 1's size (16M x 16M), not at 256M x 256M.
  python tools/bench_shipped.py [log2n]   ->  one JSON line
"""
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")


def gpu_ms(cfg, dr, ds, n, reps=5):
    ctx = rhj.Context(cfg)
    for _ in range(2):
        r = ctx.join(dr, n, ds, n)
    assert r.matches == n, r.matches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        ctx.join_async(dr, n, ds, n)
    e1.record()
    torch.cuda.synchronize()
    t = ctx.timings().asdict()
    ctx.close()
    return e0.elapsed_time(e1) / reps, t


def main():
    log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    n = 1 << log2n
    dr = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    ds = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    rhj.fill_perm(dr, 16, n, 0, n, seed=1)
    rhj.fill_perm(ds, 16, n, 0, n, seed=2)
    out = {"workload": f"{n} x {n} 16-byte tuples, unique keys, 1:1 FK, one GPU / all host cores"}
    ms, t = gpu_ms(rhj.make_config(5, 5, 27, False), dr, ds, n)
    out["gpu_shipped_one_level_5bits"] = {"ms": ms, "tuples_per_s": 2 * n / ms * 1e3, "build_probe_us": t["bp_us"]}
    ms, t = gpu_ms(rhj.make_config(5, 5, 27, True), dr, ds, n)
    out["gpu_two_level_5+5bits"] = {"ms": ms, "tuples_per_s": 2 * n / ms * 1e3, "build_probe_us": t["bp_us"]}
    ms, t = gpu_ms(rhj.make_config(8, log2n - 20, 24, True), dr, ds, n)
    out[f"gpu_two_level_8+{log2n - 20}bits"] = {"ms": ms, "tuples_per_s": 2 * n / ms * 1e3, "build_probe_us": t["bp_us"]}
    try:
        import bench
        from oracle import oracle as O
        ranks = bench.host_ranks()
        per = n // ranks
        for name, cfg in (("cpu_shipped_one_level_5bits", O.Config(5, 5, 27, 0)), ("cpu_two_level_5+5bits", O.Config(5, 5, 27, 1))):
            info = O.run_ref(cfg, ranks, per, per, mode="join", timeout=1500)
            assert info["total_matches"] == per * ranks
            out[name] = {"ms": info["join_us_max"] / 1e3, "tuples_per_s": 2 * per * ranks / (info["join_us_max"] / 1e6),
                         "ranks": ranks, "cpu_model": bench.cpu_model()}
    except Exception as e:
        out["cpu"] = {"unavailable": str(e)[:300]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
