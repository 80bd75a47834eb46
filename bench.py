#!/usr/bin/env python
"""bench.py -- join tuples/s of the radix hash join hot path on B200.

  python bench.py --gpus N --steps K --warmup W            (this repo's CUDA path)
  python bench.py --impl reference --gpus N --steps K ...  (the reference's own CPU code, oracle/_ref)

A "step" is one pass of the whole hot path (histogram -> network partitioning ->
local partitioning -> build/probe) over one pair of synthetic relations.
N=1 workload = BASELINE.json configs[1]: 256M x 256M 16-byte tuples, unique keys, 1:1 FK.
N>1 (torchrun, one rank per GPU) = configs[2]: the same per GPU (weak scaling), keys
spread over all ranks so (N-1)/N of the tuples cross NVLink in the network pass.

One JSON line on stdout (rank 0).  `value` is device-timed with inputs resident in HBM;
`e2e` times the same metric through the host-buffer entry point with the H2D copies inside.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PACKAGE = "distributed-radxi-hash-join-on-gpus_b200"

METRIC = "join tuples/sec (device-timed, max over ranks)"
UNIT = "tuples/s"
# algorithmic HBM bytes per input tuple, 16-byte tuples
BYTES_PER_TUPLE = {"hist": 16, "net": 24, "local": 24, "bp": 8, "total": 72}
KERNELS_PER_JOIN = None  # measured through rhj_launch_count()


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2n", type=int, default=28, help="tuples per relation per GPU = 2^log2n (default 256M)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sparse", action="store_true", help="skip the sparse-key run (stages.bp_sparse)")
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--net-bits", type=int, default=0, help="network radix bits (0 = by GPU count)")
    ap.add_argument("--loc-bits", type=int, default=0, help="local radix bits (0 = by GPU count)")
    ap.add_argument("--agg", action="store_true", help="warp-aggregated histogram atomics")
    ap.add_argument("--build-cap", type=int, default=0, help="build slice capacity (0 = library default)")
    ap.add_argument("--overlap", type=int, default=-1, help="SMs of the outer network pass beside the inner local pass (-1 = by GPU count, 0 = off)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "staged", "direct"])
    ap.add_argument("--debug-env", action="store_true", help="let the library read its RHJ_* ablation switches")
    return ap.parse_args()


# ------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed regions."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device, self.proc, self.lines = device, None, []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], 0, 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax = max(smax, float(f[1]))
                power = max(power, float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        # samples under load: the upper half of the observed SM clocks (idle samples sit at the floor)
        sm_sorted = sorted(sm)
        busy = sm_sorted[len(sm_sorted) // 2:] if sm_sorted else []
        med = busy[len(busy) // 2] if busy else None
        return {"sm_mhz": med, "sm_max_mhz": smax or None, "power_w_max": power or None, "samples": len(sm),
                "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------
# CPU baseline: the reference's own code (oracle/_ref), bounded sample
# ------------------------------------------------------------------------------------
def cpu_reference_run(per_rank: int, ranks: int):
    """One run of the unmodified reference (two-level on, shipped 5+5 radix bits = BASELINE
    configs[0] knobs) as `ranks` single-threaded ranks; returns (tuples/s, description)."""
    from oracle import oracle as O
    cfg = O.Config(5, 5, 27, 1)
    info = O.run_ref(cfg, ranks, per_rank, per_rank, mode="join", timeout=900)
    total = 2 * per_rank * ranks
    assert info["total_matches"] == per_rank * ranks, info
    secs = info["join_us_max"] / 1e6
    return total / secs, info


def host_ranks() -> int:
    n = os.cpu_count() or 1
    r = 1
    while r * 2 <= min(n, 32):
        r *= 2
    return r


def cpu_model() -> str:
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def mem_available_bytes() -> int:
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable"):
                return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return 0


def reference_sample(total_per_relation: int):
    """(ranks, tuples per rank) for the reference's CPU run of `total_per_relation` x the same:
    all host cores the reference can use (one single-threaded rank per core, a power of two so
    that p % n balances the 32 shipped partitions).  The whole relation pair is used unless the
    host lacks the memory for it (~80 bytes per tuple pair of a rank: relations, pool, windows)."""
    ranks = host_ranks()
    per = total_per_relation // ranks
    avail = mem_available_bytes()
    note = None
    while avail and per * ranks * 96 > avail * 0.8 and per > (1 << 20):
        per //= 2
        note = f"host has {avail >> 30} GiB available: sample reduced to {per * ranks} tuples per relation"
    return ranks, per, note


def cpu_baseline_block(total_per_relation: int) -> dict:
    """The reference's own CPU code on this box: (1) the bench workload on all host cores,
    (2) BASELINE configs[0] as stated: 16M x 16M, 2 ranks, two-level partitioning on."""
    ranks, per_rank, note = reference_sample(total_per_relation)
    try:
        v, info = cpu_reference_run(per_rank, ranks)
        block = {"value": v, "unit": UNIT, "cores": ranks, "kind": "reference",
                 "sample": f"{ranks} ranks x ({per_rank} x {per_rank}) 16-byte tuples = {ranks * per_rank} x {ranks * per_rank}, "
                           f"unique keys, reference generator + distribute, two-level on, 5+5 radix bits; "
                           f"JTOTAL max over ranks {info['join_us_max']} us" + (f"; {note}" if note else ""),
                 "host_cpus": os.cpu_count(), "cpu_model": cpu_model()}
        try:
            v1, info1 = cpu_reference_run(1 << 23, 2)
            block["config1"] = {"value": v1, "unit": UNIT, "cores": 2,
                                "sample": "BASELINE configs[0]: 16M x 16M 16-byte tuples, 2 ranks x (8M x 8M), unique keys, "
                                          f"two-level on, 5+5 radix bits; JTOTAL max over ranks {info1['join_us_max']} us"}
        except Exception as e:
            block["config1"] = {"unavailable": str(e)[:200]}
        return block
    except Exception as e:  # the reference binary did not travel / failed: fall back to the oracle port
        from oracle import oracle as O
        n = 1 << 21
        inner, outer = [O.fill_perm(n, 0, n, 1)], [O.fill_perm(n, 0, n, 2)]
        t0 = time.time()
        res = O.join(inner, outer, O.Config(5, 5, 27, 1))
        dt = time.time() - t0
        assert res[0].matches == n
        return {"value": 2 * n / dt, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": f"oracle port, 1 rank, {n} x {n} tuples (oracle/_ref unavailable: {e})"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # the GPU arm's N=1 workload (256M x 256M) in full; at N>1 the same 256M x 256M as a bounded
    # sample of the N x larger global relations, so that K+W runs still end within minutes
    n_total = 1 << args.log2n
    ranks, per_rank, note = reference_sample(n_total)
    vals = []
    info = None
    for i in range(args.warmup + args.steps):
        v, info = cpu_reference_run(per_rank, ranks)
        if i >= args.warmup:
            vals.append(v)
    value = len(vals) / sum(1.0 / v for v in vals)  # harmonic mean == total tuples / total time
    total = 2 * per_rank * ranks
    sample = (f"{ranks} reference ranks (1 thread each) x ({per_rank} x {per_rank}) 16-byte tuples per rank = "
              f"{ranks * per_rank} x {ranks * per_rank}, unique keys, reference generator + distribute, two-level on, "
              f"5+5 radix bits (BASELINE configs[0] knobs)" + (f"; {note}" if note else ""))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total / value * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": f"{ranks * per_rank} x {ranks * per_rank} 16-byte tuples, unique keys, 1:1 FK "
                                   f"(BASELINE configs[{1 if args.gpus == 1 else 2}] size{'' if not note else ', reduced'}): " + sample,
                       "same_size_as_gpu_arm": ranks * per_rank == (1 << args.log2n) * max(1, args.gpus)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": ranks, "kind": "reference", "sample": sample,
                             "host_cpus": os.cpu_count(), "cpu_model": cpu_model()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def parity_prologue(rhj, torch, dist, rank, world, local_rank, net_bits, loc_bits, per=150000):
    """Checker use of the oracle outside every timed region: all five result words of THIS rank
    must equal oracle.join(...)[rank] on the reference generator's relations."""
    import numpy as np
    from oracle import oracle as O
    inner, outer = O.generate_reference(world, per, per)
    ocfg = O.Config(net_bits, loc_bits, 32 - net_bits, 1)
    want = O.join(inner, outer, ocfg)[rank].astuple()
    cfg = rhj.make_config(net_bits, loc_bits, 32 - net_bits, True)
    ri = rhj.Relation.from_numpy(inner[rank], per * world, device=local_rank)
    ro = rhj.Relation.from_numpy(outer[rank], per * world, device=local_rank)
    hj = rhj.HashJoin(world, rank, ri, ro, config=cfg, device=local_rank, window_caps=(per * world, per * world))
    got = hj.join().astuple()
    hj.close()
    ok = torch.tensor([1 if got == want else 0], dtype=torch.int32, device=f"cuda:{local_rank}")
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if not int(ok.item()):
        raise AssertionError(f"rank {rank}: CUDA {got} != oracle {want}")
    return {"vs": "oracle", "ranks": world, "ok": True, "tuples_per_rank": per,
            "workload": "reference generator + distribute (main.cpp:94-104)", "words": 5}


# ------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    rhj = importlib.import_module(PACKAGE)
    rhj.build()
    L = rhj.lib()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n = 1 << args.log2n          # tuples per relation per GPU
    G = n * world                # global relation size
    flags = (rhj.FLAG_HIST_WARP_AGGREGATE if args.agg else 0) | (0 if args.no_e2e else rhj.FLAG_STAGE_HOST_INPUTS)
    flags |= rhj.FLAG_DEBUG_ENV if args.debug_env else 0
    flags |= {"auto": 0, "staged": rhj.FLAG_STAGED_EXCHANGE, "direct": rhj.FLAG_DIRECT_EXCHANGE}[args.exchange]
    # radix bits: sub-partitions of ~4096 build tuples (one shared-memory table) need
    # log2(global size / 4096) bits in total; weak scaling adds log2(N) bits
    default_bits = {1: (8, 8), 2: (8, 9), 4: (9, 9), 8: (9, 10)}.get(world, (9, 10))
    net_bits = args.net_bits or default_bits[0]
    loc_bits = args.loc_bits or default_bits[1]
    cfg = rhj.make_config(net_bits, loc_bits, 32 - net_bits, True, tuple_bytes=16, build_capacity=args.build_cap,
                          chunk_tuples=args.chunk, flags=flags)

    # synthetic relations, generated on the device (harness generator, bit-identical to the oracle's)
    d_inner = torch.empty(2 * n, dtype=torch.int64, device=dev)
    d_outer = torch.empty(2 * n, dtype=torch.int64, device=dev)
    rhj.fill_perm(d_inner, 16, n, rank * n, G, seed=1, start_rid=rank * n)
    rhj.fill_perm(d_outer, 16, n, (world - 1 - rank) * n, G, seed=2, start_rid=rank * n)
    torch.cuda.synchronize()

    ctx = rhj.Context(cfg, rank, world, local_rank)
    cap = int(n * 1.1) + 4096
    ctx.plan(n, n, cap, cap)
    inner = rhj.Relation(d_inner, n, G)
    outer = rhj.Relation(d_outer, n, G)
    rhj.hashjoin.exchange_windows(ctx)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        if world == 1:
            ctx.join_async(d_inner, n, d_outer, n)
        else:
            rhj.join_multi_rank(ctx, inner, outer, overlap=None if args.overlap < 0 else args.overlap)

    from tools import checksum as CK
    m64 = CK.M64
    want = CK.expected_result(torch, dist, d_inner, n, d_outer, n, G, world, rank, dev)
    closed = (G * (G - 1) // 2) & m64
    assert want[0] == G and want[1] == closed and want[2] == closed and want[3] == closed, want

    def check(res):
        got = CK.reduce_result(torch, dist, res, world, dev)
        assert got == want, f"result {got} != expected {want} (count, sum key, sum rid_R, sum rid_S, sum mix)"

    # untimed prologue: the reference's own workload (main.cpp:94-104 generator + distribute) at
    # oracle size through the same launcher, every rank against the oracle's per-rank result
    parity = parity_prologue(rhj, torch, dist, rank, world, local_rank, net_bits, loc_bits)

    for _ in range(args.warmup):
        step()
    check(ctx.result())

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.2)
    launches0 = L.rhj_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_us = {"jhist_us": 0.0, "jmpi_us": 0.0, "lp_us": 0.0, "bp_us": 0.0, "jtotal_us": 0.0}
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = L.rhj_launch_count() - launches0
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    res = ctx.result()
    check(res)
    tm = ctx.timings().asdict()  # stage split of the last step (cudaEvents inside the library)
    for k in stage_us:
        stage_us[k] = tm[k]

    value = 2.0 * G * args.steps / (ms_total / 1e3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md 6.65 TB/s)"

    # roofline of the dominant kernel: k_scatter of the network pass (2 launches per join: inner, outer)
    net_launch_us = stage_us["jmpi_us"] / 2.0
    net_bytes = BYTES_PER_TUPLE["net"] * n
    achieved = net_bytes / (net_launch_us * 1e-6) / 1e9 if net_launch_us > 0 else 0.0
    # DRAM bytes per launch from an ncu capture of exactly this configuration, else null
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic_r02.json")))
        traffic = tr.get(f"n{world}_bits{net_bits}+{loc_bits}_log2n{args.log2n}", {}).get("k_scatter_wc_network_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_scatter_wc<InputSource<16>> (network partitioning pass)",
                "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": net_bytes, "launch_us": net_launch_us}
    step_us = ms_total * 1e3 / args.steps
    stages = {}
    for key, us_key in (("hist", "jhist_us"), ("net", "jmpi_us"), ("local", "lp_us"), ("bp", "bp_us")):
        us = stage_us[us_key]
        gbs = BYTES_PER_TUPLE[key] * 2 * n / (us * 1e-6) / 1e9 if us > 0 else 0.0
        stages[key] = {"us": us, "GBps": gbs, "frac_of_hbm_peak": gbs / hbm_peak}
    whole = BYTES_PER_TUPLE["total"] * 2 * n / (step_us * 1e-6) / 1e9
    stages["join"] = {"us": step_us, "GBps": whole, "frac_of_hbm_peak": whole / hbm_peak}
    # SNETCOMPL: waiting for the other ranks' stores (barrier), part of `local` in the reference's keys too
    stages["exchange_wait"] = {"us": tm["snetcompl_us"]}
    roofline_nvlink = None
    if world > 1:
        # the exchange is fused into the network pass: every tuple bound for another rank crosses
        # NVLink as part of a full 128-byte line stored by k_scatter_wc
        nv_bytes = 2 * 8 * n * (world - 1) / world  # both relations, per GPU and direction
        # staged exchange: the window in which this rank's pushes run; direct: the network pass itself
        nv_us = tm["push_us"] if ctx.is_staged() and tm["push_us"] > 0 else stage_us["jmpi_us"]
        nv_gbs = nv_bytes / (nv_us * 1e-6) / 1e9 if nv_us > 0 else 0.0
        roofline_nvlink = {"bound": "nvlink", "bytes_per_gpu_per_direction": nv_bytes, "us": nv_us,
                           "achieved": nv_gbs, "peak": 900.0, "unit": "GB/s", "frac": nv_gbs / 900.0,
                           "frac_of_measured_peer_copy_770": nv_gbs / 770.0,
                           "how": "8 B x tuples x (N-1)/N per relation / device time (cudaEvents) of the network pass (direct exchange) "
                                  "or of first push start to last push end (staged exchange), rank 0"}

    # ---- end to end: host buffers in, result out -------------------------------------------
    e2e = None
    if not args.no_e2e:
        nbytes = n * 16
        pin_i, pin_o = rhj.PinnedBuffer(nbytes), rhj.PinnedBuffer(nbytes)
        rhj.check(L.rhj_memcpy_d2h(pin_i.ptr, d_inner.data_ptr(), nbytes))
        rhj.check(L.rhj_memcpy_d2h(pin_o.ptr, d_outer.data_ptr(), nbytes))

        def e2e_step():
            if world == 1:
                return ctx.join_host(pin_i.ptr, n, pin_o.ptr, n)
            rhj.join_multi_rank(ctx, inner, outer, host=(pin_i.ptr, pin_o.ptr),
                                overlap=None if args.overlap < 0 else args.overlap)
            return ctx.result()
        r = e2e_step()  # warm-up
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            r = e2e_step()
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        check(r)
        # what the host can deliver at all: the same pinned buffers copied to the device and nothing else,
        # all ranks at once (the e2e step cannot be faster than this)
        barrier()
        t0 = time.perf_counter()
        for _ in range(2):
            rhj.check(L.rhj_memcpy_h2d(d_inner.data_ptr(), pin_i.ptr, nbytes))
            rhj.check(L.rhj_memcpy_h2d(d_outer.data_ptr(), pin_o.ptr, nbytes))
        barrier()
        dc = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dc, op=dist.ReduceOp.MAX)
        h2d_only_ms = float(dc.item()) * 1e3 / 2
        e2e = {"value": 2.0 * G * args.e2e_steps / float(dt.item()), "unit": UNIT,
               "h2d_bytes_per_step": 2 * nbytes * world, "d2h_bytes_per_step": 40 * world,
               "steps": args.e2e_steps, "ms_per_step": float(dt.item()) * 1e3 / args.e2e_steps,
               "h2d_only_ms_per_step": h2d_only_ms,
               "h2d_only_GBps_per_gpu": 2 * nbytes / (h2d_only_ms * 1e-3) / 1e9,
               "step_over_h2d_only": float(dt.item()) * 1e3 / args.e2e_steps / h2d_only_ms,
               "api": "rhj_join_host (pinned host buffers)" if world == 1 else
                      "rhj_stage_host_inputs (pinned host buffers) + the stage entry points, one process per GPU"}
    clocks = sampler.stop() if rank == 0 else None

    # ---- the same join on sparse unique keys (one GPU): sub-partition sizes scatter around their mean and
    # hashed keys chain in the build table, which the dense bench workload never shows
    if world == 1 and not args.no_sparse:
        rhj.fill_perm(d_inner, 16, n, 0, 8 * n, seed=7)
        rhj.fill_perm(d_outer, 16, n, 0, 8 * n, seed=7)
        perm = torch.randperm(n, device=dev)
        d_outer.view(-1, 2)[:, 0] = d_outer.view(-1, 2)[perm, 0]
        del perm
        for _ in range(3):
            ctx.join_async(d_inner, n, d_outer, n)
        rs = ctx.result()
        assert rs.matches == n, rs.matches
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        s0.record()
        for _ in range(5):
            ctx.join_async(d_inner, n, d_outer, n)
        s1.record()
        torch.cuda.synchronize()
        ts = ctx.timings().asdict()
        stages["bp_sparse"] = {"us": ts["bp_us"], "join_ms": s0.elapsed_time(s1) / 5,
                               "workload": f"{n} x {n} unique keys drawn from [0, {8 * n}), probe side permuted"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_block(n)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": {"workload": f"{n} x {n} 16-byte tuples per GPU (global {G} x {G}), unique keys, 1:1 FK, "
                                       f"counter-based generator; BASELINE configs[{1 if world == 1 else 2}]",
                           "radix_bits": f"{net_bits} network + {loc_bits} local", "tuple_bytes": 16,
                           "l2": "inputs (8 GiB per GPU) are larger than L2; no flush needed",
                           "exchange": "none (1 GPU)" if world == 1 else
                           ("send buffers pushed to the owners' windows as sequential runs over NVLink (CUDA IPC), group of "
                            "chunks by group beside the scatter; arrival flags instead of a barrier" if ctx.is_staged() else
                            "fused: the scatter kernel stores full 128-byte lines into the owners' windows over NVLink "
                            "(CUDA IPC); arrival flags instead of a barrier")},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "roofline_nvlink": roofline_nvlink,
                "stages": stages, "cpu_baseline": cpu,
                "result": {"matches": G, "checksum_ok": True, "checksum_words": 5}, "parity": parity}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
