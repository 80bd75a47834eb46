"""ctypes/numpy front-end to the CPU oracle (oracle/hpcjoin_oracle.c).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package never
imports this module.

Also wraps oracle/_ref (the unmodified reference compiled by build_ref.sh) as a
subprocess, for pinning the restatement and for the timed CPU baseline.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_DIR = os.path.join(HERE, "_ref")

TUPLE = np.dtype([("key", "<u8"), ("rid", "<u8")])


class Config(C.Structure):
    """core/Configuration.h:28-37 knobs that change results."""

    _fields_ = [("net_bits", C.c_uint32), ("loc_bits", C.c_uint32),
                ("payload_bits", C.c_uint32), ("two_level", C.c_uint32), ("balanced", C.c_uint32)]

    def tag(self) -> str:
        return f"n{self.net_bits}_l{self.loc_bits}_p{self.payload_bits}_t{self.two_level}"


class Result(C.Structure):
    _fields_ = [("matches", C.c_uint64), ("sum_key", C.c_uint64), ("sum_rid_r", C.c_uint64),
                ("sum_rid_s", C.c_uint64), ("sum_mix", C.c_uint64)]

    def astuple(self):
        return (self.matches, self.sum_key, self.sum_rid_r, self.sum_rid_s, self.sum_mix)


def build(force: bool = False) -> str:
    """Compile liboracle.so (gcc) if missing or stale."""
    src = os.path.join(HERE, "hpcjoin_oracle.c")
    hdr = os.path.join(HERE, "hpcjoin_oracle.h")
    stale = (not os.path.exists(LIB_PATH)
             or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(src), os.path.getmtime(hdr)))
    if force or stale:
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-std=c11", "-shared", "-o", LIB_PATH, src, "-lm"])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
        cfgp = C.POINTER(Config)
        L.orc_rng_new.restype = vp
        L.orc_rng_new.argtypes = [C.c_uint]
        L.orc_rng_free.argtypes = [vp]
        L.orc_rng_next.restype = C.c_int32
        L.orc_rng_next.argtypes = [vp]
        L.orc_fill_unique.argtypes = [vp, vp, u64, u64, u64]
        L.orc_fill_modulo.argtypes = [vp, vp, u64, u64, u64, u64]
        L.orc_random_order.argtypes = [vp, vp, u64]
        L.orc_distribute.argtypes = [vp, vp, u64, u32]
        L.orc_distribute_perm.argtypes = [vp, u64, u32, u64]
        L.orc_distribute_perm.restype = C.c_int
        L.orc_perm.restype = u64
        L.orc_perm.argtypes = [u64, u64, u64]
        L.orc_fill_perm.argtypes = [vp, u64, u64, u64, u64, u64, u64, u64]
        L.orc_fill_zipf.argtypes = [vp, u64, u64, u64, C.c_double, u64, u64]
        L.orc_local_histogram.argtypes = [vp, u64, cfgp, vp]
        L.orc_global_histogram.argtypes = [vp, u32, u32, vp]
        L.orc_assignment.argtypes = [u32, u32, vp]
        L.orc_assignment_balanced.argtypes = [vp, vp, u32, u32, vp]
        L.orc_offsets.argtypes = [vp, vp, vp, u32, u32, u32, vp, vp, vp]
        L.orc_window_size.restype = u64
        L.orc_window_size.argtypes = [vp, vp, u32, u32]
        L.orc_compress.restype = u64
        L.orc_compress.argtypes = [u64, u64, cfgp]
        L.orc_network_partition.argtypes = [vp, u64, cfgp, vp, vp, vp]
        L.orc_local_part_histogram.argtypes = [vp, u64, cfgp, vp]
        L.orc_local_part_prefix.argtypes = [vp, u32, vp]
        L.orc_local_part_scatter.argtypes = [vp, u64, cfgp, vp, vp]
        L.orc_mix.restype = u64
        L.orc_mix.argtypes = [u64, u64, u64]
        L.orc_build_probe.argtypes = [vp, u64, vp, u64, cfgp, u32, C.POINTER(Result)]
        L.orc_join.restype = C.c_int
        L.orc_join.argtypes = [vp, vp, vp, vp, u32, cfgp, vp]
        L.orc_join_pairs.restype = C.c_int
        L.orc_join_pairs.argtypes = [vp, vp, vp, vp, u32, cfgp, vp, vp]
        L.orc_pack8.restype = u64
        L.orc_pack8.argtypes = [u64, u64, u32]
        L.orc_bruteforce_join.argtypes = [vp, u64, vp, u64, C.POINTER(Result)]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


# --------------------------------------------------------------------------
# generators
# --------------------------------------------------------------------------

class Rng:
    """One glibc rand() stream (main.cpp:94 seeds 1234 + rank)."""

    def __init__(self, seed: int):
        self.h = lib().orc_rng_new(seed)

    def next(self) -> int:
        return lib().orc_rng_next(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_rng_free(self.h)
            self.h = None


def generate_reference(nranks: int, inner_per_rank: int, outer_per_rank: int, gen: str = "unique",
                       seed: int = 1234):
    """main.cpp:70-104 for all ranks: returns (inner[], outer[]) lists of TUPLE arrays."""
    L = lib()
    gi, go = inner_per_rank * nranks, outer_per_rank * nranks
    rngs = [Rng(seed + r) for r in range(nranks)]
    inner = [np.zeros(inner_per_rank, TUPLE) for _ in range(nranks)]
    outer = [np.zeros(outer_per_rank, TUPLE) for _ in range(nranks)]
    for r in range(nranks):
        L.orc_fill_unique(rngs[r].h, _p(inner[r]), inner_per_rank, r * (gi // nranks), r * (gi // nranks))
        if gen == "unique":
            L.orc_fill_unique(rngs[r].h, _p(outer[r]), outer_per_rank, (nranks - r - 1) * (go // nranks),
                              r * (go // nranks))
        elif gen == "modulo":
            L.orc_fill_modulo(rngs[r].h, _p(outer[r]), outer_per_rank, (nranks - r - 1) * (gi // nranks),
                              r * (go // nranks), inner_per_rank)
        else:
            raise ValueError(gen)
    if nranks > 1:
        for rels, n in ((inner, inner_per_rank), (outer, outer_per_rank)):
            hs = (C.c_void_p * nranks)(*[g.h for g in rngs])
            ps = (C.c_void_p * nranks)(*[a.ctypes.data for a in rels])
            assert L.orc_distribute(hs, ps, n, nranks) == 0
    return inner, outer


def fill_perm(n: int, first: int, domain: int, seed: int, modulo: int = 0, start_key: int = 0,
              start_rid: int = 0) -> np.ndarray:
    t = np.zeros(n, TUPLE)
    lib().orc_fill_perm(_p(t), n, first, domain, seed, modulo, start_key, start_rid)
    return t


def distribute_perm(rels: list, seed: int) -> None:
    """In place: Relation::distribute for the harness generator (orc_distribute_perm)."""
    n = len(rels[0])
    assert all(len(r) == n for r in rels)
    ps = (C.c_void_p * len(rels))(*[a.ctypes.data for a in rels])
    assert lib().orc_distribute_perm(ps, n, len(rels), seed) == 0


def fill_zipf(n: int, first: int, domain: int, theta: float, seed: int, start_rid: int = 0) -> np.ndarray:
    t = np.zeros(n, TUPLE)
    lib().orc_fill_zipf(_p(t), n, first, domain, theta, seed, start_rid)
    return t


# --------------------------------------------------------------------------
# stages
# --------------------------------------------------------------------------

def local_histogram(t: np.ndarray, cfg: Config) -> np.ndarray:
    h = np.zeros(1 << cfg.net_bits, np.uint64)
    lib().orc_local_histogram(_p(t), len(t), C.byref(cfg), _p(h))
    return h


@dataclass
class Plan:
    """Everything tasks/HistogramComputation.cpp computes, for all ranks."""
    locals_: np.ndarray      # [nranks, P]
    global_: np.ndarray      # [P]
    assignment: np.ndarray   # [P] u32
    base: np.ndarray         # [P]
    rel: np.ndarray          # [nranks, P]
    abs_: np.ndarray         # [nranks, P]
    window_size: np.ndarray  # [nranks]


def balanced_assignment(global_inner: np.ndarray, global_outer: np.ndarray, nranks: int) -> np.ndarray:
    gi = np.ascontiguousarray(global_inner, np.uint64)
    go = np.ascontiguousarray(global_outer, np.uint64)
    assign = np.zeros(len(gi), np.uint32)
    lib().orc_assignment_balanced(_p(gi), _p(go), nranks, len(gi), _p(assign))
    return assign


def plan(locals_: np.ndarray, nranks: int, assignment: np.ndarray | None = None) -> Plan:
    """Histogram phase of one relation; `assignment` (e.g. from balanced_assignment) overrides p % n."""
    L = lib()
    locals_ = np.ascontiguousarray(locals_, np.uint64).reshape(nranks, -1)
    parts = locals_.shape[1]
    glob = np.zeros(parts, np.uint64)
    L.orc_global_histogram(_p(locals_), nranks, parts, _p(glob))
    assign = np.zeros(parts, np.uint32)
    if assignment is None:
        L.orc_assignment(nranks, parts, _p(assign))
    else:
        assign[:] = assignment
    base = np.zeros(parts, np.uint64)
    rel = np.zeros((nranks, parts), np.uint64)
    ab = np.zeros((nranks, parts), np.uint64)
    for r in range(nranks):
        L.orc_offsets(_p(locals_), _p(glob), _p(assign), nranks, r, parts, _p(base), _p(rel[r]), _p(ab[r]))
    ws = np.array([L.orc_window_size(_p(glob), _p(assign), parts, r) for r in range(nranks)], np.uint64)
    return Plan(locals_, glob, assign, base, rel, ab, ws)


def network_partition(rels: list, cfg: Config, pl: Plan) -> list:
    """Windows (one u64 array per rank) after every rank partitioned `rels[rank]`."""
    L = lib()
    nranks = len(rels)
    wins = [np.zeros(int(pl.window_size[r]), np.uint64) for r in range(nranks)]
    ptrs = (C.c_void_p * nranks)(*[w.ctypes.data for w in wins])
    for r in range(nranks):
        L.orc_network_partition(_p(rels[r]), len(rels[r]), C.byref(cfg), _p(pl.assignment), _p(pl.abs_[r]), ptrs)
    return wins


def local_partition(part: np.ndarray, cfg: Config):
    """tasks/LocalPartitioning.cpp for one partition: (hist, padded prefix, output)."""
    L = lib()
    sub = 1 << cfg.loc_bits
    part = np.ascontiguousarray(part, np.uint64)
    hist = np.zeros(sub, np.uint64)
    prefix = np.zeros(sub, np.uint64)
    L.orc_local_part_histogram(_p(part), len(part), C.byref(cfg), _p(hist))
    L.orc_local_part_prefix(_p(hist), sub, _p(prefix))
    out = np.full(len(part) + sub * 8, 0xFFFFFFFFFFFFFFFF, np.uint64)
    L.orc_local_part_scatter(_p(part), len(part), C.byref(cfg), _p(prefix), _p(out))
    return hist, prefix, out


def build_probe(r: np.ndarray, s: np.ndarray, cfg: Config, partition: int, acc: Result | None = None) -> Result:
    acc = acc if acc is not None else Result()
    r = np.ascontiguousarray(r, np.uint64)
    s = np.ascontiguousarray(s, np.uint64)
    lib().orc_build_probe(_p(r), len(r), _p(s), len(s), C.byref(cfg), partition, C.byref(acc))
    return acc


def join(inner: list, outer: list, cfg: Config) -> list:
    """operators/HashJoin.cpp:45-220 for all ranks; returns one Result per rank."""
    nranks = len(inner)
    ip = (C.c_void_p * nranks)(*[a.ctypes.data for a in inner])
    op = (C.c_void_p * nranks)(*[a.ctypes.data for a in outer])
    in_n = (C.c_uint64 * nranks)(*[len(a) for a in inner])
    on_n = (C.c_uint64 * nranks)(*[len(a) for a in outer])
    res = (Result * nranks)()
    rc = lib().orc_join(ip, in_n, op, on_n, nranks, C.byref(cfg), res)
    assert rc == 0
    return list(res)


class PairSink(C.Structure):
    _fields_ = [("pairs", C.c_void_p), ("cap", C.c_uint64), ("n", C.c_uint64)]


def join_pairs(inner: list, outer: list, cfg: Config, cap: int) -> tuple:
    """orc_join with materialised output: returns (results per rank, pairs per rank) where pairs is
    a u64 array [n, 2] of (rid_R, rid_S) in the reference's probe order (at most `cap` rows per rank)
    and the true number of matches is results[r].matches."""
    nranks = len(inner)
    ip = (C.c_void_p * nranks)(*[a.ctypes.data for a in inner])
    op = (C.c_void_p * nranks)(*[a.ctypes.data for a in outer])
    in_n = (C.c_uint64 * nranks)(*[len(a) for a in inner])
    on_n = (C.c_uint64 * nranks)(*[len(a) for a in outer])
    res = (Result * nranks)()
    bufs = [np.zeros((max(cap, 1), 2), np.uint64) for _ in range(nranks)]
    sinks = (PairSink * nranks)(*[PairSink(b.ctypes.data, cap, 0) for b in bufs])
    rc = lib().orc_join_pairs(ip, in_n, op, on_n, nranks, C.byref(cfg), res, sinks)
    assert rc == 0
    return list(res), [bufs[r][: min(sinks[r].n, cap)] for r in range(nranks)]


def join_hot_split(inner: list, outer: list, cfg: Config, inv_share: int = 4) -> tuple:
    """Per-rank results under hot-partition splitting, stated independently of the launcher: a partition p is
    hot when global_outer[p] * nranks * inv_share > sum(global_outer); the cold partitions are joined as
    orc_join does, and rank r additionally joins ALL ranks' inner tuples of the hot partitions with ITS OWN
    outer tuples of them.  Returns (results per rank, hot mask)."""
    n = len(inner)
    P = 1 << cfg.net_bits
    glob = np.zeros(P, np.uint64)
    for t in outer:
        glob += local_histogram(t, cfg)
    hot = (glob.astype(object) * (n * inv_share) > int(glob.sum())) if inv_share > 0 else np.zeros(P, bool)
    hot = np.array(hot, dtype=bool)
    pid = lambda t: (t["key"] & np.uint64(P - 1)).astype(np.int64)
    cold_i = [np.ascontiguousarray(t[~hot[pid(t)]]) for t in inner]
    cold_o = [np.ascontiguousarray(t[~hot[pid(t)]]) for t in outer]
    res = join(cold_i, cold_o, cfg)
    hot_r = np.ascontiguousarray(np.concatenate([t[hot[pid(t)]] for t in inner]))
    m = (1 << 64) - 1
    out = []
    for r in range(n):
        hot_s = np.ascontiguousarray(outer[r][hot[pid(outer[r])]])
        h = join([hot_r], [hot_s], cfg)[0]
        out.append(tuple((a + b) & m for a, b in zip(res[r].astuple(), h.astuple())))
    return out, hot


def total(results) -> tuple:
    m = (1 << 64) - 1
    acc = [0, 0, 0, 0, 0]
    for r in results:
        for i, v in enumerate(r.astuple()):
            acc[i] = (acc[i] + v) & m
    return tuple(acc)


def bruteforce(r: np.ndarray, s: np.ndarray) -> Result:
    out = Result()
    lib().orc_bruteforce_join(_p(r), len(r), _p(s), len(s), C.byref(out))
    return out


def mix(key: int, rid_r: int, rid_s: int) -> int:
    return lib().orc_mix(key, rid_r, rid_s)


# --------------------------------------------------------------------------
# oracle/_ref: the reference itself
# --------------------------------------------------------------------------

def ref_binary(cfg: Config) -> str | None:
    """Path of the prebuilt reference for these knobs; builds it when the reference tree is here."""
    path = os.path.join(REF_DIR, "hpcjoin_ref_" + cfg.tag())
    if not os.path.exists(path) and os.path.isdir(os.environ.get("REF", "/root/reference")):
        subprocess.check_call([os.path.join(HERE, "build_ref.sh"), str(cfg.net_bits), str(cfg.loc_bits),
                               str(cfg.payload_bits), str(cfg.two_level)], stdout=subprocess.DEVNULL)
    return path if os.path.exists(path) else None


def run_ref(cfg: Config, ranks: int, inner_per_rank: int, outer_per_rank: int, mode: str = "join",
            gen: str = "unique", dump_dir: str | None = None, seed: int = 1234, timeout: float = 600) -> dict:
    exe = ref_binary(cfg)
    if exe is None:
        raise FileNotFoundError("oracle/_ref binary for %s not built" % cfg.tag())
    cmd = [exe, "--ranks", str(ranks), "--inner", str(inner_per_rank), "--outer", str(outer_per_rank),
           "--mode", mode, "--gen", gen, "--seed", str(seed)]
    if dump_dir is not None:
        cmd += ["--dir", dump_dir]
        if gen != "file":
            cmd += ["--dump"]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True, timeout=timeout).stdout
    for line in out.splitlines():
        if line.startswith("REFJSON "):
            return json.loads(line[len("REFJSON "):])
    raise RuntimeError("no REFJSON line in reference output:\n" + out)
