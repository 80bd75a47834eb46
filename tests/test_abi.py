"""The C ABI without a GPU: the library builds for sm_100a, loads, exports every symbol
include/rhj.h declares, validates configurations, fails loudly without a device, and its
host-side planning logic equals the oracle."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "rhj.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rhj_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree(rhj):
    syms = header_symbols()
    assert len(syms) >= 35
    assert set(syms) == set(rhj.SYMBOLS), set(syms) ^ set(rhj.SYMBOLS)


def test_library_exports_every_declared_symbol(rhj):
    out = subprocess.run(["nm", "-D", "--defined-only", rhj.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (rhj_[a-z0-9_]+)", out))
    missing = set(header_symbols()) - exported
    assert not missing, missing


def test_library_contains_sm100a_code_only(rhj):
    out = subprocess.run(["cuobjdump", "--list-elf", rhj.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert not re.search(r"sm_(?!100a)\d+", out), out


def test_config_validation_messages(rhj):
    L = rhj.lib()
    h = C.c_void_p()
    bad = [rhj.make_config(0, 5, 27), rhj.make_config(13, 5, 27), rhj.make_config(5, 0, 27),
           rhj.make_config(5, 5, 60), rhj.make_config(5, 5, 27, tuple_bytes=12),
           rhj.make_config(5, 5, 27, build_capacity=3000), rhj.make_config(5, 5, 27, chunk_tuples=1000),
           rhj.make_config(5, 5, 27, allocation_factor=0.5)]
    for cfg in bad:
        assert L.rhj_init(C.byref(h), 0, 1, 0, C.byref(cfg)) == -1
        assert L.rhj_last_error()
    cfg = rhj.make_config()
    assert L.rhj_init(C.byref(h), 2, 2, 0, C.byref(cfg)) == -1  # rank out of range
    assert L.rhj_init(C.byref(h), 0, 65, 0, C.byref(cfg)) == -1


def test_no_silent_cpu_fallback(rhj):
    """Without a CUDA device rhj_init must fail with RHJ_ERR_CUDA, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rhj.RhjError) as e:
        rhj.Context(rhj.make_config())
    assert e.value.code == -2 and "no CPU path" in str(e.value)
    out = np.zeros(32, np.uint64)
    rc = rhj.lib().rhj_local_histogram_noalloc(out.ctypes.data, 16, 10, 5, out.ctypes.data, None)
    assert rc == -2


def test_workspace_bytes(rhj):
    cfg = rhj.make_config(8, 8, 24, True)
    n = 1 << 28
    b = rhj.lib().rhj_workspace_bytes(C.byref(cfg), 1, n, n, 0, 0)
    # two windows + two local outputs of ~1.1 n tuples, 8 bytes each, plus the per-tile offset
    # matrices (6 bytes per (tile, bin): ~100 MB per relation and pass) and small change
    assert 4 * 8 * n * 1.1 < b < 4 * 8 * n * 1.1 + (768 << 20)


@pytest.mark.parametrize("nranks,parts", [(1, 32), (2, 32), (3, 32), (4, 256), (8, 256), (5, 8), (16, 8)])
def test_host_offsets_equal_oracle(rhj, oracle, nranks, parts):
    """rhj_host_assignment / rhj_host_offsets == AssignmentMap + OffsetMap restated in the oracle."""
    rng = np.random.default_rng(nranks * 1000 + parts)
    loc = rng.integers(0, 10000, (nranks, parts)).astype(np.uint64)
    pl = oracle.plan(loc, nranks)
    L = rhj.lib()
    a = np.zeros(parts, np.uint32)
    L.rhj_host_assignment(nranks, parts, a.ctypes.data)
    assert np.array_equal(a, pl.assignment)
    for r in range(nranks):
        g, b, rel, ab = (np.zeros(parts, np.uint64) for _ in range(4))
        ws = np.zeros(nranks, np.uint64)
        L.rhj_host_offsets(loc.ctypes.data, nranks, r, parts, g.ctypes.data, b.ctypes.data, rel.ctypes.data,
                           ab.ctypes.data, ws.ctypes.data)
        assert np.array_equal(g, pl.global_) and np.array_equal(b, pl.base)
        assert np.array_equal(rel, pl.rel[r]) and np.array_equal(ab, pl.abs_[r])
        assert np.array_equal(ws, pl.window_size)


@pytest.mark.parametrize("nranks,parts", [(2, 32), (3, 32), (4, 256), (8, 512), (5, 8), (40, 64)])
def test_host_balanced_assignment_equals_oracle(rhj, oracle, nranks, parts):
    """RHJ_FLAG_BALANCED_ASSIGNMENT's rule on the host (rhj_host_assignment_balanced,
    rhj_host_offsets_assigned) against the oracle's statement of it, on skewed histograms with ties."""
    rng = np.random.default_rng(nranks * 7 + parts)
    loc_i = rng.integers(0, 50, (nranks, parts)).astype(np.uint64)          # many ties
    loc_o = (rng.zipf(1.3, (nranks, parts)) % 100000).astype(np.uint64)     # a few very heavy partitions
    gi, go = loc_i.sum(0, dtype=np.uint64), loc_o.sum(0, dtype=np.uint64)
    want = oracle.balanced_assignment(gi, go, nranks)
    L = rhj.lib()
    a = np.zeros(parts, np.uint32)
    L.rhj_host_assignment_balanced(gi.ctypes.data, go.ctypes.data, nranks, parts, a.ctypes.data)
    assert np.array_equal(a, want)
    # every rank keeps the number of partitions p % n gives it
    assert np.array_equal(np.bincount(a, minlength=nranks), np.bincount(np.arange(parts) % nranks, minlength=nranks))
    # and the heaviest rank is no heavier than under p % n
    w = (gi + go).astype(np.float64)
    assert np.bincount(a, w, nranks).max() <= np.bincount(np.arange(parts) % nranks, w, nranks).max()
    for loc in (loc_i, loc_o):
        pl = oracle.plan(loc, nranks, want)
        for r in range(nranks):
            g, b, rel, ab = (np.zeros(parts, np.uint64) for _ in range(4))
            ws = np.zeros(nranks, np.uint64)
            L.rhj_host_offsets_assigned(loc.ctypes.data, a.ctypes.data, nranks, r, parts, g.ctypes.data, b.ctypes.data,
                                        rel.ctypes.data, ab.ctypes.data, ws.ctypes.data)
            assert np.array_equal(g, pl.global_) and np.array_equal(b, pl.base)
            assert np.array_equal(rel, pl.rel[r]) and np.array_equal(ab, pl.abs_[r])
            assert np.array_equal(ws, pl.window_size)
