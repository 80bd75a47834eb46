"""BASELINE.json config 2 at full size through size-independent properties (the oracle
would need minutes): count == |S|, sums == G(G-1)/2, histograms uniform."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("log2n", [24, 28])
def test_unique_fk_closed_form(rhj, log2n):
    import torch
    free, _ = torch.cuda.mem_get_info()
    n = 1 << log2n
    if free < n * 16 * 2 * 2.5:
        pytest.skip("not enough device memory")
    cfg = rhj.make_config(8, 8, 24, True)
    di = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    do = torch.empty(2 * n, dtype=torch.int64, device="cuda")
    rhj.fill_perm(di, 16, n, 0, n, seed=1)
    rhj.fill_perm(do, 16, n, 0, n, seed=2)
    ctx = rhj.Context(cfg)
    r = ctx.join(di, n, do, n)
    g = n * (n - 1) // 2 % (1 << 64)
    assert r.matches == n
    assert r.sum_key == g and r.sum_rid_s == g and r.sum_rid_r == g
    for which in (0, 1):
        assert np.all(ctx.get_global_histogram(which) == n // 256)
        h, pf = ctx.get_local_partition(which, 3)
        assert np.all(h == n // 65536)
    # idempotence: the same context joins again with identical result
    r2 = ctx.join(di, n, do, n)
    assert r2.astuple() == r.astuple()
    ctx.close()


def test_asymmetric_packed_1_to_16(rhj):
    """BASELINE config 5 shape on one GPU: packed 8-byte tuples, 16 M build x 256 M probe,
    every build key matched 16 times (fillModuloValues pattern, data/Relation.cpp:75-85)."""
    import torch
    nr, ns = 1 << 24, 1 << 28
    free, _ = torch.cuda.mem_get_info()
    if free < (nr + ns) * 8 * 4:
        pytest.skip("not enough device memory")
    cfg = rhj.make_config(8, 8, 24, True, tuple_bytes=8)
    dr = torch.empty(nr, dtype=torch.int64, device="cuda")
    ds = torch.empty(ns, dtype=torch.int64, device="cuda")
    rhj.fill_perm(dr, 8, nr, 0, nr, seed=3)
    rhj.fill_perm(ds, 8, ns, 0, ns, seed=4, modulo=nr)
    ctx = rhj.Context(cfg)
    r = ctx.join(dr, nr, ds, ns)
    assert r.matches == ns
    assert r.sum_rid_s == ns * (ns - 1) // 2 % (1 << 64)
    assert r.sum_key == 16 * (nr * (nr - 1) // 2) % (1 << 64)
    assert r.sum_rid_r == 16 * (nr * (nr - 1) // 2) % (1 << 64)
    ctx.close()


def test_zipf_probe_side_is_not_pathological(rhj, oracle):
    """BASELINE config 4 shape on one GPU: Zipf(1.0) probe keys.  The hot sub-partition is split
    into probe slices (work items), so the join must stay within a small factor of the uniform case."""
    import time
    import torch
    nr, ns = 1 << 22, 1 << 24
    inner = oracle.fill_perm(nr, 0, nr, seed=5)
    outer = oracle.fill_zipf(ns, 0, nr, 1.0, seed=6)
    di = torch.from_numpy(inner.view(np.int64).copy()).cuda()
    do = torch.from_numpy(outer.view(np.int64).copy()).cuda()
    cfg = rhj.make_config(6, 6, 26, True)
    ctx = rhj.Context(cfg)
    ctx.join(di, nr, do, ns)  # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = ctx.join(di, nr, do, ns)
    dt = time.perf_counter() - t0
    assert r.matches == ns  # every probe key exists exactly once in the build side
    assert r.sum_rid_s == ns * (ns - 1) // 2
    assert dt < 0.05, f"skewed join took {dt * 1e3:.1f} ms"
    ctx.close()
