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
