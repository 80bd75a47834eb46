"""rhj_local_histogram_noalloc: the histogram entry point for callers that cannot allocate.

Some users of this library cannot call cudaMalloc at all: an engine whose buffer manager
already owns every byte of the GPU (a library-side allocation would fail), or a caller on a
latency-critical path that must not risk the device-wide stalls an allocation can cause.
For them the library offers one entry point whose scratch lives in a fixed array inside the
library image.  This test calls it through the C ABI and checks it, bit for bit (integer
counts, so the tolerance is zero, as in the other GPU tests), against the existing
rhj_histogram path -- twice in a row, with different inputs, because the array is shared by
all calls and keeps the previous call's contents.

The file name sorts last on purpose: a run that stops at the first failure has by then run
every other test.  It must also pass alone in a fresh process (the first call does the
per-device set-up).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _reference_path_histogram(rhj, cfg, dev_tuples, n):
    """u64[P] through the existing entry points (pool-backed scratch)."""
    ctx = rhj.Context(cfg)
    ctx.plan(max(n, 1), 1)
    import torch
    empty = torch.empty(0, dtype=torch.int64, device="cuda")
    ctx.histogram(dev_tuples, n, empty, 0)
    out = ctx.get_local_histogram(0)
    ctx.close()
    return out


def _noalloc_histogram(rhj, dev_tuples, tuple_bytes, n, bits):
    import torch
    out = torch.full((1 << bits,), -1, dtype=torch.int64, device="cuda")  # caller-owned output
    rc = rhj.lib().rhj_local_histogram_noalloc(dev_tuples.data_ptr() if n else 0, tuple_bytes, n, bits,
                                               out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rhj.lib().rhj_last_error().decode()
    torch.cuda.synchronize()
    return out.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("tuple_bytes", [16, 8])
def test_noalloc_histogram_matches_pool_backed_path_twice(rhj, oracle, tuple_bytes):
    import torch
    bits = 8
    cfg = rhj.make_config(bits, 8, 24, True, tuple_bytes=tuple_bytes)
    # two different relations back to back: the second is smaller and differently distributed, so
    # counts left in the shared array by the first call would show up as a mismatch
    first = oracle.fill_perm(300001, 0, 300001, seed=5)
    second = oracle.fill_zipf(70001, 0, 50000, 1.0, seed=6)
    for rel in (first, second):
        n = len(rel)
        host = rel if tuple_bytes == 16 else ((rel["key"] << np.uint64(32)) | rel["rid"]).astype(np.uint64)
        dev = torch.from_numpy(np.ascontiguousarray(host).view(np.int64).copy()).cuda()
        got = _noalloc_histogram(rhj, dev, tuple_bytes, n, bits)
        want = _reference_path_histogram(rhj, cfg, dev, n)
        assert np.array_equal(got, want)
        assert int(got.sum()) == n
        # and against the CPU oracle, for good measure
        assert np.array_equal(got, oracle.local_histogram(rel, oracle.Config(bits, 8, 24, 1)))


def test_noalloc_histogram_empty_input_and_limit(rhj):
    import torch
    got = _noalloc_histogram(rhj, torch.empty(0, dtype=torch.int64, device="cuda"), 16, 0, 5)
    assert got.shape == (32,) and not got.any()
    # one chunk too many for the built-in array at 12 radix bits: refused, nothing launched
    out = torch.zeros(1 << 12, dtype=torch.int64, device="cuda")
    n = (1 << 24) + 1
    rc = rhj.lib().rhj_local_histogram_noalloc(out.data_ptr(), 16, n, 12, out.data_ptr(), None)
    assert rc == -4
    assert b"scratch cells" in rhj.lib().rhj_last_error()
