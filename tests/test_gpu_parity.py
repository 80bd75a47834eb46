"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs.  Integer work only -> every comparison is bit-exact (no tolerance)."""
import numpy as np
import pytest

from _util import GpuRun, check_against_oracle, ocfg

pytestmark = pytest.mark.gpu


def cfgs(rhj):
    return {
        "shipped_one_level": rhj.make_config(5, 5, 27, False),            # core/Configuration.h as shipped
        "shipped_two_level": rhj.make_config(5, 5, 27, True),             # BASELINE config 1 knobs
        "gpu_8_8": rhj.make_config(8, 8, 24, True),                        # the B200 fan-out
        "tiny_2_3": rhj.make_config(2, 3, 30, True, chunk_tuples=8192),
        "wide_10_6": rhj.make_config(10, 6, 22, True, build_capacity=1024),
    }


@pytest.mark.parametrize("name", ["shipped_one_level", "shipped_two_level", "gpu_8_8", "tiny_2_3", "wide_10_6"])
@pytest.mark.parametrize("nranks", [1, 2, 4])
def test_reference_generator_all_stages(rhj, oracle, name, nranks):
    """main.cpp:94-104 workload (unique keys, 1:1 FK, distribute) at oracle-friendly size."""
    cfg = cfgs(rhj)[name]
    per = 60000 if name != "shipped_one_level" else 20000
    inner, outer = oracle.generate_reference(nranks, per, per)
    res = check_against_oracle(rhj, oracle, cfg, inner, outer)
    assert sum(r.matches for r in res) == nranks * per  # |S| matches


@pytest.mark.parametrize("nranks", [1, 3])
def test_modulo_outer_duplicates(rhj, oracle, nranks):
    """fillModuloValues (Relation.cpp:75-85): every inner key matched several times."""
    cfg = rhj.make_config(4, 4, 28, True)
    inner, outer = oracle.generate_reference(nranks, 5000, 40000, gen="modulo")
    check_against_oracle(rhj, oracle, cfg, inner, outer)


def test_duplicate_build_keys_and_misses(rhj, oracle):
    """Non-unique build keys (chains longer than 1) and probe keys without a partner."""
    rng = np.random.default_rng(7)
    inner = np.zeros(30000, oracle.TUPLE)
    inner["key"] = rng.integers(0, 5000, len(inner))
    inner["rid"] = np.arange(len(inner))
    outer = np.zeros(50000, oracle.TUPLE)
    outer["key"] = rng.integers(0, 10000, len(outer))
    outer["rid"] = np.arange(len(outer)) + 1000000
    for cfg in (rhj.make_config(3, 3, 29, True), rhj.make_config(3, 3, 29, False)):
        res = check_against_oracle(rhj, oracle, cfg, [inner], [outer])
        assert res[0].astuple() == oracle.bruteforce(inner, outer).astuple()


def test_heavy_skew_single_key(rhj, oracle):
    """All probe tuples hit one key: one sub-partition holds everything (slices of S, hot bin)."""
    inner = oracle.fill_perm(20000, 0, 20000, seed=3)
    outer = np.zeros(70000, oracle.TUPLE)
    outer["key"] = 4242
    outer["rid"] = np.arange(len(outer))
    cfg = rhj.make_config(4, 4, 28, True)
    res = check_against_oracle(rhj, oracle, cfg, [inner], [outer])
    assert res[0].matches == 70000


def test_build_side_larger_than_shared_memory_slice(rhj, oracle):
    """A build sub-partition larger than build_capacity is joined slice by slice."""
    inner = oracle.fill_perm(50000, 0, 50000, seed=5)
    outer = oracle.fill_perm(50000, 0, 50000, seed=6, start_rid=7)
    cfg = rhj.make_config(2, 2, 30, True, build_capacity=1024)  # 16 sub-partitions of ~3125 > 1024
    res = check_against_oracle(rhj, oracle, cfg, [inner], [outer])
    assert res[0].matches == 50000


@pytest.mark.parametrize("ni,no", [(0, 0), (0, 100), (100, 0), (1, 1), (7, 5), (4097, 4095)])
def test_empty_and_ragged(rhj, oracle, ni, no):
    inner = oracle.fill_perm(ni, 0, max(ni, 1), seed=1)
    outer = oracle.fill_perm(no, 0, max(ni, 1), seed=2)
    for two in (True, False):
        check_against_oracle(rhj, oracle, rhj.make_config(3, 2, 29, two), [inner], [outer])


def test_ragged_ranks(rhj, oracle):
    """Ranks with different local sizes, one of them empty; nranks not dividing P evenly."""
    sizes_i, sizes_o = [5000, 0, 12345], [7000, 3000, 1]
    inner = [oracle.fill_perm(s, sum(sizes_i[:i]), 17345, seed=11) for i, s in enumerate(sizes_i)]
    outer = [oracle.fill_perm(s, sum(sizes_o[:i]), 17345, seed=12, start_rid=5) for i, s in enumerate(sizes_o)]
    check_against_oracle(rhj, oracle, rhj.make_config(4, 3, 28, True), inner, outer)


def test_packed_8_byte_input(rhj, oracle):
    """BASELINE config 5 layout: 8-byte tuples (key<<32 | rid), 1:16 build:probe."""
    inner = oracle.fill_perm(8000, 0, 8000, seed=21)
    outer = oracle.fill_perm(128000, 0, 128000, seed=22, modulo=8000)
    cfg = rhj.make_config(5, 5, 27, True, tuple_bytes=8)
    res = check_against_oracle(rhj, oracle, cfg, [inner], [outer])
    assert res[0].matches == 128000


def test_warp_aggregated_histogram_flag(rhj, oracle):
    inner, outer = oracle.generate_reference(2, 30000, 30000)
    cfg = rhj.make_config(6, 5, 26, True, flags=rhj.FLAG_HIST_WARP_AGGREGATE)
    check_against_oracle(rhj, oracle, cfg, inner, outer)


def test_oversized_rid_carries_like_reference(rhj, oracle):
    """rid >= 2^(F_net+PAYLOAD_BITS) corrupts the key field through `+` exactly as
    tasks/NetworkPartitioning.cpp:128-129 does."""
    inner = oracle.fill_perm(3000, 0, 3000, seed=31)
    outer = oracle.fill_perm(3000, 0, 3000, seed=32)
    outer["rid"][::7] += np.uint64(1 << 33)
    check_against_oracle(rhj, oracle, rhj.make_config(3, 3, 29, True), [inner], [outer])


def test_rhj_join_single_call_and_repeat(rhj, oracle):
    """The fused entry point (no host sync between stages) twice on one context."""
    import torch
    cfg = rhj.make_config(8, 8, 24, True)
    inner, outer = oracle.generate_reference(1, 200000, 200000)
    want = oracle.join(inner, outer, ocfg(oracle, cfg))[0].astuple()
    ctx = rhj.Context(cfg)
    di = torch.from_numpy(inner[0].view(np.int64).copy()).cuda()
    do = torch.from_numpy(outer[0].view(np.int64).copy()).cuda()
    for _ in range(2):
        assert ctx.join(di, 200000, do, 200000).astuple() == want
    t = ctx.timings()
    assert t.jtotal_us > 0 and t.jhist_us > 0 and t.jmpi_us > 0 and t.jproc_us > 0
    ctx.close()


def test_join_host_buffers(rhj, oracle):
    cfg = rhj.make_config(6, 6, 26, True, flags=rhj.FLAG_STAGE_HOST_INPUTS)
    inner, outer = oracle.generate_reference(1, 100000, 100000)
    want = oracle.join(inner, outer, ocfg(oracle, cfg))[0].astuple()
    ctx = rhj.Context(cfg)
    assert ctx.join_host(inner[0], 100000, outer[0], 100000).astuple() == want
    ctx.close()


def test_join_host_pipeline_many_pieces(rhj, oracle, monkeypatch):
    """rhj_join_host copies in pieces and partitions the inner relation while the outer one is still on its
    way; with small pieces (ablation switch) every relation crosses in dozens of them, ragged tail included."""
    monkeypatch.setenv("RHJ_H2D_PIECE_KB", "256")
    cfg = rhj.make_config(6, 6, 26, True, chunk_tuples=8192, flags=rhj.FLAG_STAGE_HOST_INPUTS | rhj.FLAG_DEBUG_ENV)
    inner = oracle.fill_perm(300001, 0, 300001, seed=5)
    outer = oracle.fill_zipf(700003, 0, 300001, 1.0, seed=6)
    want = oracle.join([inner], [outer], ocfg(oracle, cfg))[0].astuple()
    ctx = rhj.Context(cfg)
    pin_i, pin_o = rhj.PinnedBuffer(inner.nbytes), rhj.PinnedBuffer(outer.nbytes)
    pin_i.array[:] = inner.view(np.uint8)
    pin_o.array[:] = outer.view(np.uint8)
    for _ in range(2):
        assert ctx.join_host(pin_i.ptr, len(inner), pin_o.ptr, len(outer)).astuple() == want
    t = ctx.timings()
    assert t.h2d_us > 0 and t.jtotal_us >= t.h2d_us * 0.5
    # device-resident inputs on the same context afterwards
    import torch
    di = torch.from_numpy(inner.view(np.int64).copy()).cuda()
    do = torch.from_numpy(outer.view(np.int64).copy()).cuda()
    assert ctx.join(di, len(inner), do, len(outer)).astuple() == want
    assert ctx.timings().h2d_us == 0
    ctx.close()
    pin_i.free()
    pin_o.free()


@pytest.mark.parametrize("two_level", [True, False])
def test_probe_side_in_pieces_larger_than_the_plan(rhj, oracle, two_level):
    """an outer relation several times larger than the planned maximum and the receive
    window, streamed from host memory in pieces against an inner relation that is partitioned once; count and
    all four sums equal the oracle's join of the whole relations (the join is additive over the probe side)."""
    import torch
    cfg = rhj.make_config(5, 4, 27, two_level, chunk_tuples=8192)
    inner = oracle.fill_perm(40000, 0, 40000, seed=5)
    outer = oracle.fill_zipf(500003, 0, 40000, 1.0, seed=6)
    want = oracle.join([inner], [outer], ocfg(oracle, cfg))[0].astuple()
    piece = 65536
    ctx = rhj.Context(cfg)
    ctx.plan(len(inner), piece)                       # windows and maxima for ONE piece of the outer relation
    with pytest.raises(rhj.RhjError):                 # the whole outer relation does not fit this plan
        ctx.histogram(torch.zeros(2, dtype=torch.int64, device="cuda"), 1,
                      torch.zeros(2, dtype=torch.int64, device="cuda"), len(outer))
    di = torch.from_numpy(inner.view(np.int64).copy()).cuda()
    rel = rhj.Relation(di, len(inner), len(inner))
    for _ in range(2):                                # twice: every piece restarts the outer-side state
        got = rhj.join_probe_pieces(ctx, rel, rhj.host_pieces(outer, piece, "cuda:0"))
        assert got.astuple() == want
    # an empty outer relation, and pieces that are already on the device
    assert rhj.join_probe_pieces(ctx, rel, iter(())).matches == 0
    do = torch.from_numpy(outer.view(np.int64).copy()).cuda()
    dev_pieces = ((do[2 * b:], min(piece, len(outer) - b)) for b in range(0, len(outer), piece))
    assert rhj.join_probe_pieces(ctx, rel, dev_pieces).astuple() == want
    ctx.close()


def test_hot_split_join_single_rank(rhj, oracle):
    """HotSplitJoin on one rank: the hot partitions go through the second context, the rest through the first;
    the sum equals the plain join (and the oracle's statement of the rule, which on one rank is the same)."""
    import torch
    inner = oracle.fill_perm(50000, 0, 50000, seed=9)
    outer = oracle.fill_zipf(200000, 0, 50000, 1.0, seed=21)
    cfg = rhj.make_config(6, 4, 26, True)
    ri = rhj.Relation.from_numpy(inner, 50000)
    ro = rhj.Relation.from_numpy(outer, 200000)
    hj = rhj.HotSplitJoin(1, 0, ri, ro, config=cfg, device=0, inv_share=32)
    assert bool(hj.hot.any()) and hj.s_hot.shape[0] > 0
    want, hot = oracle.join_hot_split([inner], [outer], ocfg(oracle, cfg), 32)
    assert np.array_equal(hj.hot.cpu().numpy(), hot)
    assert hj.join().astuple() == want[0] == oracle.join([inner], [outer], ocfg(oracle, cfg))[0].astuple()
    hj.close()


def test_window_overflow_is_reported(rhj, oracle):
    inner, outer = oracle.generate_reference(1, 50000, 50000)
    run = GpuRun(rhj, rhj.make_config(4, 4, 28, True), inner, outer, window_caps=(1000, 1000))
    with pytest.raises(rhj.RhjError) as e:
        run.run_all()
    assert e.value.code == -4
    run.close()


def test_default_window_capacity_under_skew_is_refused_exact_sizes_pass(rhj, oracle):
    """include/rhj.h, rhj_plan: the default window capacity (allocation_factor x local size) only holds for
    near-uniform partitions.  Two ranks, every probe tuple in one partition: the owner's window overflows
    -> RHJ_ERR_CAPACITY at rhj_get_result; with the exact sizes of rhj_host_offsets (what the reference
    computes after its histogram phase, data/Window.cpp:160-178) the same join equals the oracle."""
    cfg = rhj.make_config(4, 4, 28, True)
    inner = [oracle.fill_perm(20000, r * 20000, 40000, seed=3) for r in range(2)]
    outer = []
    for r in range(2):
        o = np.zeros(30000, oracle.TUPLE)
        o["key"] = 4242
        o["rid"] = np.arange(len(o)) + r * 30000
        outer.append(o)
    run = GpuRun(rhj, cfg, inner, outer, window_caps=(0, 0))
    with pytest.raises(rhj.RhjError) as e:
        run.run_all()
    assert e.value.code == -4
    run.close()
    P = 16
    caps = []
    for rels in (inner, outer):
        loc = np.ascontiguousarray(np.stack([oracle.local_histogram(t, ocfg(oracle, cfg)) for t in rels]))
        ws = np.zeros(2, np.uint64)
        rhj.lib().rhj_host_offsets(loc.ctypes.data, 2, 0, P, None, None, None, None, ws.ctypes.data)
        caps.append(int(ws.max()))
    run = GpuRun(rhj, cfg, inner, outer, window_caps=tuple(caps))
    res = run.run_all()
    want = oracle.join(inner, outer, ocfg(oracle, cfg))
    assert [r.astuple() for r in res] == [w.astuple() for w in want]
    run.close()


def test_device_generator_matches_oracle(rhj, oracle):
    import torch
    n, dom = 100000, 77777
    for tb in (16, 8):
        t = torch.empty(n * tb // 8, dtype=torch.int64, device="cuda")
        rhj.fill_perm(t, tb, n, 5, dom, seed=99, modulo=1000, start_key=3, start_rid=11)
        want = oracle.fill_perm(n, 5, dom, 99, modulo=1000, start_key=3, start_rid=11)
        got = t.cpu().numpy().view(np.uint64)
        if tb == 16:
            assert np.array_equal(got, want.view(np.uint64))
        else:
            assert np.array_equal(got, (want["key"] << np.uint64(32)) | want["rid"])


def test_apply_perm_matches_oracle(rhj, oracle):
    """rhj_apply_perm == orc_perm (the bijection orc_fill_zipf sends its ranks through)."""
    import torch
    dom, seed = 123457, 4177
    vals = np.random.default_rng(3).integers(0, dom, 50000).astype(np.uint64)
    t = torch.from_numpy(np.stack([vals, np.arange(len(vals), dtype=np.uint64)], axis=1).view(np.int64).copy()).cuda()
    rhj.apply_perm(t, len(vals), 2, dom, seed)
    got = t.cpu().numpy().view(np.uint64).reshape(-1, 2)
    want = np.array([oracle.lib().orc_perm(int(v), dom, seed) for v in vals], np.uint64)
    assert np.array_equal(got[:, 0], want) and np.array_equal(got[:, 1], np.arange(len(vals), dtype=np.uint64))


def test_zipf_outer(rhj, oracle):
    """BASELINE config 4 shape at oracle size: Zipf(1.0) probe keys."""
    n = 200000
    inner = oracle.fill_perm(n, 0, n, seed=41)
    outer = oracle.fill_zipf(n, 0, n, 1.0, seed=42)
    check_against_oracle(rhj, oracle, rhj.make_config(6, 6, 26, True), [inner], [outer], deep=False)


@pytest.mark.parametrize("nranks", [2, 3, 4])
@pytest.mark.parametrize("two_level", [True, False])
def test_balanced_assignment_all_stages(rhj, oracle, nranks, two_level):
    """RHJ_FLAG_BALANCED_ASSIGNMENT: assignment, offsets, windows, local partitions and per-rank results
    equal the oracle's under the same rule (skewed probe side, so the assignment differs from p % n)."""
    cfg = rhj.make_config(4, 3, 28, two_level, flags=rhj.FLAG_BALANCED_ASSIGNMENT)
    inner = [oracle.fill_perm(6000, r * 6000, 6000 * nranks, seed=9) for r in range(nranks)]
    outer = [oracle.fill_zipf(20000, r * 20000, 6000 * nranks, 1.0, seed=10 + r) for r in range(nranks)]
    res = check_against_oracle(rhj, oracle, cfg, inner, outer)
    plain = oracle.join(inner, outer, oracle.Config(4, 3, 28, two_level, 0))
    assert oracle.total(res) == oracle.total(plain)
    assert [r.matches for r in res] != [r.matches for r in plain]


def test_balanced_assignment_ties_and_uniform_sizes(rhj, oracle):
    """All partitions equally heavy (dense unique keys): every step of the greedy rule is a tie."""
    cfg = rhj.make_config(5, 4, 27, True, flags=rhj.FLAG_BALANCED_ASSIGNMENT)
    inner, outer = oracle.generate_reference(3, 30000, 30000)
    check_against_oracle(rhj, oracle, cfg, inner, outer)


@pytest.mark.parametrize("knobs", [(12, 3, 20, True), (3, 12, 20, True), (12, 12, 8, True), (12, 5, 20, False)])
def test_fan_out_of_4096_partitions(rhj, oracle, knobs):
    """The widest fan-out the configuration accepts (12 bits per pass) does not fit the offset-driven
    scatter's shared memory; the library switches to its first-generation kernels by itself."""
    cfg = rhj.make_config(*knobs)
    inner, outer = oracle.generate_reference(2, 50000, 50000)
    check_against_oracle(rhj, oracle, cfg, inner, outer, deep=False)
    inner = [oracle.fill_perm(40000, 0, 40000, seed=4)]
    outer = [oracle.fill_zipf(90000, 0, 40000, 1.0, seed=6)]
    check_against_oracle(rhj, oracle, cfg, inner, outer, deep=(knobs != (12, 12, 8, True)))
