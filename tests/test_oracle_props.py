"""Properties of the CPU oracle that do not need the reference: independent sort-merge
join, closed forms of the unique 1:1 FK workload, generator bijection."""
import numpy as np
import pytest


def test_perm_is_a_bijection(oracle):
    for dom in (1, 2, 3, 17, 1000, 4096, 100003):
        t = oracle.fill_perm(dom, 0, dom, seed=dom + 1)
        assert np.array_equal(np.sort(t["key"]), np.arange(dom, dtype=np.uint64))
        assert np.array_equal(t["rid"], np.arange(dom, dtype=np.uint64))


def test_rng_is_glibc_rand(oracle):
    """srand(1234); rand() first values of glibc's TYPE_3 generator."""
    import ctypes
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(1234)
    want = [libc.rand() for _ in range(16)]
    g = oracle.Rng(1234)
    assert [g.next() for _ in range(16)] == want


@pytest.mark.parametrize("knobs", [(5, 5, 27, 1), (5, 5, 27, 0), (8, 8, 24, 1), (2, 3, 30, 1), (10, 6, 22, 1)])
@pytest.mark.parametrize("nranks", [1, 2, 3])
def test_unique_fk_closed_form(oracle, knobs, nranks):
    per = 20000
    inner, outer = oracle.generate_reference(nranks, per, per)
    res = oracle.join(inner, outer, oracle.Config(*knobs))
    G = nranks * per
    m, sk, sr, ss, _ = oracle.total(res)
    assert m == G
    assert sk == sr == ss == G * (G - 1) // 2
    P = 1 << knobs[0]
    if P % nranks == 0 and G % P == 0:
        # every partition holds exactly G/P tuples; rank r owns P/nranks of them
        assert [r.matches for r in res] == [G // nranks] * nranks


def test_join_equals_sort_merge_with_duplicates_and_misses(oracle):
    rng = np.random.default_rng(1)
    for knobs in ((3, 3, 29, 1), (3, 3, 29, 0), (6, 4, 26, 1)):
        r = np.zeros(3000, oracle.TUPLE)
        s = np.zeros(5000, oracle.TUPLE)
        r["key"] = rng.integers(0, 700, len(r))
        s["key"] = rng.integers(0, 1400, len(s))
        r["rid"] = np.arange(len(r))
        s["rid"] = np.arange(len(s)) + 12345
        res = oracle.join([r], [s], oracle.Config(*knobs))
        assert res[0].astuple() == oracle.bruteforce(r, s).astuple()
        # three ranks, arbitrary split
        res3 = oracle.join([r[:1000], r[1000:1500], r[1500:]], [s[:10], s[10:4000], s[4000:]], oracle.Config(*knobs))
        assert oracle.total(res3) == oracle.bruteforce(r, s).astuple()


def test_compressed_value_round_trip(oracle):
    """rid = value & mask, key = ((value >> shift) << F_net) | p."""
    cfg = oracle.Config(5, 5, 27, 1)
    t = oracle.fill_perm(10000, 0, 1 << 30, seed=9, start_rid=77)
    pl = oracle.plan(oracle.local_histogram(t, cfg)[None, :], 1)
    win = oracle.network_partition([t], cfg, pl)[0]
    keys, rids = [], []
    for p in range(32):
        part = win[int(pl.base[p]):int(pl.base[p] + pl.global_[p])]
        keys.append(((part >> np.uint64(32)) << np.uint64(5)) | np.uint64(p))
        rids.append(part & np.uint64(0xFFFFFFFF))
    got = np.sort(np.stack([np.concatenate(keys), np.concatenate(rids)], 1), 0)
    want = np.sort(np.stack([t["key"], t["rid"]], 1), 0)
    assert np.array_equal(got, want)


def test_empty_relations(oracle):
    e = np.zeros(0, oracle.TUPLE)
    t = oracle.fill_perm(100, 0, 100, seed=1)
    for a, b in ((e, e), (e, t), (t, e)):
        for tl in (0, 1):
            assert oracle.join([a], [b], oracle.Config(3, 2, 29, tl))[0].astuple() == (0, 0, 0, 0, 0)


@pytest.mark.parametrize("two_level", [False, True])
def test_materialised_pairs_equal_an_independent_equi_join(oracle, two_level):
    """join_pairs (the probe loop of BuildProbe.cpp:97-106 with an append) against pairs built
    independently with numpy from the raw tuples; 3 ranks, duplicates on both sides, misses."""
    rng = np.random.default_rng(5)
    inner, outer = [], []
    for r in range(3):
        a = np.zeros(4000, oracle.TUPLE)
        a["key"] = rng.integers(0, 3000, len(a))
        a["rid"] = np.arange(len(a)) + r * 10000
        b = np.zeros(7000, oracle.TUPLE)
        b["key"] = rng.integers(0, 5000, len(b))
        b["rid"] = np.arange(len(b)) + 500000 + r * 10000
        inner.append(a)
        outer.append(b)
    cfg = oracle.Config(3, 3, 29, two_level)
    res, pairs = oracle.join_pairs(inner, outer, cfg, 200000)
    R, S = np.concatenate(inner), np.concatenate(outer)
    want = []
    by_key = {}
    for k, rid in zip(R["key"].tolist(), R["rid"].tolist()):
        by_key.setdefault(k, []).append(rid)
    for k, rid in zip(S["key"].tolist(), S["rid"].tolist()):
        for rr in by_key.get(k, ()):
            want.append((rr, rid))
    got = sorted(tuple(x) for p in pairs for x in p.tolist())
    assert got == sorted(want)
    assert sum(r.matches for r in res) == len(want)
    for r in range(3):  # a pair lands on the rank that owns its key's partition (AssignmentMap.cpp:41-43)
        assert len(pairs[r]) == res[r].matches
    # capacity smaller than the output: everything is counted, only `cap` rows are stored
    res2, pairs2 = oracle.join_pairs(inner, outer, cfg, 10)
    assert [r.astuple() for r in res2] == [r.astuple() for r in res]
    assert all(len(p) == min(10, r.matches) for p, r in zip(pairs2, res2))


@pytest.mark.parametrize("nranks", [2, 3, 4])
def test_balanced_assignment_moves_partitions_not_results(oracle, nranks):
    """Under the size-balanced assignment (orc_config.balanced) each rank joins other partitions, so
    per-rank results change, but their sum is the join of the same relations: equal to p % n's sum and
    to the sort-merge checker; on a skewed probe side the busiest rank gets less than under p % n."""
    inner = [oracle.fill_perm(6000, r * 6000, 6000 * nranks, seed=9) for r in range(nranks)]
    outer = [oracle.fill_zipf(20000, r * 20000, 6000 * nranks, 1.0, seed=10 + r) for r in range(nranks)]
    rr = oracle.join(inner, outer, oracle.Config(4, 3, 28, True, 0))
    bal = oracle.join(inner, outer, oracle.Config(4, 3, 28, True, 1))
    assert oracle.total(rr) == oracle.total(bal)
    assert oracle.total(bal) == oracle.bruteforce(np.concatenate(inner), np.concatenate(outer)).astuple()
    assert max(r.matches for r in bal) <= max(r.matches for r in rr)
    assert [r.matches for r in bal] != [r.matches for r in rr]


def test_hot_split_total_equals_plain_join_and_balances(oracle):
    """oracle.join_hot_split: the split changes who counts a match, never the sum; the busiest rank's share drops."""
    world = 4
    cfg = oracle.Config(6, 3, 26, 1)
    inner = [oracle.fill_perm(6000, r * 6000, 6000 * world, seed=9) for r in range(world)]
    outer = [oracle.fill_zipf(30000, r * 30000, 6000 * world, 1.0, seed=20 + r) for r in range(world)]
    plain = oracle.join(inner, outer, cfg)
    split, hot = oracle.join_hot_split(inner, outer, cfg, 8)
    assert hot.any() and not hot.all()
    m = (1 << 64) - 1
    assert tuple(sum(s[i] for s in split) & m for i in range(5)) == oracle.total(plain)
    assert max(s[0] for s in split) < max(p.matches for p in plain)
    none, hot0 = oracle.join_hot_split(inner, outer, cfg, 0)
    assert not hot0.any() and [tuple(x) for x in none] == [p.astuple() for p in plain]
