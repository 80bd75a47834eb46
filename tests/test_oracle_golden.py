"""The CPU oracle against golden vectors dumped from the UNMODIFIED reference
(tests/golden/*.npz, made by tests/golden/make_golden.py from oracle/_ref).  This is what
pins the oracle: generator output, histograms, offsets, window contents, local-pass output
and match counts must all be bit-identical to what the reference's own getters returned."""
import glob
import os

import numpy as np
import pytest

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def load(path):
    z = np.load(path)
    nb, lb, pb, tl, ranks, ni, no = (int(x) for x in z["meta"])
    gen = os.path.basename(path)[:-4].split("_")[-1]
    return z, (nb, lb, pb, tl), ranks, ni, no, gen


def relations(z, ranks, O):
    inner = [z[f"rank{r}_inner"].view(O.TUPLE) for r in range(ranks)]
    outer = [z[f"rank{r}_outer"].view(O.TUPLE) for r in range(ranks)]
    return inner, outer


def test_golden_present():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=os.path.basename)
def test_generator_reproduces_reference_relations(oracle, path):
    """data/Relation.cpp:63-141 + main.cpp:94-104 (glibc rand() Fisher-Yates, distribute)."""
    z, knobs, ranks, ni, no, gen = load(path)
    inner, outer = oracle.generate_reference(ranks, ni, no, gen=gen)
    rin, rout = relations(z, ranks, oracle)
    for r in range(ranks):
        assert np.array_equal(inner[r], rin[r])
        assert np.array_equal(outer[r], rout[r])


@pytest.mark.parametrize("path", GOLDEN, ids=os.path.basename)
def test_every_stage_matches_reference_dump(oracle, path):
    z, knobs, ranks, ni, no, gen = load(path)
    cfg = oracle.Config(*knobs)
    rels = dict(zip(("inner", "outer"), relations(z, ranks, oracle)))
    P = 1 << cfg.net_bits
    for name, rel in rels.items():
        loc = np.stack([oracle.local_histogram(t, cfg) for t in rel])
        pl = oracle.plan(loc, ranks)
        wins = oracle.network_partition(rel, cfg, pl)
        for r in range(ranks):
            assert np.array_equal(z[f"rank{r}_{name}_local_hist"], loc[r])
            assert np.array_equal(z[f"rank{r}_{name}_global_hist"], pl.global_)
            assert np.array_equal(z[f"rank{r}_{name}_base"], pl.base)
            assert np.array_equal(z[f"rank{r}_{name}_write"], pl.abs_[r])
            assert np.array_equal(z[f"rank{r}_assignment"], pl.assignment)
            assert np.array_equal(z[f"rank{r}_{name}_window"], wins[r])
            if cfg.two_level:
                for p in range(P):
                    if pl.assignment[p] != r:
                        continue
                    part = wins[r][int(pl.base[p]):int(pl.base[p] + pl.global_[p])]
                    h, pf, out = oracle.local_partition(part, cfg)
                    assert np.array_equal(z[f"rank{r}_{name}_p{p}_lhist"], h)
                    assert np.array_equal(z[f"rank{r}_{name}_p{p}_lprefix"], pf)
                    assert np.array_equal(z[f"rank{r}_{name}_p{p}_lout"], out)


@pytest.mark.parametrize("path", GOLDEN, ids=os.path.basename)
def test_match_counts_equal_reference(oracle, path):
    """HashJoin::RESULT_COUNTER per rank (operators/HashJoin.cpp:28, tasks/BuildProbe.cpp:115)."""
    z, knobs, ranks, ni, no, gen = load(path)
    inner, outer = relations(z, ranks, oracle)
    res = oracle.join(inner, outer, oracle.Config(*knobs))
    assert [r.matches for r in res] == [int(m) for m in z["matches"]]
    # and the checksum agrees with an independent sort-merge join over all ranks' tuples
    bf = oracle.bruteforce(np.concatenate(inner), np.concatenate(outer))
    assert oracle.total(res) == bf.astuple()


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=os.path.basename)
def test_cuda_path_on_golden_inputs(rhj, oracle, path):
    """The CUDA path on the reference-generated relations gives the reference's own counts."""
    from _util import check_against_oracle
    z, knobs, ranks, ni, no, gen = load(path)
    inner, outer = relations(z, ranks, oracle)
    cfg = rhj.make_config(knobs[0], knobs[1], knobs[2], bool(knobs[3]), chunk_tuples=8192)
    res = check_against_oracle(rhj, oracle, cfg, [a.copy() for a in inner], [a.copy() for a in outer])
    assert [r.matches for r in res] == [int(m) for m in z["matches"]]
