"""The oracle against the reference itself (oracle/_ref, built from /root/reference by
oracle/build_ref.sh).  Skipped where neither the prebuilt binary nor the reference tree
exists; tests/golden/ covers the same ground with committed dumps."""
import glob
import os
import tempfile

import numpy as np
import pytest


def need_ref(oracle, cfg):
    if oracle.ref_binary(cfg) is None:
        pytest.skip("oracle/_ref not built and /root/reference absent")


@pytest.mark.parametrize("knobs,ranks,per", [((5, 5, 27, 1), 1, 200000), ((5, 5, 27, 1), 2, 100000),
                                             ((5, 5, 27, 0), 2, 20000), ((8, 8, 24, 1), 4, 100000),
                                             ((2, 3, 30, 1), 3, 7000)])
def test_counts_and_timing_fields(oracle, knobs, ranks, per):
    cfg = oracle.Config(*knobs)
    need_ref(oracle, cfg)
    info = oracle.run_ref(cfg, ranks, per, per, mode="join")
    inner, outer = oracle.generate_reference(ranks, per, per)
    res = oracle.join(inner, outer, cfg)
    assert info["matches"] == [r.matches for r in res]
    assert info["total_matches"] == ranks * per
    assert info["failed_ranks"] == 0 and info["join_us_max"] > 0


def test_stage_dump_matches_oracle_at_gpu_knobs(oracle):
    """8+8 radix bits (the B200 fan-out), 4 ranks: every intermediate the reference exposes."""
    cfg = oracle.Config(8, 8, 24, 1)
    need_ref(oracle, cfg)
    ranks, per = 4, 50000
    with tempfile.TemporaryDirectory() as d:
        info = oracle.run_ref(cfg, ranks, per, per, mode="stages", dump_dir=d)
        inner, outer = oracle.generate_reference(ranks, per, per)
        for name, rel in (("inner", inner), ("outer", outer)):
            loc = np.stack([oracle.local_histogram(t, cfg) for t in rel])
            pl = oracle.plan(loc, ranks)
            wins = oracle.network_partition(rel, cfg, pl)
            for r in range(ranks):
                g = lambda s, dt=np.uint64: np.fromfile(os.path.join(d, f"rank{r}_{name}_{s}.bin"), dt)
                assert np.array_equal(g("local_hist"), loc[r])
                assert np.array_equal(g("write"), pl.abs_[r])
                assert np.array_equal(g("window"), wins[r])
                for path in glob.glob(os.path.join(d, f"rank{r}_{name}_p*_lout.bin")):
                    p = int(os.path.basename(path).split("_p")[1].split("_")[0])
                    part = wins[r][int(pl.base[p]):int(pl.base[p] + pl.global_[p])]
                    assert np.array_equal(np.fromfile(path, np.uint64), oracle.local_partition(part, cfg)[2])
        assert info["total_matches"] == ranks * per
