"""The C++ host mirror (distributed-radxi-hash-join-on-gpus_b200/host): HashJoin / Relation /
Pool / Task classes with the reference's names, driving the CUDA path through the C ABI.
`rhj_main` plays main.cpp: reference generator (srand(1234), rand() Fisher-Yates) -> join ->
[RESULTS] lines.  Its count and checksum must equal the oracle's on the same generated data."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "distributed-radxi-hash-join-on-gpus_b200", "host")


def build_host(rhj):
    subprocess.check_call(["make", "-s", "-C", HOST])
    return os.path.join(HOST, "rhj_main")


def test_host_mirror_builds_and_fails_loudly_without_gpu(rhj):
    import torch
    exe = build_host(rhj)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([exe, "1000"], capture_output=True, text=True)
    assert p.returncode == 1 and "rhj error -2" in p.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("knobs", [(5, 5, 27, 0), (5, 5, 27, 1), (8, 8, 24, 1)])
def test_cpp_hashjoin_matches_oracle(rhj, oracle, knobs, tmp_path):
    exe = build_host(rhj)
    n = 200000 if knobs[3] else 30000
    p = subprocess.run([exe, str(n)] + [str(k) for k in knobs], capture_output=True, text=True, cwd=tmp_path)
    assert p.returncode == 0, p.stderr
    tuples = int(re.search(r"\[RESULTS\] Tuples:\t(\d+)", p.stdout).group(1))
    chk = tuple(int(x) for x in re.search(r"\[RESULTS\] Checksum:\t(\d+)\t(\d+)\t(\d+)\t(\d+)", p.stdout).groups())
    inner, outer = oracle.generate_reference(1, n, n)
    want = oracle.join(inner, outer, oracle.Config(*knobs))[0].astuple()
    assert (tuples,) + chk == want
    perf = open(tmp_path / "experiment-1" / "0.perf").read()
    for key in ("JTOTAL", "JHIST", "JMPI", "JPROC"):  # performance/Measurements.cpp:136-142
        assert re.search(rf"^{key}\t\d+\tus$", perf, re.M)


@pytest.mark.gpu
@pytest.mark.parametrize("knobs", [(5, 5, 27, 0), (8, 8, 24, 1)])
def test_cpp_hashjoin_materialises_pairs(rhj, oracle, knobs, tmp_path):
    """HashJoin::RESULT_PAIR_CAPACITY: join() also writes (rid_R, rid_S); rhj_main folds the pairs it
    copied back: their number, the two rid sums (must equal the checksum's) and how many pairs have
    different keys at the two rids (none)."""
    exe = build_host(rhj)
    n = 100000 if knobs[3] else 30000
    p = subprocess.run([exe, str(n)] + [str(k) for k in knobs] + [str(n + 10)], capture_output=True, text=True,
                       cwd=tmp_path)
    assert p.returncode == 0, p.stderr
    chk = tuple(int(x) for x in re.search(r"\[RESULTS\] Checksum:\t(\d+)\t(\d+)\t(\d+)\t(\d+)", p.stdout).groups())
    cnt, sr, ss, bad = (int(x) for x in re.search(r"\[RESULTS\] Pairs:\t(\d+)\t(\d+)\t(\d+)\t(\d+)", p.stdout).groups())
    assert cnt == n and bad == 0 and (sr, ss) == (chk[1], chk[2])
    # a buffer smaller than the output: the count stays complete, `capacity` pairs come back
    p = subprocess.run([exe, str(n)] + [str(k) for k in knobs] + ["1000"], capture_output=True, text=True, cwd=tmp_path)
    assert p.returncode == 0, p.stderr
    assert int(re.search(r"\[RESULTS\] Tuples:\t(\d+)", p.stdout).group(1)) == n
    cnt, _, _, bad = (int(x) for x in re.search(r"\[RESULTS\] Pairs:\t(\d+)\t(\d+)\t(\d+)\t(\d+)", p.stdout).groups())
    assert cnt == 1000 and bad == 0
