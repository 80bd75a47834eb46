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
    # the reference's per-rank 10-tuple (performance/Measurements.cpp:548-565): matches first, then times
    ten = [int(x) for x in re.search(r"\[RESULTS\] Rank 0:((?:\t\d+){10})", p.stdout).group(1).split()]
    assert ten[0] == tuples and ten[1] > 0 and ten[5] == 0 and ten[7] == 0
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


def test_host_mirror_two_ranks_fail_loudly_without_gpu(rhj, tmp_path):
    """Two forked ranks (data::ProcessComm): both reach rhj_init, both report the missing device, exit code 1."""
    import torch
    exe = build_host(rhj)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([exe, "1000", "5", "5", "27", "1", "0", "2"], capture_output=True, text=True, cwd=tmp_path,
                       timeout=120)
    assert p.returncode == 1
    assert p.stderr.count("rhj error -2") >= 1


@pytest.mark.gpu
@pytest.mark.parametrize("knobs", [(5, 5, 27, 1), (8, 8, 24, 1)])
def test_cpp_hashjoin_two_ranks_match_oracle(rhj, oracle, knobs, tmp_path):
    """rhj_main with 2 ranks = 2 forked processes, one GPU each, as `mpirun -np 2` of the reference:
    reference generator + distribute on the host, histograms all-gathered through shared memory,
    windows opened over CUDA IPC.  Per-rank counts and checksums must equal the oracle's."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    exe = build_host(rhj)
    n = 150000
    p = subprocess.run([exe, str(n)] + [str(k) for k in knobs] + ["0", "2"], capture_output=True, text=True,
                       cwd=tmp_path, timeout=300)
    assert p.returncode == 0, p.stderr
    counts = [int(x) for x in re.search(r"\[RESULTS\] Tuples:\t(\d+)\t(\d+)", p.stdout).groups()]
    inner, outer = oracle.generate_reference(2, n, n)
    want = oracle.join(inner, outer, oracle.Config(*knobs))
    for r in range(2):
        chk = tuple(int(x) for x in re.search(rf"\[RESULTS\] Checksum rank {r}:\t(\d+)\t(\d+)\t(\d+)\t(\d+)", p.stdout).groups())
        assert (counts[r],) + chk == want[r].astuple()
    assert sum(counts) == 2 * n
