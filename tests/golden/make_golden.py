"""Regenerate tests/golden/ref_n2_l3_p30_t1_r2.npz from the UNMODIFIED reference (oracle/_ref).

Runs the reference as 2 ranks (fork + shared-memory MPI shim) on its own generator
(main.cpp:94-104, seeds 1234+rank) with knobs F_net=2, F_loc=3, PAYLOAD_BITS=30, two-level on,
3000 x 3000 tuples per rank, dumps every intermediate its getters expose, and stores them in one
compressed file.  Needs /root/reference (or a prebuilt oracle/_ref binary); the .npz travels.

    python tests/golden/make_golden.py
"""
import glob
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

CASES = [
    # (net, loc, payload, two_level, ranks, inner_per_rank, outer_per_rank, gen)
    (2, 3, 30, 1, 2, 3000, 3000, "unique"),
    (5, 5, 27, 1, 1, 5000, 5000, "unique"),
    (5, 5, 27, 0, 4, 1000, 2000, "modulo"),
]


def main():
    for nb, lb, pb, tl, ranks, ni, no, gen in CASES:
        cfg = O.Config(nb, lb, pb, tl)
        with tempfile.TemporaryDirectory() as d:
            info = O.run_ref(cfg, ranks, ni, no, mode="stages", gen=gen, dump_dir=d)
            arrays = {"matches": np.array(info["matches"], np.uint64),
                      "meta": np.array([nb, lb, pb, tl, ranks, ni, no], np.uint64)}
            for path in sorted(glob.glob(os.path.join(d, "rank*_*.bin"))):
                name = os.path.basename(path)[:-4]
                dt = np.uint32 if name.endswith("assignment") else np.uint64
                arrays[name] = np.fromfile(path, dt)
        out = os.path.join(ROOT, "tests", "golden", f"ref_{cfg.tag()}_r{ranks}_{gen}.npz")
        np.savez_compressed(out, **arrays)
        print(out, os.path.getsize(out), "bytes", info["matches"])


if __name__ == "__main__":
    main()
