"""SASS opcode histogram of the hot kernels in the built library (cuobjdump, no GPU needed).
  python tools/sass_histogram.py > profiles/sass_r02_opcodes.txt
Evidence for what the kernels are made of: bulk asynchronous copies (UBLKCP = cp.async.bulk, the
non-tensor TMA form), L2 bulk prefetch (UBLKPF), mbarriers (SYNCS), shared-memory atomics (ATOMS),
16-byte global accesses (LDG/STG .128); no tensor-core instructions (the path has no contraction).
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-radxi-hash-join-on-gpus_b200", "librhj_b200.so")
HOT = [("K1 network histogram", r"k_histogramINS_11InputSourceILi16EEENS_12ChunkMapFlatELb0E"),
       ("K2 network scatter", r"k_scatter_wcINS_11InputSourceILi16EEENS_12ChunkMapFlatELi512ELi8ELi48E"),
       ("K3a local histogram", r"k_histogramINS_11ValueSourceENS_11ChunkMapSegELb0E"),
       ("K3b local scatter", r"k_scatter_wcINS_11ValueSourceENS_11ChunkMapSegELi512ELi8ELi48E"),
       ("K4 build/probe", r"k_build_probe3ILi5120ELi512ELb1ELb0E"),
       ("Kx push (staged exchange)", r"k_pushEPKm")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], check=True, capture_output=True, text=True).stdout
    funcs = {}
    name = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            funcs[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            funcs[name][m.group(1)] += 1
    print(f"# SASS opcode histogram, {os.path.basename(LIB)} (sm_100a), by tools/sass_histogram.py")
    for label, pat in HOT:
        for fn, cnt in funcs.items():
            if re.search(pat, fn):
                total = sum(cnt.values())
                print(f"\n## {label}: {fn}  ({total} instructions)")
                groups = collections.Counter()
                for op, c in cnt.items():
                    groups[op.split(".")[0]] += c
                print("   by mnemonic: " + ", ".join(f"{op} {c}" for op, c in groups.most_common()))
                notable = [op for op in cnt if re.match(r"(UBLKCP|UBLKPF|SYNCS|ATOMS|ATOMG|RED|LDG\.E.*128|STG\.E.*128|LDS\.128|BAR|UTMA|UTC|HMMA|MATCH|VOTE|SHFL)", op)]
                for op in sorted(notable):
                    print(f"   {op:40s} {cnt[op]}")
                break
        else:
            print(f"\n## {label}: no function matches {pat}")
    tensor = sum(c for f in funcs.values() for op, c in f.items() if re.match(r"(UTC|HMMA|HGMMA|UTMALDG)", op))
    print(f"\n# tensor-core / tensor-map instructions in the whole library: {tensor}")


if __name__ == "__main__":
    sys.exit(main())
