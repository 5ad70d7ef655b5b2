"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libref_oracle.so, built by
oracle/Makefile from /root/reference).  Run in the build container:  python tests/golden/make_golden.py
Each fixture carries the sha256 of its seeded inputs (tests/golden_cases.problem_digest)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import golden_cases as G  # noqa: E402
from oracle.loader import Checker  # noqa: E402


def main():
    ref = Checker("ref")
    for case in G.CASES:
        p = G.problem(case)
        out = G.compute(ref, p)
        out["digest"] = np.frombuffer(G.problem_digest(p).encode(), dtype=np.uint8)
        f = os.path.join(G.GOLDEN_DIR, case + ".npz")
        np.savez_compressed(f, **out)
        print(case, "N", p.N, "L", p.L, "iters", out["solve_iters"].tolist(), "chain", out["chain_iters"].tolist(),
              os.path.getsize(f) // 1024, "KiB")


if __name__ == "__main__":
    main()
