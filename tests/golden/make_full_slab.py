"""Generates tests/golden/full_<case>.npz: the answer of the UNMODIFIED reference (oracle/_ref/libref_oracle.so, built by
oracle/Makefile from /root/reference) for EVERY layer of the BASELINE.json slabs at full size (tests/full_slab_cases.py).
Run in the build container (minutes of CPU time per case):
    python tests/golden/make_full_slab.py [case ...]
Each fixture carries the sha256 of its seeded inputs (tests/golden_cases.problem_digest): a generator change that alters
the inputs invalidates the fixture loudly.

Beside the answer, every fixture records how SENSITIVE TO ROUNDING each layer's answer is in the reference itself: the same
run repeated with the probes of oracle/shim/linear_algebra.h (the dense solve with fused multiply-subtracts -- same
algorithm, same order, one rounding instead of two -- and with the matrix entries moved by one unit in the last place),
`sens_*` = the largest deviation of any probe run from the answer.  The reference's own lu_matrix_solve is not in its
repository, so its rounding is unknown; a layer whose answer moves under these probes by more than
the parity tolerance cannot be pinned to 1e-6 by ANY implementation of the solve."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import full_slab_cases as F  # noqa: E402
import golden_cases as G  # noqa: E402
from oracle.loader import Checker  # noqa: E402


def main():
    ref = Checker("ref")
    cases = sys.argv[1:] or list(F.CASES)
    nthreads = os.cpu_count() or 1
    for case in cases:
        p = F.problem(case)
        mode = F.CASES[case][1]
        t = time.time()
        if mode == "chained":
            pop, st, _ = ref.calc_molecular_populations(p)  # the reference function itself (serial over layers)
        else:
            pop, st, _ = ref.solve(p, nthreads=nthreads)    # every layer from its own start, layers spread over the cores
        dt = time.time() - t
        # rounding-sensitivity map: probe runs against the answer above
        L = p.L
        s_pop, s_it, s_fl = np.zeros(L), np.zeros(L, dtype=np.int32), np.zeros(L, dtype=np.int32)
        probes = F.sens_probes(case)
        for v in probes:
            ref.set_lu_variant(v)
            try:
                if mode == "chained":
                    pop_v, st_v, _ = ref.calc_molecular_populations(p)
                else:
                    pop_v, st_v, _ = ref.solve(p, nthreads=nthreads)
            finally:
                ref.set_lu_variant(0)
            s_it = np.maximum(s_it, np.abs(st_v["iters"].astype(np.int64) - st["iters"]).astype(np.int32))
            s_fl |= (st_v["converged"] != st["converged"]).astype(np.int32)
            for l in range(L):
                big = np.isfinite(pop[l]) & (pop[l] > 1e-30)
                if not big.any() or not np.isfinite(pop_v[l]).all():
                    s_pop[l] = np.inf
                    continue
                s_pop[l] = max(s_pop[l], float(np.max(np.abs(pop_v[l] - pop[l])[big] / pop[l][big])))
        out = dict(pop=pop, iters=st["iters"].astype(np.int32), converged=st["converged"].astype(np.int32),
                   eq_error=st["eq_error"], rel_error=st["rel_error"], cpu_seconds=np.array(dt), threads=np.array(nthreads),
                   sens_pop=s_pop, sens_iters=s_it, sens_flag=s_fl, sens_probes=np.array(probes, dtype=np.int32),
                   digest=np.frombuffer(G.problem_digest(p).encode(), dtype=np.uint8))
        f = F.fixture_path(case)
        np.savez_compressed(f, **out)
        print(f"{case}: N {p.N} L {p.L} mode {mode}: converged {int(st['converged'].sum())}/{p.L}, iters mean "
              f"{st['iters'].mean():.1f} max {st['iters'].max()}, non-finite layers "
              f"{np.flatnonzero(~np.isfinite(pop).all(axis=1)).tolist()}, rounding-sensitive layers "
              f"{np.flatnonzero(F.sensitive(out)).tolist()}, {dt:.0f} s on {nthreads} threads, "
              f"{os.path.getsize(f) // 1024} KiB", flush=True)


if __name__ == "__main__":
    main()
