"""Ad-hoc GPU-vs-oracle report (development aid; the judged parity tests live in tests/)."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radiative_transfer_b200 import synthetic as S, solver, abi
from oracle.loader import Checker


def rel(a, b, floor=0.0):
    den = np.maximum(np.abs(a), np.abs(b))
    mask = den > floor
    if not mask.any():
        return 0.0
    return float(np.max(np.abs(a - b)[mask] / den[mask]))


def main():
    O = Checker("port")
    names = sys.argv[1:] or ["generic", "oh", "oh_overlap", "ch3oh_small", "ch3oh_mid", "h2o_small"]
    for name in names:
        p = S.small(name, L=6)
        m = solver.Model(p)
        b_o = O.boundary(p)
        b_g = m.boundary_layer_populations()
        Mo, dfo = O.assemble(p, b_o)
        Mg, dfg = m.assemble(b_o)
        scale = np.abs(Mo).max(axis=2, keepdims=True)
        no, eo = O.step(p, b_o)
        ng, eg = m.calc_new_pop(b_o)
        t = time.time()
        pg, sg = m.calculate_populations()
        tg = time.time() - t
        po, so, _ = O.solve(p, nthreads=8)
        print(f"{name:12s} N={p.N} bnd {rel(b_o, b_g, 1e-30):.1e} M {np.max(np.abs(Mo - Mg) / scale):.1e} df {np.max(np.abs(dfo - dfg)):.1e} "
              f"step {rel(no, ng, 1e-30):.1e} eq {rel(eo, eg):.1e} solve {[('%.0e' % rel(po[l], pg[l], 1e-30)) for l in range(p.L)]} "
              f"iters gpu {sg['iters'].tolist()} ora {so['iters'].tolist()} conv {sg['converged'].tolist()} {so['converged'].tolist()} t={tg:.2f}s", flush=True)
        m.close()


if __name__ == "__main__":
    main()
