"""One calc_new_pop (intensity + assemble + LU) over a full-size slab: the short command ncu profiles.
usage: python scripts/profile_step.py [config] [layers] [reps]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radiative_transfer_b200 import solver, synthetic

cfg = sys.argv[1] if len(sys.argv) > 1 else "ch3oh"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 200
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
p = {"ch3oh": synthetic.config_ch3oh, "h2o": synthetic.config_h2o}[cfg](L=L) if cfg in ("ch3oh", "h2o") else synthetic.config_oh(L=L, overlap=(cfg == "oh_overlap"))
m = solver.Model(p)
b = m.boundary_layer_populations()
for _ in range(reps):
    t = time.time()
    new, eq = m.calc_new_pop(b)
    print(f"calc_new_pop N={p.N} L={p.L}: {1e3 * (time.time() - t):.1f} ms (incl. host copies), sum-1 = {abs(new.sum(axis=1) - 1).max():.1e}")
m.close()
