"""Per-kernel-class CUDA-event times of one full-batch calc_new_pop (development aid).
usage: python scripts/time_step.py [config] [layers] [reps]   (RT_B200_LIB selects a library variant)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radiative_transfer_b200 import solver, synthetic

cfg = sys.argv[1] if len(sys.argv) > 1 else "ch3oh"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 200
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
p = {"ch3oh": synthetic.config_ch3oh, "h2o": synthetic.config_h2o}[cfg](L=L) if cfg in ("ch3oh", "h2o") else synthetic.config_oh(L=L, overlap=(cfg == "oh_overlap"))
m = solver.Model(p)
b = m.boundary_layer_populations()
ref, _ = m.calc_new_pop(b)
m.profile_reset(timing=True)
for _ in range(reps):
    new, eq = m.calc_new_pop(b)
prof = m.profile_get()
tot = sum(v["ms"] for v in prof.values())
print(os.environ.get("RT_B200_LIB", "default"), f"N={p.N} L={p.L} total {tot / reps:.3f} ms/step  sum-1 {abs(new.sum(axis=1) - 1).max():.1e} checksum {float(np.abs(new).sum()):.12f}")
print("   " + "  ".join(f"{k} {v['ms'] / reps:.3f}" for k, v in prof.items() if v["launches"]))
m.close()
