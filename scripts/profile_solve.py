"""One full solve of a slab (the short command ncu profiles for the fused small-N kernel).
usage: python scripts/profile_solve.py [oh|oh_overlap|h2o150] [layers]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radiative_transfer_b200 import solver, synthetic
cfg = sys.argv[1] if len(sys.argv) > 1 else "oh"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 100
p = synthetic.config_h2o(L=L, N=150) if cfg == "h2o150" else synthetic.config_oh(L=L, overlap=(cfg == "oh_overlap"))
m = solver.Model(p)
t = time.time()
pop, st = m.calculate_populations()
print(f"{cfg} N={p.N} L={p.L}: {1e3 * (time.time() - t):.1f} ms, iters max {st['iters'].max()} mean {st['iters'].mean():.0f}, converged {st['converged'].sum()}")
m.close()
