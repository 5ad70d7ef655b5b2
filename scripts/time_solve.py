"""Un-profiled timing of the device-resident solve (development aid): ms per full-batch iteration from the difference of
two capped solves (all layers still active), and the whole converged solve.
usage: python scripts/time_solve.py [config] [layers] [reps]   (RT_B200_LIB / RT_B200_LU_GROUPS select variants)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radiative_transfer_b200 import solver, synthetic

cfg = sys.argv[1] if len(sys.argv) > 1 else "ch3oh"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 200
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
if cfg in ("ch3oh", "h2o"):
    p = {"ch3oh": synthetic.config_ch3oh, "h2o": synthetic.config_h2o}[cfg](L=L)
elif cfg == "h2o150":
    p = synthetic.config_h2o(L=L, N=150)
else:
    p = synthetic.config_oh(L=L, overlap=(cfg == "oh_overlap"))
m = solver.Model(p)

def timed(max_iter):
    best = 1e30
    for _ in range(reps):
        m.timer_begin()
        m.solve_device(max_iter=max_iter)
        best = min(best, m.timer_end())
    return best

timed(3)
t4, t12 = timed(4), timed(12)
tfull = timed(None)
st = m.get_status()
pop = m.get_populations()
print(os.environ.get("RT_B200_LIB", "default"), "groups", os.environ.get("RT_B200_LU_GROUPS", "default"),
      f"N={p.N} L={p.L}: {(t12 - t4) / 8:.3f} ms per full-batch iteration; converged solve {tfull:.2f} ms, "
      f"iters max {int(st['iters'].max())} mean {st['iters'].mean():.2f} converged {int(st['converged'].sum())} "
      f"checksum {float(np.abs(pop).sum()):.12f} {float((pop * np.arange(1, p.N + 1)).sum()):.9f}")
m.close()
