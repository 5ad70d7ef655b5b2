"""All layers of a BASELINE slab against the compiled reference (oracle/_ref) on the host cores: iteration counts, convergence
flags and populations (development aid; minutes of CPU time).  usage: python scripts/parity_full_slab.py [ch3oh|oh_overlap] [out.json]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radiative_transfer_b200 import solver, synthetic
from oracle.loader import Checker, ref_available

cfg = sys.argv[1] if len(sys.argv) > 1 else "ch3oh"
p = synthetic.config_ch3oh(L=200) if cfg == "ch3oh" else synthetic.config_oh(L=400, overlap=True)
m = solver.Model(p)
pop, st = m.calculate_populations()
m.close()
chk = Checker("ref" if ref_available() else "port")
t = time.time()
pop_o, st_o, _ = chk.solve(p, nthreads=os.cpu_count())
dt = time.time() - t
dit = np.abs(st["iters"].astype(int) - st_o["iters"].astype(int))
both = (st["converged"] == 1) & (st_o["converged"] == 1)
rel = np.full(p.L, np.nan)
bad = []
for l in range(p.L):
    big = np.isfinite(pop_o[l]) & (pop_o[l] > 1e-30)
    if not big.any() or not np.isfinite(pop[l]).all():
        bad.append(l)
        continue
    rel[l] = np.max(np.abs(pop[l] - pop_o[l])[big] / pop_o[l][big])
both &= np.isfinite(rel)
out = dict(layers_without_reference_answer=bad, config=cfg, N=int(p.N), L=int(p.L), checker=chk.kind, cpu_seconds=dt, threads=os.cpu_count(),
           flags_equal=int((st["converged"] == st_o["converged"]).sum()),
           iter_diff_le1=int((dit <= 1).sum()), iter_diff_max=int(dit.max()),
           layers_iter_diff_gt1=[(int(l), int(st["iters"][l]), int(st_o["iters"][l])) for l in np.flatnonzero(dit > 1)],
           pop_rel_err_le_1e6=int((rel[both] <= 1e-6).sum()), both_converged=int(both.sum()),
           pop_rel_err_max=float(rel[both].max()), pop_rel_err_median=float(np.median(rel[both])),
           layers_pop_err_gt_1e6=[(int(l), float(rel[l])) for l in np.flatnonzero(both & (rel > 1e-6))])
print(json.dumps(out))
if len(sys.argv) > 2:
    json.dump(out, open(sys.argv[2], "w"), indent=1)
