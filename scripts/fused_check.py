"""fused vs multi-kernel path on the small problems (development aid): run twice with/without RT_B200_NO_FUSED."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radiative_transfer_b200 import solver, synthetic, abi
for name in sys.argv[1:] or ["h2o_small", "oh_overlap", "ch3oh_small"]:
    p = synthetic.small(name, L=3)
    m = solver.Model(p)
    t = time.time(); pop, st = m.calculate_populations(); t1 = time.time() - t
    t = time.time(); pop2, st2 = m.calc_molecular_populations(acceleration=p.accel); t2 = time.time() - t
    print(name, "fused" if not os.environ.get("RT_B200_NO_FUSED") else "multi", st["iters"], st["converged"], st["rel_error"], f"{t1:.3f}s | retry:", st2["iters"], st2["converged"], f"{t2:.3f}s", flush=True)
    np.save(f"gpurun_out/pop_{name}_{'multi' if os.environ.get('RT_B200_NO_FUSED') else 'fused'}.npy", pop2)
    m.close()
