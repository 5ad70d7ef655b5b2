"""Development aid: rt_lim_luminosity at BASELINE size against the checker, per output array."""
import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import full_slab_cases as F, golden_cases as G
from oracle.loader import Checker, ref_available
from radiative_transfer_b200 import solver
for case in ("ch3oh","h2o","oh_overlap"):
    p=F.problem(case); gold=F.load(case); pop=np.array(gold["pop"])
    for l in np.flatnonzero(~np.isfinite(pop).all(axis=1)): pop[l]=pop[l-1]
    up,low=G.maser_lines(p,pop)
    chk=Checker("ref" if ref_available() else "port")
    want=chk.maser_postpass(p,pop,up,low)
    m=solver.Model(p); got=m.maser_postpass(pop,up,low); lim=m.lim_luminosity(pop,up,low,beaming_factor=1.0,lay_nb_hg=got["scalars"][:,4]); m.close()
    wl=chk.lim_luminosity(p,pop,up,low,beaming_factor=1.0,lay_nb_hg=want["scalars"][:,4])
    for k in ("lum_arr","loss_rate_arr","pump_rate_arr","lum","tau_sat","temp_sat"):
        a,b=lim[k],wl[k]; den=np.maximum(np.abs(b),1e-300); r=np.abs(a-b)/den
        print(case,k,"gpu-vs-ref max",np.nanmax(r) if r.size else 0)
