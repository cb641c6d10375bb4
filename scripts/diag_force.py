import sys; sys.path.insert(0, '.')
import numpy as np
from cr_mech_coli_b200 import _capi
from cr_mech_coli_b200.testing import *
from oracle import oracle
lib = _capi.product()
n = 2000
L = lattice_domain(n); nvx = max(1, int(L // 40))
cfg = _capi.make_config(domain_size=(L, L), n_voxels=(nvx, nvx))
pos = lattice_colony(n, seed=7)
for pot in ["morse", "mie", "mie_real"]:
    par = {"morse": morse_params(n, radius=1.5, cutoff=6.0), "mie": mie_params(n), "mie_real": mie_params(n, en=0.94, em=0.22)}[pot]
    o = oracle.sim(cfg); o.set_agents(pos, np.zeros_like(pos), par); fo, no = o.compute_forces()
    for exact in [1, 0]:
        g = _capi.Sim(lib, cfg); lib.lib.crm_sim_set_exact_math(g.h, exact)
        g.set_agents(pos, np.zeros_like(pos), par); fg, ng = g.compute_forces()
        d = np.abs(fg - fo); k = np.unravel_index(d.argmax(), d.shape)
        print(pot, "exact" if exact else "fast", "max|F|", np.abs(fo).max(), "maxabs err", d.max(), "rel", d.max()/np.abs(fo).max(), "at", k, fg[k], fo[k], "nn eq", (ng==no).all(),
              "pairs", g.read_timers(), o.L.fn("sim_pairs_in_cutoff")(o.h))
        print("   err quantiles", np.quantile(d, [0.5, 0.9, 0.99, 0.999, 0.9999]))
