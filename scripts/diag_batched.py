"""Diagnostic: forces of a batched ensemble at evolved states vs the oracle, and trajectory
divergence batched-GPU vs standalone-GPU vs oracle for one sim."""
import sys
import numpy as np
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from cr_mech_coli_b200 import _capi
from oracle import oracle
from test_gpu_batched import _ensemble

lib = _capi.product()
n_sims = 24
cfg = _capi.make_config(domain_size=(200.0, 200.0), n_voxels=(2, 2), t_max=25.0, dt=0.1, n_saves=4)
sims = _ensemble(n_sims)
states = []
for m, (p0, v0, par0) in enumerate(sims):
    o = oracle.sim(cfg)
    o.set_agents(p0, v0, par0)
    o.step(180)
    p, v, par, ids, _ = o.get_state()
    fo, no = o.compute_forces()
    states.append((p, v, par, fo, no))
pos = np.concatenate([s[0] for s in states]); vel = np.concatenate([s[1] for s in states]); par = np.concatenate([s[2] for s in states])
sim_of = np.concatenate([np.full(len(s[0]), m, np.uint32) for m, s in enumerate(states)])
g = _capi.Sim(lib, cfg)
g.set_agents_batched(pos, vel, par, sim_of, n_sims)
fg, ng = g.compute_forces()
o0 = 0
for m, s in enumerate(states):
    n = len(s[0]); f = fg[o0:o0 + n]; sc = np.abs(s[3]).max()
    err = np.abs(f - s[3]).max() / (sc if sc > 0 else 1)
    print(f"sim {m:2d} rods {n:2d} force scale {sc:.3e} rel err {err:.2e} nn ok {(ng[o0:o0+n] == s[4]).all()}")
    o0 += n
# trajectory: sim 17 standalone GPU vs oracle at several horizons
p0, v0, par0 = sims[17]
for steps in (100, 150, 200):
    o = oracle.sim(cfg); o.set_agents(p0, v0, par0); o.step(steps)
    s1 = _capi.Sim(lib, cfg); s1.set_agents(p0, v0, par0); s1.step(steps)
    s2 = _capi.Sim(lib, cfg); s2.set_skin(0.0); s2.set_agents(p0, v0, par0); s2.step(steps)
    s3 = _capi.Sim(lib, cfg); s3.set_exact_math(True); s3.set_agents(p0, v0, par0); s3.step(steps)
    po = o.get_state()[0]
    print(steps, "rods", len(po), "gpu-oracle", np.abs(s1.get_state()[0] - po).max(), "skin0-oracle", np.abs(s2.get_state()[0] - po).max(),
          "exact-oracle", np.abs(s3.get_state()[0] - po).max())
