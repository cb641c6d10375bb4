import sys; sys.path.insert(0, '.')
import numpy as np
from cr_mech_coli_b200 import _capi
from cr_mech_coli_b200.testing import *
from oracle import oracle
lib = _capi.product()
def compare(cfg, pos, par, label):
    g = _capi.Sim(lib, cfg); o = oracle.sim(cfg)
    g.set_agents(pos, np.zeros_like(pos), par); o.set_agents(pos, np.zeros_like(pos), par)
    g.run(); o.run()
    for (it, pg, vg, parg, idg, prg), (_, po, vo, paro, ido, pro) in zip(g.snapshots(), o.snapshots()):
        same = pg.shape == po.shape
        print(label, it, len(pg), len(po), "ids", bool(same and (idg == ido).all() and (prg == pro).all()),
              "maxdiff", float(np.abs(pg - po).max()) if same else None, "sl eq", bool(same and (parg["spring_length"] == paro["spring_length"]).all()))
cfg = _capi.make_config(n_saves=19)
pos = np.load('tests/golden/readme_positions_seed0.npy')
compare(cfg, pos, morse_params(4), "readme")
n, L = 36, 400.0
cfg = _capi.make_config(domain_size=(L, L), n_voxels=(3, 3), t_max=30.0, dt=0.1, n_saves=14)
pos = lattice_colony(n, spacing=(60.0, 40.0), margin=40.0, seed=11, angle_std=0.5)
par = morse_params(n, growth_rate=0.08)
rng = np.random.default_rng(4)
par["growth_rate"] = rng.uniform(0.05, 0.1, n); par["growth_setter_kind"] = 1; par["growth_setter_a"] = 0.09; par["growth_setter_b"] = 0.06
par["spring_length"] = rng.uniform(2.5, 3.5, n)
compare(cfg, pos, par, "growth")
n = 2000; L = lattice_domain(n); nvx = int(L // 40)
cfg = _capi.make_config(domain_size=(L, L), n_voxels=(nvx, nvx), t_max=1.0, dt=0.01, n_saves=9)
compare(cfg, lattice_colony(n, seed=7), mie_params(n, growth_rate=0.0), "mie-lattice")
