"""C5 (BASELINE config 5): crm_fit ensemble, N_SIMS independent 64-rod sims batched on one GPU.
Prints one JSON line: sim-steps/s, agent-steps/s, ms per batched step; CPU oracle on a sample."""
import argparse
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from cr_mech_coli_b200 import _capi
from cr_mech_coli_b200.testing import lattice_colony, morse_params

ap = argparse.ArgumentParser()
ap.add_argument("--sims", type=int, default=512)
ap.add_argument("--rods", type=int, default=64)
ap.add_argument("--steps", type=int, default=200)
ap.add_argument("--cpu-sims", type=int, default=4)
ap.add_argument("--full-cost", action="store_true",
                help="also time one whole population cost evaluation (t_max 17.5, dt 0.002 = 8750 steps, device-side costs)")
args = ap.parse_args()
lib = _capi.product()
cfg = _capi.make_config(domain_size=(200.0, 200.0), n_voxels=(1, 1), dt=0.002, t_max=1e6, n_saves=0)
rng = np.random.default_rng(0)
pos, par, sim_of = [], [], []
for m in range(args.sims):
    p = lattice_colony(args.rods, domain=200.0, seed=1000 + m, spacing=(24.0, 4.0))
    q = morse_params(args.rods, radius=float(rng.uniform(1.2, 1.8)), strength=float(rng.uniform(0.05, 0.2)),
                     potential_stiffness=float(rng.uniform(0.3, 0.8)), cutoff=6.0, damping=float(rng.uniform(0.5, 2.0)),
                     growth_rate=0.0)
    q["spring_length_threshold"] = 1e30
    pos.append(p); par.append(q); sim_of.append(np.full(args.rods, m, np.uint32))
P, Q, S = np.concatenate(pos), np.concatenate(par), np.concatenate(sim_of)
g = _capi.Sim(lib, cfg)
g.set_agents_batched(P, np.zeros_like(P), Q, S, args.sims)
g.step(5)
ms = g.step_timed(args.steps)
stats = g.list_stats()
out = {"workload": f"C5: {args.sims} sims x {args.rods} rods, Morse, dt 0.002, batched on one GPU",
       "ms_per_batched_step": ms / args.steps, "sim_steps_per_s": args.sims * args.steps / (ms * 1e-3),
       "agent_steps_per_s": len(P) * args.steps / (ms * 1e-3), "neighbour_lists": stats}
if args.cpu_sims:
    from oracle import oracle
    t0 = time.perf_counter()
    for m in range(args.cpu_sims):
        o = oracle.sim(cfg)
        o.set_agents(pos[m], np.zeros_like(pos[m]), par[m])
        o.step(args.steps)
    dt = time.perf_counter() - t0
    out["cpu_oracle_sim_steps_per_s_1core"] = args.cpu_sims * args.steps / dt
g.close()
if args.full_cost:
    from cr_mech_coli_b200.crm_fit import costs_on_device
    cfg2 = _capi.make_config(domain_size=(200.0, 200.0), n_voxels=(1, 1), dt=0.002, t_max=17.5, n_saves=8)
    data = np.repeat(pos[0][None, :, :, :2], 3, axis=0).astype(np.float32)
    t0 = time.perf_counter()
    costs, status = costs_on_device(lib, cfg2, P, np.zeros_like(P), Q, S, args.sims, data, [3, 6, 9])
    dt = time.perf_counter() - t0
    out["population_cost_eval"] = {"seconds": dt, "steps": 8750, "members": args.sims, "failed_members": int((status != 0).sum()),
                                   "members_per_s": args.sims / dt, "finite": bool(np.isfinite(costs).all())}
print(json.dumps(out))
