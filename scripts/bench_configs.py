"""BASELINE.json configs[1] (C2: 2D colony growing from 16 rods to >= 1e4, Morse, one GPU) and configs[3]
(C4: 3D multilayer, 2e5 rods in stacked layers, produce_ml_config parameters: gel pressure, surface
friction, Wiener term, neighbour-reduced growth).  One JSON line per config; the CPU oracle (OpenMP) is
timed beside each on a bounded sample.  Workload definitions: SURVEY.md §8(d)."""
import argparse
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np

from cr_mech_coli_b200 import _capi
from cr_mech_coli_b200.testing import morse_params, multilayer_colony, multilayer_params, random_colony

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="c2,c4")
ap.add_argument("--c2-target", type=int, default=10_000)
ap.add_argument("--c2-growth", default="defaults:0.01,relaxed:0.03,relaxed:0.015,relaxed:0.007,relaxed:0.003")
ap.add_argument("--c4-rods", type=int, default=200_000)
ap.add_argument("--c4-steps", type=int, default=300)
ap.add_argument("--cpu-seconds", type=float, default=15.0)
ap.add_argument("--no-cpu", action="store_true")
args = ap.parse_args()
lib = _capi.product()


def cpu_rate(cfg, pos, par, seconds):
    """agent-steps/s of the OpenMP oracle from the same state, stepped for about ``seconds``."""
    import os
    from oracle import oracle
    o = oracle.sim(cfg, openmp=True)
    o.set_agents(pos, np.zeros_like(pos), par)
    o.step(1)
    t0 = time.perf_counter()
    done = 0
    while time.perf_counter() - t0 < seconds:
        o.step(1)
        done += 1
    dt = time.perf_counter() - t0
    return {"value": o.n_rods * done / dt, "unit": "agent-steps/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"{o.n_rods} rods x {done} steps ({dt:.1f} s)"}


MECH = {"defaults": dict(dt=0.1, z=1.25, kw={}),
        # produce_ml_config's mechanics in an exactly planar colony (z = 0: no out-of-plane rounding noise to amplify)
        "relaxed": dict(dt=0.05, z=0.0, kw=dict(damping=0.02, spring_tension=10.0, rigidity=20.0, potential_stiffness=1.0,
                                                 strength=0.03))}


def c2_attempt(mech, growth):
    L = 3000.0
    m = MECH[mech]
    cfg = _capi.make_config(domain_size=(L, L), domain_height=2.5, n_voxels=(60, 60), dt=m["dt"], t_max=1e6, n_saves=0)
    n0 = 16
    pos = random_colony(n0, 300.0, seed=3, z_spread=0.0)
    pos[:, :, :2] += np.float32(0.5 * (L - 300.0))
    pos[:, :, 2] = m["z"]
    par = morse_params(n0, growth_rate=growth, **m["kw"])
    par["growth_setter_kind"] = _capi.SETTER_EXPLICIT  # daughters get (1.1 g, 0.8 g): divisions desynchronise
    par["growth_setter_a"] = 1.1 * growth
    par["growth_setter_b"] = 0.8 * growth
    par["spring_length"] = np.random.default_rng(1).uniform(3.0, 5.5, n0)
    sim = _capi.Sim(lib, cfg, 0)
    sim.set_agents(pos, np.zeros_like(pos), par)
    chunk = max(10, int(1.0 / growth) // 10 * 10)
    curve, steps, t0 = [], 0, time.perf_counter()
    try:
        while sim.n_rods < args.c2_target and steps < 400000:
            n_before = sim.n_rods
            ms = sim.step_timed(chunk)
            steps += chunk
            curve.append({"step": steps, "rods": sim.n_rods, "ms_per_step": ms / chunk,
                          "agent_steps_per_s": 0.5 * (n_before + sim.n_rods) * chunk / (ms * 1e-3)})
        grow_wall = time.perf_counter() - t0
        n_a = sim.n_rods
        sim.enable_timers(True)
        ms = sim.step_timed(200)
        n_b = sim.n_rods
    except _capi.SimulationError as e:
        sim.close()
        return {"mechanics": mech, "growth_rate": growth, "rods_reached": curve[-1]["rods"] if curve else n0, "steps": steps,
                "error": str(e)[:60] + " ... " + str(e)[-40:]}, None
    tm = sim.read_timers()
    p, v, q = sim.get_state()[:3]
    seg = np.linalg.norm(np.diff(p, axis=1), axis=2)
    out = {"mechanics": mech, "mechanics_values": m["kw"] or "AgentSettings() defaults", "dt": m["dt"], "growth_rate": growth,
           "segment_length_range": [float(seg.min()), float(seg.max())], "colony_extent_um": float(np.ptp(p[:, :, 0])),
           "steps_to_target": steps, "wall_seconds_to_target": grow_wall, "rods_timed": [n_a, n_b],
           "ms_per_step": ms / 200, "agent_steps_per_s": 0.5 * (n_a + n_b) * 200 / (ms * 1e-3),
           "force_ms": tm["force_ms"] / max(tm["force_launches"], 1), "neighbour_lists": sim.list_stats(),
           "patched_divisions": sim.patch_count(), "z_range": [float(p[:, :, 2].min()), float(p[:, :, 2].max())],
           "curve": curve[:: max(1, len(curve) // 12)]}
    sim.close()
    return out, (cfg, p, q)


def c2():
    """With AgentSettings() defaults the reference's model cannot carry a colony of this size: growth
    compresses the interior faster than damping 1 lets it relax, rods buckle out of the 2.5-um-high domain
    (BoundaryError) or, kept in plane, collapse (CalcError) - in the oracle at the same iteration, for every
    growth rate from 0.1 down to 0.001 (profiles/configs_r01_c2_buckling.jsonl).  (mechanics, growth rate)
    pairs are tried in order; the first that reaches the target is the timed one, failures are reported."""
    out = {"workload": f"C2: 2D colony grown from 16 rods (Morse, Explicit 6/6 thresholds, domain 3000^2 x 2.5) until "
                       f">= {args.c2_target} rods, then 200 timed steps", "attempts": []}
    for spec in args.c2_growth.split(","):
        mech, growth = spec.split(":")
        res, state = c2_attempt(mech, float(growth))
        if state is None:
            out["attempts"].append(res)
            continue
        out.update(res)
        if not args.no_cpu:
            cfg, p, q = state
            keep = np.arange(min(len(p), 2000))  # the first 2000 slots: the oldest lineages
            out["cpu_baseline"] = cpu_rate(cfg, p[keep], q[keep], args.cpu_seconds)
        break
    return out


def c4():
    pos, L, H = multilayer_colony(args.c4_rods)
    nvx = max(1, int(L // 40.0))
    cfg = _capi.make_config(domain_size=(L, L), domain_height=H, n_voxels=(nvx, nvx), dt=0.025, t_max=1e6, n_saves=0,
                            gel_pressure=0.1, surface_friction=0.03, surface_friction_distance=0.3, rng_seed=0)
    par = multilayer_params(len(pos))
    par["spring_length_threshold"] = np.random.default_rng(5).uniform(4.0, 20.0, len(par)).astype(np.float32)
    sim = _capi.Sim(lib, cfg, 0)
    sim.set_agents(pos, np.zeros_like(pos), par)
    sim.step(20)
    sim.compute_forces()  # instrumented launch: ordered pairs within the cutoff
    pairs = sim.read_timers()
    sim.enable_timers(True)
    n_a = sim.n_rods
    ms = sim.step_timed(args.c4_steps)
    n_b = sim.n_rods
    tm = sim.read_timers()
    p = sim.get_state()[0]
    out = {"workload": f"C4: 3D multilayer, {len(pos)} rods in 3 stacked staggered layers, produce_ml_config parameters "
                       f"(Morse r=3 stiffness 1 strength 0.03 cutoff 10, damping 0.02, tension 10, rigidity 20, gel pressure "
                       f"0.1, surface friction 0.03, diffusion 0.03, neighbour reduction (200, 0.5)), dt 0.025, domain "
                       f"{L:.0f}^2 x {H:.1f}",
           "steps": args.c4_steps, "rods_timed": [n_a, n_b], "ms_per_step": ms / args.c4_steps,
           "agent_steps_per_s": 0.5 * (n_a + n_b) * args.c4_steps / (ms * 1e-3),
           "force_ms": tm["force_ms"] / max(tm["force_launches"], 1), "neighbour_lists": sim.list_stats(),
           "ordered_pairs_in_cutoff": pairs["pairs_in_cutoff"], "neighbours_per_rod": pairs["pairs_in_cutoff"] / n_a,
           "force_kernel_algorithmic_tflops": pairs["pairs_in_cutoff"] * 56 * 59.0 /
                                              (tm["force_ms"] / max(tm["force_launches"], 1) * 1e-3) * 1e-12,
           "patched_divisions": sim.patch_count(), "z_range_after": [float(p[:, :, 2].min()), float(p[:, :, 2].max())]}
    sim.close()
    if not args.no_cpu:
        ps, Ls, Hs = multilayer_colony(3000)
        cfgs = _capi.make_config(domain_size=(Ls, Ls), domain_height=H, n_voxels=(max(1, int(Ls // 40)),) * 2, dt=0.025,
                                 t_max=1e6, n_saves=0, gel_pressure=0.1, surface_friction=0.03,
                                 surface_friction_distance=0.3, rng_seed=0)
        out["cpu_baseline"] = cpu_rate(cfgs, ps, multilayer_params(len(ps)), args.cpu_seconds)
    return out


for name in args.config.split(","):
    try:
        print(json.dumps({"c2": c2, "c4": c4}[name]()), flush=True)
    except Exception as e:  # a failing config is reported, the other one still runs
        print(json.dumps({"config": name, "error": f"{type(e).__name__}: {e}"}), flush=True)
