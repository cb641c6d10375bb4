"""The reference's own benchmark scenario (reference src/bin/crm_bench.rs:7-80) run through the CPU oracle,
with the open restatement decisions of SURVEY §9 flipped one at a time.  Records WHEN the run ends and how
(rods alive, error kind), which says which upstream facts matter for the "default mechanics cannot carry a
dense colony" finding.  CPU only; test infrastructure (uses oracle/)."""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from cr_mech_coli_b200 import _capi  # noqa: E402
from cr_mech_coli_b200.testing import morse_params  # noqa: E402
from oracle import oracle  # noqa: E402
from oracle.generate_positions import generate_positions  # noqa: E402


def scenario():
    n, nv = 16, 8
    L = float(np.sqrt(np.float32(n)) * np.float32(75.0))
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(4, 4), t_max=150.0, dt=0.1, n_saves=10)
    pos = np.asarray(generate_positions(n, 3.0, (L, L), 2.5, rng_seed=0, dx=(0.0, 0.0), randomize_positions=0.1,
                                        n_vertices=nv), dtype=np.float32)
    par = morse_params(n, growth_rate=0.1)  # AgentSettings() mechanics, Morse(3, 0.5, 10, 0.1), thresholds 6/6
    return cfg, pos, par


def run(label, order=2, spring_w=1.0, mie_textbook=0, max_rods=20000):
    cfg, pos, par = scenario()
    lib = oracle.library(openmp=True)
    lib.fn("set_mie_textbook")(mie_textbook)
    o = oracle.sim(cfg, openmp=True)
    lib.check(lib.fn("sim_set_solver_order")(o.h, order))
    lib.check(lib.fn("sim_set_spring_end_weight")(o.h, spring_w))
    o.set_agents(pos, np.zeros_like(pos), par)
    n_it = int(round(cfg.t_max / cfg.dt))
    out = {"variant": label, "stored_increments": order, "spring_end_weight": spring_w, "iterations_planned": n_it}
    hist = []
    try:
        while o.iteration < n_it:
            o.step(10)
            if o.iteration % 100 == 0:
                hist.append([o.iteration, o.n_rods])
            if o.n_rods > max_rods:
                out["stopped"] = f"more than {max_rods} rods (CPU budget)"
                break
        else:
            out["stopped"] = "reached t_max"
    except _capi.SimulationError as e:
        out["stopped"] = str(e)[:160]
    out["iteration"] = o.iteration
    out["t"] = o.iteration * float(cfg.dt)
    out["rods"] = o.n_rods
    out["rods_by_iteration"] = hist
    pos_now = o.get_state()[0]
    out["z_range"] = [float(pos_now[:, :, 2].min()), float(pos_now[:, :, 2].max())]
    lib.fn("set_mie_textbook")(0)
    o.close()
    return out


if __name__ == "__main__":
    rows = [run("as restated (AB3, full spring force per end)"),
            run("AB2 (one stored increment)", order=1),
            run("Euler", order=0),
            run("half spring force per end", spring_w=0.5),
            run("AB2 + half spring force", order=1, spring_w=0.5)]
    for r in rows:
        print(json.dumps(r))
