"""Growth variant of C3: the same 1e6-rod lattice with growth on and division thresholds spread so that
about 0.03 % of the rods divide in EVERY step (the regime of an asynchronously growing colony: the
persistent neighbour structure never survives a step).  Prints one JSON line."""
import argparse, json, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench
from cr_mech_coli_b200 import _capi

ap = argparse.ArgumentParser()
ap.add_argument("--rods", type=int, default=1_000_000)
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--warmup", type=int, default=10)
args = ap.parse_args()
lib = _capi.product()
cfg, pos, par = bench.workload(args.rods, seed=12345)
rng = np.random.default_rng(7)
par["growth_rate"] = 0.1
par["growth_setter_kind"] = _capi.SETTER_EXPLICIT
par["growth_setter_a"] = 0.1
par["growth_setter_b"] = 0.1
par["spring_length_threshold"] = rng.uniform(3.0, 6.0, len(par)).astype(np.float32)
par["threshold_setter_kind"] = _capi.SETTER_EXPLICIT
par["threshold_setter_a"] = 6.0
par["threshold_setter_b"] = 6.0
sim = _capi.Sim(lib, cfg, 0)
sim.set_agents(pos, np.zeros_like(pos), par)
sim.step(args.warmup)
n0 = sim.n_rods
sim.enable_timers(True)
t0 = time.perf_counter()
ms = sim.step_timed(args.steps)
wall = time.perf_counter() - t0
n1 = sim.n_rods
tm = sim.read_timers()
print(json.dumps({"workload": "C3 growth variant: divisions in every step", "rods_start": n0, "rods_end": n1,
                  "divisions_per_step": (n1 - n0) / args.steps, "ms_per_step_device": ms / args.steps,
                  "ms_per_step_wall": wall * 1e3 / args.steps, "force_ms": tm["force_ms"] / max(tm["force_launches"], 1),
                  "agent_steps_per_s": 0.5 * (n0 + n1) * args.steps / wall, "neighbour_lists": sim.list_stats()}))
