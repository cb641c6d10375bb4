#!/usr/bin/env python
"""bench.py — agent-steps/s of the colony-mechanics step (BASELINE.json metric).

A "step" is one mechanics step (binning where due, rod-rod forces, springs/bending/integrator,
growth/division, boundary) over the whole synthetic colony.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c3|c2|c4|c5] [--impl reference]

Workloads (BASELINE.json configs); the default and the headline is c3:
    c3  2D monolayer, 1e6 rods x 8 vertices per GPU (weak) and ONE 1e6-rod colony over the N GPUs (strong,
        second value of the N > 1 line), Mie n=6 m=3, voxel slabs along x
    c2  2D colony grown from 16 rods to >= 1e4 by division (Morse), then K timed steps; single GPU
    c4  3D multilayer, 2e5 rods per GPU in three stacked layers (gel pressure, surface friction, Wiener
        term, neighbour-reduced growth), voxel slabs along x with halo exchange
    c5  crm_fit ensemble: 4096 independent 64-rod sims, 4096 / N per GPU, no communication (replicas)

Prints ONE JSON line (rank 0).  `value` = rods*K / device time with the colony resident in HBM;
`e2e` = the same through the C ABI with host buffers (upload + K steps + download in the timed region,
median of 3); `roofline` = the force kernel against the measured FP32 peak; `cpu_baseline` = the CPU
oracle on a bounded sample of the same workload.
"""

from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "agent-steps/sec at 1e6 rods on 1/2/4/8 B200; ms per mechanics step"
UNIT = "agent-steps/s"
# algorithmic work of the force kernel (SURVEY.md §8d / DESIGN.md): one point-segment
# evaluation = 65 FLOP (Mie) / 59 FLOP (Morse); 56 evaluations per ordered interacting pair
FLOP_PER_EVAL = {"mie": 65.0, "morse": 59.0}
EVALS_PER_ORDERED_PAIR = 56.0
MECH_BYTES_PER_ROD_STEP = 836.0
DEFAULT_RODS = {"c3": 1_000_000, "c2": 10_000, "c4": 200_000, "c5": 64}
C5_SIMS = 4096


def workload(n_rods: int, seed: int = 12345):
    """C3 on one GPU: (config, positions, parameters)."""
    from cr_mech_coli_b200 import _capi
    from cr_mech_coli_b200.testing import lattice_colony, lattice_domain, mie_params

    L = lattice_domain(n_rods)
    nvx = max(1, int(L // 40.0))  # reference voxels: edge >= rod length + cutoff
    cfg = _capi.make_config(domain_size=(L, L), domain_height=2.5, n_voxels=(nvx, nvx), dt=0.01, t_max=1e6,
                            n_saves=0)
    pos = lattice_colony(n_rods, seed=seed)
    par = mie_params(n_rods, growth_rate=0.0)
    par["spring_length_threshold"] = 1e30
    return cfg, pos, par


WORKLOAD_TEXT = {
    "c3": "C3: 2D monolayer, jittered nematic lattice (24x4 um spacing), Mie n=6 m=3, r=1.5, cutoff 6, "
          "8 vertices/rod, dt=0.01, growth 0",
    "c2": "C2: 2D colony grown from 16 rods by division to >= 1e4 rods (Morse r=3 cutoff 10; produce_ml_config "
          "mechanics: damping 0.02, tension 10, rigidity 20, stiffness 1, strength 0.03; growth 0.015/min, daughters "
          "1.1 g / 0.8 g, thresholds 6/6, dt 0.05, colony in the z = 0 plane), then the timed steps of the growing colony",
    "c4": "C4: 3D multilayer, three stacked staggered layers, produce_ml_config parameters (Morse r=3 stiffness 1 "
          "strength 0.03 cutoff 10, damping 0.02, tension 10, rigidity 20, gel pressure 0.1, surface friction 0.03, "
          "diffusion 0.03, neighbour reduction (200, 0.5), growth 0.005/min with thresholds 4..20), dt 0.025",
    "c5": "C5: crm_fit ensemble, independent 64-rod sims (Morse, per-sim radius / strength / stiffness / damping), "
          "dt 0.002, batched in one device state per GPU, no communication",
}


def workload_config(name: str, world: int, rods_per_gpu: int, rods_total: int, scaling: str) -> dict:
    """The `config` object of the line: identical for the GPU arm and the reference arm of one (config, N)."""
    if name in ("c3", "c4"):
        par = "1 GPU" if world == 1 else f"{world} voxel slabs along x, one per GPU, per-step NCCL halo exchange"
    elif name == "c5":
        par = f"{C5_SIMS} sims, {C5_SIMS // world} per GPU, replicas only (no communication)"
    else:
        par = "1 GPU" if world == 1 else f"{world} independent replicas (the config is a single-GPU one)"
    return {"workload": WORKLOAD_TEXT[name], "name": name, "rods_per_gpu": int(rods_per_gpu),
            "rods_total": int(rods_total), "parallelism": par, "scaling": scaling,
            "l2": "state larger than the 126 MB L2 (c3, c4) / resident in L2 (c2, c5); no explicit flush"}


def rods_per_gpu(name: str, rods: int, world: int) -> int:
    """Rods one GPU holds for `rods` requested (the generators round up to whole lattice columns)."""
    if name == "c3" and world > 1:
        ny = int(np.ceil(np.sqrt(world * rods * 24.0 / 4.0)))
        return int(np.ceil(rods / ny)) * ny
    if name == "c4":
        per = int(np.ceil(rods / 3))
        ny = int(np.ceil(np.sqrt(world * per * 27.0 / 6.5)))
        return 3 * int(np.ceil(per / ny)) * ny
    if name == "c5":
        return rods * (C5_SIMS // world)
    return rods


def make_workload(name: str, rods: int, rank: int, world: int, strong: bool = False):
    """This rank's part of the workload: dict(cfg, pos, vel, par, edges, g0, n_total, pot, sim_of, n_sims)."""
    from cr_mech_coli_b200 import _capi
    from cr_mech_coli_b200.testing import (lattice_colony, lattice_slab, mie_params, morse_params, multilayer_params,
                                           multilayer_slab, random_colony)

    w = {"sim_of": None, "n_sims": 1, "edges": [], "g0": 0, "grow_to": 0}
    if name == "c3":
        per = rods // world if strong else rods
        if world == 1:
            cfg, pos, par = workload(per)
        else:
            pos, domain, w["edges"], w["g0"] = lattice_slab(per, rank, world)
            nvx = (max(1, int(domain[0] // 40.0)), max(1, int(domain[1] // 40.0)))
            cfg = _capi.make_config(domain_size=domain, domain_height=2.5, n_voxels=nvx, dt=0.01, t_max=1e6, n_saves=0)
            par = mie_params(len(pos), growth_rate=0.0)
            par["spring_length_threshold"] = 1e30
        w.update(cfg=cfg, pos=pos, par=par, pot="mie")
    elif name == "c4":
        pos, domain, H, w["edges"], w["g0"] = multilayer_slab(rods, rank, world)
        nvx = (max(1, int(domain[0] // 40.0)), max(1, int(domain[1] // 40.0)))
        cfg = _capi.make_config(domain_size=domain, domain_height=H, n_voxels=nvx, dt=0.025, t_max=1e6, n_saves=0,
                                gel_pressure=0.1, surface_friction=0.03, surface_friction_distance=0.3, rng_seed=0)
        par = multilayer_params(len(pos))
        par["spring_length_threshold"] = np.random.default_rng(5 + rank).uniform(4.0, 20.0, len(par)).astype(np.float32)
        w.update(cfg=cfg, pos=pos, par=par, pot="morse")
    elif name == "c2":
        L = 3000.0
        cfg = _capi.make_config(domain_size=(L, L), domain_height=2.5, n_voxels=(60, 60), dt=0.05, t_max=1e6, n_saves=0)
        n0, growth = 16, 0.015
        pos = random_colony(n0, 300.0, seed=3 + rank, z_spread=0.0)
        pos[:, :, :2] += np.float32(0.5 * (L - 300.0))
        pos[:, :, 2] = 0.0
        par = morse_params(n0, growth_rate=growth, damping=0.02, spring_tension=10.0, rigidity=20.0,
                           potential_stiffness=1.0, strength=0.03)
        par["growth_setter_kind"] = _capi.SETTER_EXPLICIT  # daughters get (1.1 g, 0.8 g): divisions desynchronise
        par["growth_setter_a"] = 1.1 * growth
        par["growth_setter_b"] = 0.8 * growth
        par["spring_length"] = np.random.default_rng(1).uniform(3.0, 5.5, n0)
        w.update(cfg=cfg, pos=pos, par=par, pot="morse", grow_to=rods)
    elif name == "c5":
        n_sims = C5_SIMS // world
        cfg = _capi.make_config(domain_size=(200.0, 200.0), n_voxels=(1, 1), dt=0.002, t_max=1e6, n_saves=0)
        rng = np.random.default_rng(rank)
        P, Q, S = [], [], []
        for m in range(n_sims):
            p = lattice_colony(rods, domain=200.0, seed=1000 + rank * n_sims + m, spacing=(24.0, 4.0))
            q = morse_params(rods, radius=float(rng.uniform(1.2, 1.8)), strength=float(rng.uniform(0.05, 0.2)),
                             potential_stiffness=float(rng.uniform(0.3, 0.8)), cutoff=6.0,
                             damping=float(rng.uniform(0.5, 2.0)), growth_rate=0.0)
            q["spring_length_threshold"] = 1e30
            P.append(p), Q.append(q), S.append(np.full(rods, m, np.uint32))
        w.update(cfg=cfg, pos=np.concatenate(P), par=np.concatenate(Q), pot="morse", sim_of=np.concatenate(S),
                 n_sims=n_sims)
    else:
        raise SystemExit(f"unknown config {name}")
    w["vel"] = np.zeros_like(w["pos"])
    w["n_total"] = len(w["pos"]) * world
    return w


class ClockSampler:
    """SM clock and throttle reasons of this rank's GPU during the timed region (B200_PROFILING.md's clocks
    line), polled in-process through NVML every 5 ms; falls back to one `nvidia-smi -lms` child when the NVML
    binding is missing."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, gpu_index: int):
        self.rows = []  # (sm MHz, max MHz, set of reasons)
        self.proc = None
        self.nvml = None
        self.idx = int(gpu_index)
        cvd = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        try:  # NVML and nvidia-smi count physical devices
            self.idx = int(cvd.split(",")[self.idx]) if cvd else self.idx
        except (ValueError, IndexError):
            pass
        self.halt = threading.Event()
        self.t = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nvml
        reasons_fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self.halt.is_set():
            try:
                mask = int(reasons_fn(self.handle))
                self.rows.append((float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)), self.max_mhz,
                                  {n for n, b in self.BITS.items() if mask & b}))
            except Exception:
                pass
            self.halt.wait(0.005)

    def _read(self):
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            try:
                self.rows.append((float(r[1]), float(r[2]),
                                  {n for n, v in zip(self.NAMES, r[5:9]) if v.lower().startswith("active")}))
            except Exception:
                pass

    def stop(self):
        if self.nvml is None and not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no NVML binding and no nvidia-smi"]}
        if self.nvml is not None:
            self.halt.set()
            self.t.join(timeout=1.0)
            source = "nvml, 5 ms"
        else:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
            source = "nvidia-smi -lms 100"
        sm = [r[0] for r in self.rows]
        mx = [r[1] for r in self.rows]
        reasons = set().union(*[r[2] for r in self.rows]) if self.rows else set()
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": source}


def pinned_copy(lib, a):
    b = lib.pinned_empty(a.shape, a.dtype)
    b[...] = a
    return b


def state_buffers(lib, n, nv):
    """Page-locked (pos, vel, params, ids, parents) for Sim.get_state(out=...)."""
    from cr_mech_coli_b200 import _capi
    return (lib.pinned_empty((n, nv, 3), np.float32), lib.pinned_empty((n, nv, 3), np.float32),
            lib.pinned_empty(n, _capi.ROD_PARAMS_DTYPE), lib.pinned_empty(n, _capi.IDENT_DTYPE),
            lib.pinned_empty(n, _capi.IDENT_DTYPE))


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text()), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


def ncu_traffic():
    """dram bytes per launch of the force kernel from the committed ncu capture (or None)."""
    p = ROOT / "profiles" / "force_kernel_traffic.json"
    if p.exists():
        try:
            return json.loads(p.read_text()).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


# ----------------------------------------------------------------------------------------------------
# CPU arm: the oracle (OpenMP over rods = the reference's rayon-over-subdomains strategy) with the
# reference backend's pair schedule (each unordered pair visited once)

def cpu_sample(name: str, n_sample: int):
    """A bounded sample of the workload `name` for the CPU arm: same generator, same density, fewer rods."""
    if name == "c3":
        cfg, pos, par = workload(n_sample)
        return cfg, pos, par, f"{n_sample} rods of the same lattice"
    if name == "c4":
        w = make_workload("c4", n_sample, 0, 1)
        return w["cfg"], w["pos"], w["par"], f"{len(w['pos'])} rods of the same three-layer colony"
    if name == "c5":
        w = make_workload("c5", 64, 0, C5_SIMS // max(1, n_sample // 64))
        return w["cfg"], w["pos"], w["par"], f"{w['n_sims']} of the {C5_SIMS} sims (one colony, rods of different sims far apart)"
    # c2: a random dense patch with the colony's mechanics (the grown state lives on the GPU)
    from cr_mech_coli_b200 import _capi
    from cr_mech_coli_b200.testing import morse_params, random_colony
    L = float(np.sqrt(n_sample * 21.0 * 6.0 / 0.6))
    cfg = _capi.make_config(domain_size=(L, L), domain_height=2.5, n_voxels=(max(1, int(L // 40)),) * 2, dt=0.05,
                            t_max=1e6, n_saves=0)
    pos = random_colony(n_sample, L, seed=5, z_spread=0.0)
    pos[:, :, 2] = 0.0
    par = morse_params(n_sample, growth_rate=0.0, damping=0.02, spring_tension=10.0, rigidity=20.0,
                       potential_stiffness=1.0, strength=0.03)
    par["spring_length_threshold"] = 1e30
    return cfg, pos, par, f"{n_sample} rods at the colony's packing density"


def cpu_baseline(name: str, n_sample: int, steps: int, warmup: int = 1, threads: int | None = None):
    from oracle import oracle

    cfg, pos, par, what = cpu_sample(name, n_sample)
    if name == "c5":  # sims side by side in one domain: shift them apart so that they never meet
        n_s = len(pos) // 64
        side = int(np.ceil(np.sqrt(n_s)))
        for m in range(n_s):
            pos[64 * m:64 * (m + 1), :, 0] += np.float32(260.0 * (m % side))
            pos[64 * m:64 * (m + 1), :, 1] += np.float32(260.0 * (m // side))
        from cr_mech_coli_b200 import _capi
        cfg = _capi.make_config(domain_size=(260.0 * side, 260.0 * side), n_voxels=(side, side), dt=0.002, t_max=1e6,
                                n_saves=0)
    lib = oracle.library(openmp=True)
    # all host cores unless told otherwise (torchrun exports OMP_NUM_THREADS=1 to its workers)
    lib.fn("set_num_threads")(threads or os.cpu_count() or 1)
    cores = lib.fn("num_threads")()
    with oracle.sim(cfg, openmp=True) as o:
        lib.check(lib.fn("sim_set_pair_once")(o.h, 1))
        o.set_agents(pos, np.zeros_like(pos), par)
        o.step(max(1, warmup))
        t0 = time.perf_counter()
        o.step(steps)
        dt = time.perf_counter() - t0
    n = len(pos)
    return {"value": n * steps / dt, "unit": UNIT, "cores": int(cores), "kind": "port",
            "sample": f"{what} x {steps} steps ({dt:.1f} s); pair-once schedule (each unordered pair: "
                      f"calculate_force_between(A,B) and (B,A) once, like the reference backend)",
            "ms_per_step": dt / steps * 1e3, "sample_rods": n}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = args.gpus
    name = args.config
    rods = args.rods or DEFAULT_RODS[name]
    n_sample = args.ref_rods or {"c3": 100_000, "c4": 30_000, "c2": 10_000, "c5": 64 * 64}[name]
    cb = cpu_baseline(name, n_sample, max(1, args.steps), max(1, min(args.warmup, 3)))
    per_gpu = rods_per_gpu(name, rods, world)
    total = per_gpu * world
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "strong" if name == "c5" else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(name, world, per_gpu, total, "strong" if name == "c5" else "weak"),
            "sample_rods": cb["sample_rods"],
            "cpu_baseline": {"value": cb["value"], "unit": UNIT, "cores": cb["cores"], "kind": "port",
                             "sample": cb["sample"]},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference = CPU restatement of the cellular_raza path on all host cores (the Rust reference cannot be "
                    "built here: no cargo, un-vendored dependency); every step is a bounded sample of the workload "
                    "(sample_rods rods at the workload's density: agent-steps/s at fixed density is size-independent, "
                    "ms_per_step is the sample's)"}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------
# GPU arm

class Runner:
    """One rank's simulation: a plain Sim (one GPU, replicas) or a DistributedShard (voxel slabs)."""

    def __init__(self, lib, device, w, world, rank, pinned=False):
        from cr_mech_coli_b200 import _capi, sharded
        self.lib, self.w, self.world = lib, w, world
        self.slabs = world > 1 and bool(w["edges"])
        pos, vel, par = w["pos"], w["vel"], w["par"]
        if pinned:
            pos, vel, par = pinned_copy(lib, pos), pinned_copy(lib, vel), pinned_copy(lib, par)
        self.host = (pos, vel, par)
        if self.slabs:
            self.ds = sharded.DistributedShard(lib, w["cfg"], sharded.SlabPartition(w["edges"]), device)
            self.sim = None
        else:
            self.ds = None
            self.sim = _capi.Sim(lib, w["cfg"], device)

    def upload(self):
        pos, vel, par = self.host
        if self.slabs:
            gsl = np.arange(self.w["g0"], self.w["g0"] + len(pos), dtype=np.uint64)
            self.ds.set_agents_local(pos, vel, par, gsl, n_total=self.w["n_total"])
            self.sim = self.ds.shard.sim
        elif self.w["sim_of"] is not None:
            self.sim.set_agents_batched(pos, vel, par, self.w["sim_of"], self.w["n_sims"])
        else:
            self.sim.set_agents(pos, vel, par)

    def step(self, k):
        if self.slabs:
            self.ds.step(k)
        else:
            self.sim.step(k)

    def count_pairs(self):
        if self.slabs:
            self.ds._halo_ahead = False
            self.ds.rebuild_next = True
            self.ds._exchange_halos()
        f, nn = self.sim.compute_forces()
        return self.sim.read_timers(), f, nn

    def close(self):
        (self.ds or self.sim).close()


def grow_colony(sim, target):
    """C2 warm-up: the colony divides its way from 16 rods to `target` (untimed)."""
    steps = 0
    while sim.n_rods < target and steps < 400000:
        sim.step(100)
        steps += 100
    return steps


def parity_check(run, lib, world, rank, dist, torch):
    """Outside the timed region: after the K sharded steps, the forces of the rods in a 300-um window straddling
    rank 0's upper slab edge (owned by ranks 0 and 1, every one of them with ghost partners) against the oracle on
    the gathered window; all ranks: rod count and momentum balance of the whole colony."""
    from cr_mech_coli_b200.testing import force_errors
    w = run.w
    _, f, nn = run.count_pairs()
    pos, vel, par = run.sim.get_state()[:3]
    n_now = torch.tensor([len(pos)], device="cuda", dtype=torch.int64)
    dist.all_reduce(n_now)
    mom = torch.tensor(np.concatenate([f.astype(np.float64).sum(axis=(0, 1)), np.abs(f).astype(np.float64).sum(axis=(0, 1))]),
                       device="cuda", dtype=torch.float64)
    dist.all_reduce(mom)
    mom = mom.cpu().numpy()
    out = {"rods_total_ok": bool(int(n_now.item()) == w["n_total"]),
           "momentum_rel": float(np.abs(mom[:2]).max() / max(mom[3:5].max(), 1e-300))}
    c = pos[:, :, :2].mean(axis=1)
    xe, y0, margin = float(w["edges"][0]), 14.0, 30.0
    sel = (c[:, 0] > xe - 150 - margin) & (c[:, 0] < xe + 150 + margin) & (c[:, 1] > y0 - margin) & (c[:, 1] < y0 + 300 + margin)
    mine = (pos[sel], vel[sel], par[sel], f[sel], nn[sel]) if rank < 2 else None
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        from oracle import oracle
        parts = [g for g in gathered if g is not None and len(g[0])]
        P, V, Q, F, NN = (np.concatenate([p[k] for p in parts]) for k in range(5))
        cw = P[:, :, :2].mean(axis=1)
        inner = (cw[:, 0] > xe - 150) & (cw[:, 0] < xe + 150) & (cw[:, 1] > y0) & (cw[:, 1] < y0 + 300)
        with oracle.sim(w["cfg"], openmp=True) as o:
            o.set_agents(P, V, Q)
            fo, no = o.compute_forces()
            fabs = np.zeros_like(fo)
            o.L.check(o.L.fn("sim_force_abs_sum")(o.h, fabs.ctypes.data))
        e = force_errors(F[inner], fo[inner], fabs[inner])
        out.update(window_rods=int(inner.sum()), window_rods_per_rank=[int(len(p[0])) for p in parts],
                   window_force_rel_err=e["colony"], window_force_vertex_bound_used=e["vertex"],
                   window_neighbours_equal=bool((NN[inner] == no[inner]).all()))
    return out


def timed_steps(run, lib, K, barrier):
    """K steps on the device clock (CUDA events on the sim's stream); max over ranks is taken by the caller."""
    if barrier:
        barrier()
    msd = C.c_double(0)
    lib.check(lib.fn("sim_timer_start")(run.sim.h))
    run.step(K)
    lib.check(lib.fn("sim_timer_stop")(run.sim.h, C.byref(msd)))
    if barrier:
        barrier()
    return float(msd.value)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c3", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--rods", type=int, default=0, help="rods per GPU (c3, c4), target rods (c2), rods per sim (c5)")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--ref-rods", type=int, default=0, help="rods of the reference arm's sample (default per config)")
    ap.add_argument("--cpu-rods", type=int, default=0)
    ap.add_argument("--cpu-steps", type=int, default=4)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-long-run", action="store_true")
    ap.add_argument("--relax", type=int, default=400, help="c4: untimed relaxation steps before the second timing")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--e2e-reps", type=int, default=3)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    from cr_mech_coli_b200 import _capi

    name = args.config
    rods = args.rods or DEFAULT_RODS[name]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    lib = _capi.product()
    ndev = lib.fn("device_count")()
    if ndev <= 0:
        raise SystemExit("bench.py needs a CUDA device: " + lib.last_error())
    device = local_rank % ndev
    W = max(args.warmup, 3)
    K = args.steps

    # non-tensor peaks of this device (not in MEASURED_PEAKS.json)
    fp32, mufu, mhz = C.c_double(0), C.c_double(0), C.c_double(0)
    lib.check(lib.fn("microbench")(device, C.byref(fp32), C.byref(mufu), C.byref(mhz)))

    dist = torch = None
    barrier = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(device)
        dist.init_process_group("nccl")

        def barrier():
            dist.barrier()
            torch.cuda.synchronize()

    def reduce_max(vals):
        if world == 1:
            return list(vals)
        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def reduce_sum(vals):
        if world == 1:
            return list(vals)
        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        return [float(x) for x in t]

    # ---- the timed region: colony resident in HBM, K steps -------------------------------------------
    w = make_workload(name, rods, rank, world)
    run = Runner(lib, device, w, world, rank)
    run.upload()
    grow_steps = grow_colony(run.sim, w["grow_to"]) if w["grow_to"] else 0
    pairs, _, _ = run.count_pairs()  # instrumented launch: interacting ordered pairs (outside the timed region)
    halo = {"ghost_rods_rank0": int(sum(run.ds.last_counts)), "record_bytes": run.ds.shard.rb} if run.slabs else None
    run.step(W)
    run.sim.enable_timers(True)
    launches0 = run.sim.launch_count()
    rods0 = run.sim.rod_steps
    stats0 = run.sim.list_stats()
    sampler = ClockSampler(device)
    sampler.start()
    ms = timed_steps(run, lib, K, barrier)
    clocks = sampler.stop()
    launches = run.sim.launch_count() - launches0
    rod_steps = run.sim.rod_steps - rods0
    tm = run.sim.read_timers()
    st = run.sim.step_stats()
    lists = run.sim.list_stats()
    totals0 = run.sim.list_totals()
    lists = {"builds": lists["builds"] - stats0["builds"], "eval_steps": lists["eval_steps"] - stats0["eval_steps"],
             "batches": st["batches"], "note": "builds / eval_steps inside the timed region; the rest since the upload",
             "overflows": lists.get("overflows"), "overflow_reasons": lists.get("overflow_reasons"),
             "longest_list": lists.get("longest_list"),
             "rebuild_reasons_since_upload": lists.get("rebuild_reasons"), "structure_totals": totals0}
    n = run.sim.n_rods
    fused_steps = run.sim.fused_steps()
    # the same steps as two launches (k_force_eval + k_integrate), a few of them outside the timed region: what
    # the force part and the mechanics part of the fused kernel cost on their own
    two_launch = None
    if fused_steps and not run.slabs:
        run.sim.set_fused(False)
        run.sim.enable_timers(True)
        run.step(8)
        t2, s2 = run.sim.read_timers(), run.sim.step_stats()
        run.sim.set_fused(True)
        if t2["force_launches"] and s2["integrate_launches"]:
            two_launch = {"force_ms": t2["force_ms"] / t2["force_launches"],
                          "integrate_ms": s2["integrate_ms"] / s2["integrate_launches"]}
    run.sim.enable_timers(False)

    # ---- C3: 100 warm-up + 1000 timed steps, rebuild steps inside the region, and the cost of one
    long_run = None
    if name == "c3" and not args.no_long_run:
        run.step(100 - W - K if 100 > W + K else 0)
        b0 = run.sim.list_stats()
        sampler_l = ClockSampler(device)
        sampler_l.start()
        ms_long = reduce_max([timed_steps(run, lib, 1000, barrier)])[0]
        clocks_long = sampler_l.stop()
        b1 = run.sim.list_stats()
        pairs_long = run.count_pairs()[0] if not run.slabs else None  # (invalidates the structure, like the line below)
        rebuild_ms = None
        if not run.slabs:
            run.sim.compute_forces()  # invalidates the structure: the next step searches and rebuilds it
            rebuild_ms = run.sim.step_timed(1)
        # list steps at the end of the run (right behind the rebuild above / wherever the slabs are), to tell the
        # cost of the rebuild cycles from a drift of the list step itself
        run.step(2)
        b2 = run.sim.list_stats()
        ms_late = reduce_max([timed_steps(run, lib, 20, barrier)])[0]
        b3 = run.sim.list_stats()
        totals_late = run.sim.list_totals()
        diffr = {q: b1["rebuild_reasons"][q] - b0["rebuild_reasons"][q] for q in b1.get("rebuild_reasons", {})} if "rebuild_reasons" in b0 else None
        long_run = {"steps": 1000, "warmup": 100, "ms_per_step": ms_long / 1000,
                    "value": n * world * 1000 / (ms_long * 1e-3),
                    "builds": b1["builds"] - b0["builds"], "eval_steps": b1["eval_steps"] - b0["eval_steps"],
                    "rebuild_step_ms": rebuild_ms, "clocks": clocks_long, "rebuild_reasons": diffr,
                    "list_steps_at_end": {"ms_per_step": ms_late / 20, "builds_inside": b3["builds"] - b2["builds"],
                                          "longest_list": b3.get("longest_list"), "structure_totals": totals_late,
                                          "note": "the relaxing lattice queues more evaluations per pair as rods settle "
                                                  "side by side: the list step's cost follows structure_totals, not the pair count"},
                    "ordered_pairs_in_cutoff_at_end": pairs_long["pairs_in_cutoff"] if pairs_long else None}

    # ---- C4: the stacked-lattice start state is far from mechanical equilibrium (damping 0.02: a vertex moves
    # beyond skin/2 in nearly every step, so nearly every step searches).  The same K steps timed again after
    # `--relax` untimed steps, with the reasons for the searches, show what a settled colony costs.
    relaxed = None
    if name == "c4" and args.relax > 0 and not args.no_long_run:
        r0 = run.sim.list_stats()
        try:
            run.step(args.relax)
            r1 = run.sim.list_stats()
            ms_r = reduce_max([timed_steps(run, lib, 5 * K, barrier)])[0]
            r2_ = run.sim.list_stats()
            diff = lambda a, b, k: {q: b[k][q] - a[k][q] for q in b[k]} if k in b else None
            relaxed = {"relax_steps": args.relax, "steps": 5 * K, "ms_per_step": ms_r / (5 * K),
                       "value": run.sim.n_rods * world * 5 * K / (ms_r * 1e-3), "rods_rank0": int(run.sim.n_rods),
                       "builds": r2_["builds"] - r1["builds"], "eval_steps": r2_["eval_steps"] - r1["eval_steps"],
                       "rebuild_reasons": diff(r1, r2_, "rebuild_reasons"),
                       "rebuild_reasons_while_relaxing": diff(r0, r1, "rebuild_reasons")}
        except _capi.SimulationError as e:  # (the model's own verdict on the synthetic colony: reported, not fatal)
            relaxed = {"relax_steps": args.relax, "error": str(e)[:160], "iteration": int(run.sim.iteration)}

    check = parity_check(run, lib, world, rank, dist, torch) if run.slabs else None
    run.close()

    # ---- strong scaling: ONE colony of `rods` rods over the N slabs (BASELINE configs[2]) --------------
    strong = None
    if name == "c3" and world > 1 and not args.no_strong:
        ws = make_workload(name, rods, rank, world, strong=True)
        rs = Runner(lib, device, ws, world, rank)
        try:
            rs.upload()
            rs.step(W)
            ms_s = reduce_max([timed_steps(rs, lib, K, barrier)])[0]
            ms_s2 = reduce_max([timed_steps(rs, lib, 200, barrier)])[0]
            n_s = reduce_sum([rs.sim.n_rods])[0]
            strong = {"rods_total": int(n_s), "rods_per_gpu": int(rs.sim.n_rods), "ms_per_step": ms_s / K,
                      "value": n_s * K / (ms_s * 1e-3), "ms_per_step_200": ms_s2 / 200,
                      "value_200": n_s * 200 / (ms_s2 * 1e-3), "scaling": "strong"}
        except Exception as e:  # (slabs narrower than two halo widths at large N: reported, not fatal)
            strong = {"error": f"{type(e).__name__}: {e}"[:200]}
        rs.close()

    # ---- e2e: host buffers in, K steps, host buffers out, through the C ABI (page-locked host memory) ---
    reps = []
    for _ in range(max(1, args.e2e_reps)):
        w2 = make_workload(name, rods, rank, world) if w["grow_to"] else w
        if barrier:
            barrier()
        t0 = time.perf_counter()
        r2 = Runner(lib, device, w2, world, rank, pinned=False)
        t1 = time.perf_counter()
        r2.host = tuple(pinned_copy(lib, a) for a in r2.host) if not reps else pinned_host
        pinned_host = r2.host
        out = state_buffers(lib, max(len(w2["pos"]), 1) if not w["grow_to"] else 2 * w["grow_to"] + 4096,
                            w2["pos"].shape[1]) if not reps else out
        t1 = time.perf_counter()
        r2.upload()
        t2 = time.perf_counter()
        if w["grow_to"]:
            grow_colony(r2.sim, w["grow_to"])
        tg = time.perf_counter()
        r2.step(K)
        if run.slabs:
            torch.cuda.synchronize()
        t3 = time.perf_counter()
        pos_out = r2.sim.get_state(out=out)[0]
        t4 = time.perf_counter()
        assert np.isfinite(pos_out).all()
        n2 = r2.sim.n_rods
        r2.close()
        # (the creation of the sim and, for c2, the growth phase are outside: inputs = the colony to be stepped)
        reps.append({"upload_s": t2 - t1, "steps_s": t3 - tg, "download_s": t4 - t3,
                     "total_s": (t2 - t1) + (t3 - tg) + (t4 - t3), "rods": n2})
    tot = reduce_max([r["total_s"] for r in reps])
    for r, t in zip(reps, tot):
        r["total_s_max_over_ranks"] = t
    med = sorted(reps, key=lambda r: r["total_s_max_over_ranks"])[len(reps) // 2]
    nv = w["pos"].shape[1]
    h2d = len(w["pos"]) * (2 * nv * 12 + 100)
    d2h = med["rods"] * (2 * nv * 12 + 100 + 48) + K * 64

    ms = reduce_max([ms])[0]
    total_rods, total_rod_steps = reduce_sum([n, rod_steps])
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return

    peaks, peaks_src = measured_peaks()
    value = total_rod_steps / (ms * 1e-3)
    force_ms = tm["force_ms"] / max(tm["force_launches"], 1)
    integ_ms = st["integrate_ms"] / max(st["integrate_launches"], 1) if st["integrate_launches"] else None
    flops = pairs["pairs_in_cutoff"] * EVALS_PER_ORDERED_PAIR * FLOP_PER_EVAL[w["pot"]]
    achieved = flops / (force_ms * 1e-3) * 1e-12 if force_ms > 0 else 0.0
    mech_ms = ms / K - force_ms
    scaling = "strong" if name == "c5" else "weak"
    e2e_rods = med["rods"] * world
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, world, rods_per_gpu(name, rods, world), rods_per_gpu(name, rods, world) * world, scaling),
        "rods_stepped": {"rank0": int(n), "total": int(total_rods)},
        "math": "MUFU fast path (rsqrt/ex2/lg2) + bit-exact closest-point geometry",
        "e2e": {"value": e2e_rods * K / med["total_s_max_over_ranks"], "unit": UNIT, "h2d_bytes_per_step": h2d / K,
                "d2h_bytes_per_step": d2h / K, "seconds": med["total_s_max_over_ranks"],
                "median_of": len(reps), "breakdown_rank0": {k: med[k] for k in ("upload_s", "steps_s", "download_s")},
                "all_reps_s": tot, "note": "upload of the colony from page-locked host buffers + K steps + download of the "
                                           "state; sim creation (cudaMalloc of the arrays happens inside upload) excluded"},
        "gpu_launches": int(launches),
        "halo": halo,
        "neighbour_lists": lists,
        "clocks": clocks,
        "step_breakdown_ms": {"force": force_ms, "integrate": integ_ms, "fused": bool(fused_steps),
                              "idle_or_other": (ms / K - force_ms - (integ_ms or 0.0)) if (integ_ms is not None or fused_steps) else None,
                              "note": "force / integrate: CUDA events around the first launch of every batch (list steps) or "
                                      "every launch (single steps); the rest of ms_per_step is gaps between kernels, rebuild "
                                      "work, memsets and the status copy"},
        "roofline": {"kernel": ("k_force_eval<.., FUSE>: interaction forces + mechanics of a list step in ONE launch" if fused_steps
                                else "k_force_eval on list steps, k_force<BUILD> on the rebuild steps (event time averaged over both)"),
                     "fused": bool(fused_steps), "bound": "fp32", "achieved": achieved, "peak": fp32.value,
                     "unit": "TFLOP/s", "frac": achieved / fp32.value if fp32.value else None,
                     "traffic": ncu_traffic() if name == "c3" else None,
                     "peak_source": "crm_microbench FFMA loop on this device (FP32 non-tensor peak is not in "
                                    "MEASURED_PEAKS.json); nominal 148*128*2*1.965 GHz = 74.5",
                     "peak_nominal": 74.5, "frac_of_nominal": achieved / 74.5,
                     "mufu_peak_tops": mufu.value, "fma_implied_sm_mhz": mhz.value,
                     "force_ms_per_launch": force_ms, "force_share_of_step": force_ms / (ms / K),
                     "ordered_pairs_in_cutoff": pairs["pairs_in_cutoff"],
                     "candidate_pairs_evaluated": pairs["pairs_tested"],
                     "flop_per_eval": FLOP_PER_EVAL[w["pot"]], "evals_per_ordered_pair": EVALS_PER_ORDERED_PAIR,
                     "note": "achieved = algorithmic FLOP of the interaction forces / duration of the launch; the fused launch also "
                             "carries the mechanics (HBM side: roofline_mech), so its FP32 fraction is below the force-only one "
                             "(two_launch_reference)",
                     "two_launch_reference": (dict(two_launch, frac_fp32=flops / (two_launch["force_ms"] * 1e-3) * 1e-12 / fp32.value)
                                              if two_launch and fp32.value else None)},
        "roofline_mech": ({"kernels": "the same fused launch, read as a memory kernel: SURVEY's 836 algorithmic B/rod-step of the "
                                      "mechanics over the WHOLE launch duration (the force part hides it)", "bound": "hbm",
                           "achieved": n * MECH_BYTES_PER_ROD_STEP / (force_ms * 1e-3) * 1e-9, "peak": peaks["hbm_gbs"],
                           "unit": "GB/s", "peak_source": peaks_src, "ms_per_launch": force_ms,
                           "mechanics_cost_ms": (force_ms - two_launch["force_ms"]) if two_launch else None} if fused_steps else
                          {"kernels": "k_integrate (plus binning on rebuild steps): everything but the force kernel", "bound": "hbm",
                           "achieved": n * MECH_BYTES_PER_ROD_STEP / ((integ_ms or mech_ms) * 1e-3) * 1e-9 if (integ_ms or mech_ms) > 0 else None,
                           "peak": peaks["hbm_gbs"], "unit": "GB/s", "peak_source": peaks_src,
                           "ms_per_launch": integ_ms, "ms_per_step_all_but_force": mech_ms}),
    }
    if line["roofline_mech"]["achieved"]:
        line["roofline_mech"]["frac"] = line["roofline_mech"]["achieved"] / peaks["hbm_gbs"]
    if name == "c5":
        line["ensemble"] = {"sims_total": C5_SIMS, "sims_per_gpu": C5_SIMS // world, "rods_per_sim": rods,
                            "sim_steps_per_s": C5_SIMS * K / (ms * 1e-3)}
    if name == "c2":
        line["growth"] = {"steps_to_target": grow_steps, "rods_at_start_of_timed_region": int(rod_steps // K)}
    if long_run:
        line["long_run"] = long_run
    if relaxed:
        line["relaxed"] = relaxed
    if strong:
        line["strong"] = strong
    if check:
        line["parity_check"] = check
    if world == 1 and not args.no_cpu:
        cb = cpu_baseline(name, args.cpu_rods or {"c3": 50_000, "c4": 20_000, "c2": 5_000, "c5": 64 * 32}[name],
                          args.cpu_steps)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(line))


def _only_json_on_stdout():
    """Libraries write to fd 1 (NCCL prints its version there): everything but the JSON line goes to stderr."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


if __name__ == "__main__":
    _only_json_on_stdout()
    main()
