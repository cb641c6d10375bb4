#!/usr/bin/env python
"""bench.py — agent-steps/s of the colony-mechanics step (BASELINE.json metric).

A "step" is one mechanics step (binning, rod-rod forces, springs/bending/integrator,
growth/division, boundary) over the whole synthetic colony.  Workload at N=1 = BASELINE.json
configs[2]: 2D monolayer, 1e6 rods x 8 vertices, Mie potential (SURVEY.md §8d "C3").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--rods R] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = rods*K / device time with the colony resident in
HBM; `e2e` = the same through the C-ABI with host buffers (upload + K steps + download in the
timed region); `roofline` = the force kernel against the measured FP32 peak; `cpu_baseline` =
the CPU oracle on a bounded sample of the same lattice.
"""

from __future__ import annotations

import argparse
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


def workload(n_rods: int, seed: int = 12345):
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


class ClockSampler:
    """SM clock and throttle reasons of this rank's GPU during the timed region (B200_PROFILING.md's clocks
    line), polled in-process through NVML every 5 ms; falls back to one `nvidia-smi -lms` child when the NVML
    binding is missing.  (A child per rank is slow to start next to a 20-ms timed region, and killing it in the
    middle of its NVML initialisation stalled the following cudaMalloc calls of the job.)"""

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


def cpu_baseline(n_sample: int, steps: int, threads: int | None):
    """The CPU oracle (OpenMP over rods = the reference's rayon-over-subdomains strategy) timed on
    a bounded sample of the same lattice."""
    from oracle import oracle

    cfg, pos, par = workload(n_sample)
    lib = oracle.library(openmp=True)
    # all host cores unless told otherwise (torchrun exports OMP_NUM_THREADS=1 to its workers)
    lib.fn("set_num_threads")(threads or os.cpu_count() or 1)
    cores = lib.fn("num_threads")()
    with oracle.sim(cfg, openmp=True) as o:
        o.set_agents(pos, np.zeros_like(pos), par)
        o.step(1)  # warm-up
        t0 = time.perf_counter()
        o.step(steps)
        dt = time.perf_counter() - t0
    return {"value": n_sample * steps / dt, "unit": UNIT, "cores": int(cores), "kind": "port",
            "sample": f"{n_sample} rods of the same lattice x {steps} steps ({dt:.1f} s)",
            "ms_per_step": dt / steps * 1e3}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_sample = args.ref_rods
    cb = cpu_baseline(n_sample, max(1, args.steps), None)
    cfg_w = {"workload": "C3: 2D monolayer, jittered nematic lattice, Mie n=6 m=3, 8 vertices/rod",
             "rods": args.rods, "dt": 0.01, "sample_rods": n_sample}
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg_w,
            "cpu_baseline": {"value": cb["value"], "unit": UNIT, "cores": cb["cores"], "kind": "port",
                             "sample": cb["sample"]},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference = CPU restatement of the cellular_raza path (the Rust reference cannot be built "
                    "here: no cargo, un-vendored dependency); agent-steps/s at fixed density is size-independent"}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--rods", type=int, default=1_000_000, help="rods per GPU")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--ref-rods", type=int, default=20000)
    ap.add_argument("--cpu-rods", type=int, default=20000)
    ap.add_argument("--cpu-steps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    from cr_mech_coli_b200 import _capi

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
    import ctypes as C

    # non-tensor peaks of this device (not in MEASURED_PEAKS.json)
    fp32, mufu, mhz = C.c_double(0), C.c_double(0), C.c_double(0)
    lib.check(lib.fn("microbench")(device, C.byref(fp32), C.byref(mufu), C.byref(mhz)))

    if world == 1:
        cfg, pos, par = workload(args.rods, seed=12345)
        vel = np.zeros_like(pos)
        n = len(pos)
        sim = _capi.Sim(lib, cfg, device)
        sim.set_agents(pos, vel, par)
        # interacting ordered pairs of this colony (instrumented launch, outside the timed region)
        sim.compute_forces()
        pairs = sim.read_timers()
        sim.step(W)
        sim.enable_timers(True)
        launches0 = sim.launch_count()
        sampler = ClockSampler(device)
        sampler.start()
        ms = sim.step_timed(K)
        clocks = sampler.stop()
        launches = sim.launch_count() - launches0
        tm = sim.read_timers()
        lists = sim.list_stats()
        sim.enable_timers(False)
        sim.close()

        # e2e: host buffers in, K steps, host buffers out, through the C ABI
        # (inputs and outputs live in page-locked host memory, crm_host_alloc)
        pos, vel, par = pinned_copy(lib, pos), pinned_copy(lib, vel), pinned_copy(lib, par)
        out = state_buffers(lib, n, pos.shape[1])
        sim2 = _capi.Sim(lib, cfg, device)
        t0 = time.perf_counter()
        sim2.set_agents(pos, vel, par)
        sim2.step(K)
        pos_out = sim2.get_state(out=out)[0]
        e2e_s = time.perf_counter() - t0
        assert np.isfinite(pos_out).all()
        h2d = pos.nbytes + vel.nbytes + n * (32 + 16 + 16 + 48 + 4)
        d2h = pos_out.nbytes * 2 + n * 20 + K * 40
        sim2.close()
        total_rods = n
        parallelism = "1 GPU"
        halo = None
    else:
        # one slab of ONE colony per GPU: halo exchange over NCCL every step
        import torch
        import torch.distributed as dist

        from cr_mech_coli_b200 import sharded
        from cr_mech_coli_b200.testing import lattice_slab, mie_params

        torch.cuda.set_device(device)
        dist.init_process_group("nccl")
        pos, domain, edges, g0 = lattice_slab(args.rods, rank, world)
        n = len(pos)
        nvx = (max(1, int(domain[0] // 40.0)), max(1, int(domain[1] // 40.0)))
        cfg = _capi.make_config(domain_size=domain, domain_height=2.5, n_voxels=nvx, dt=0.01, t_max=1e6, n_saves=0)
        par = mie_params(n, growth_rate=0.0)
        par["spring_length_threshold"] = 1e30
        vel = np.zeros_like(pos)
        part = sharded.SlabPartition(edges)
        gsl = np.arange(g0, g0 + n, dtype=np.uint64)

        def barrier():
            dist.barrier()
            torch.cuda.synchronize()

        ds = sharded.DistributedShard(lib, cfg, part, device)
        ds.set_agents_local(pos, vel, par, gsl, n_total=n * world)
        ds._exchange_halos()
        ds.shard.sim.compute_forces()
        pairs = ds.shard.sim.read_timers()
        halo = {"ghost_rods_rank0": int(sum(ds.last_counts)), "record_bytes": ds.shard.rb}
        ds.step(W)
        ds.shard.sim.enable_timers(True)
        launches0 = ds.shard.sim.launch_count()
        sampler = ClockSampler(device)
        sampler.start()
        barrier()
        lib.check(lib.fn("sim_timer_start")(ds.shard.sim.h))
        ds.step(K)
        msd = C.c_double(0)
        lib.check(lib.fn("sim_timer_stop")(ds.shard.sim.h, C.byref(msd)))
        barrier()
        ms = float(msd.value)
        clocks = sampler.stop()
        launches = ds.shard.sim.launch_count() - launches0
        tm = ds.shard.sim.read_timers()
        lists = ds.shard.sim.list_stats()
        ds.close()

        # e2e: host buffers in, K sharded steps, host buffers out
        pos, vel, par = pinned_copy(lib, pos), pinned_copy(lib, vel), pinned_copy(lib, par)
        out = state_buffers(lib, n, pos.shape[1])
        ds2 = sharded.DistributedShard(lib, cfg, part, device)
        barrier()
        t0 = time.perf_counter()
        ds2.set_agents_local(pos, vel, par, gsl, n_total=n * world)
        ds2.step(K)
        pos_out = ds2.shard.sim.get_state(out=out)[0]
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        barrier()
        assert np.isfinite(pos_out).all()
        h2d = pos.nbytes + vel.nbytes + n * (32 + 16 + 16 + 48 + 4)
        d2h = pos_out.nbytes * 2 + n * 20 + K * 40
        ds2.close()
        t = torch.tensor([ms, e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = float(t[0]), float(t[1])
        cnt = torch.tensor([n], device="cuda", dtype=torch.int64)
        dist.all_reduce(cnt)
        total_rods = int(cnt.item())
        parallelism = f"{world} voxel slabs along x, one per GPU, per-step NCCL halo exchange"
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return

    peaks, peaks_src = measured_peaks()
    value = total_rods * K / (ms * 1e-3)
    force_ms = tm["force_ms"] / max(tm["force_launches"], 1)
    flops = pairs["pairs_in_cutoff"] * EVALS_PER_ORDERED_PAIR * FLOP_PER_EVAL["mie"]
    achieved = flops / (force_ms * 1e-3) * 1e-12 if force_ms > 0 else 0.0
    mech_ms = ms / K - force_ms
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C3: 2D monolayer, jittered nematic lattice (24x4 um spacing), Mie n=6 m=3, "
                               "r=1.5, cutoff 6, 8 vertices/rod, dt=0.01, growth 0",
                   "rods_per_gpu": n, "rods_total": total_rods,
                   "parallelism": parallelism,
                   "l2": "state (~0.9 GB/GPU) larger than the 126 MB L2; no explicit flush",
                   "math": "MUFU fast path (rsqrt/ex2/lg2) + bit-exact closest-point geometry"},
        "e2e": {"value": total_rods * K / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d / K,
                "d2h_bytes_per_step": d2h / K, "seconds": e2e_s},
        "gpu_launches": int(launches),
        "halo": halo,
        "neighbour_lists": lists,
        "clocks": clocks,
        "roofline": {"kernel": "k_force_eval<8,1> on list steps, k_force<8,1,0,1> on the rebuild steps "
                               "(event time averaged over both)", "bound": "fp32", "achieved": achieved, "peak": fp32.value,
                     "unit": "TFLOP/s", "frac": achieved / fp32.value if fp32.value else None,
                     "traffic": ncu_traffic(),
                     "peak_source": "crm_microbench FFMA loop on this device (FP32 non-tensor peak is not in "
                                    "MEASURED_PEAKS.json); nominal 148*128*2*1.965 GHz = 74.5",
                     "mufu_peak_tops": mufu.value, "fma_implied_sm_mhz": mhz.value,
                     "force_ms_per_launch": force_ms, "force_share_of_step": force_ms / (ms / K),
                     "ordered_pairs_in_cutoff": pairs["pairs_in_cutoff"],
                     "candidate_pairs_evaluated": pairs["pairs_tested"],
                     "flop_per_eval": FLOP_PER_EVAL["mie"], "evals_per_ordered_pair": EVALS_PER_ORDERED_PAIR},
        "roofline_mech": {"kernels": "binning + k_integrate (everything but the force kernel)", "bound": "hbm",
                          "achieved": n * MECH_BYTES_PER_ROD_STEP / (mech_ms * 1e-3) * 1e-9 if mech_ms > 0 else None,
                          "peak": peaks["hbm_gbs"], "unit": "GB/s", "peak_source": peaks_src,
                          "ms_per_step": mech_ms},
    }
    if line["roofline_mech"]["achieved"]:
        line["roofline_mech"]["frac"] = line["roofline_mech"]["achieved"] / peaks["hbm_gbs"]
    if world == 1 and not args.no_cpu:
        cb = cpu_baseline(args.cpu_rods, args.cpu_steps, None)
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
