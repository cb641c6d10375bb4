"""ctypes binding of the C ABI declared in ``include/crm_b200.h``.

The product library ``libcrm_b200.so`` (CUDA kernels + C++ host orchestration)
is loaded from this package directory.  There is **no CPU fallback**: if the
library is missing, or was built without its CUDA objects, importing the
simulation entry points raises.  The same binding class is reused by the tests
to drive the CPU oracle (``oracle/libcrm_oracle.so``, prefix ``oracle_``), which
exports the same signatures.
"""

from __future__ import annotations

import ctypes as C
import os
import weakref
from pathlib import Path

import numpy as np

_PKG_DIR = Path(__file__).resolve().parent
LIB_NAME = "libcrm_b200.so"


class CrmConfig(C.Structure):
    """``crm_config`` (replaces ``Configuration``, reference src/config.rs:395-465)."""

    _fields_ = [
        ("t0", C.c_float),
        ("dt", C.c_float),
        ("t_max", C.c_float),
        ("n_saves", C.c_uint32),
        ("domain_size", C.c_float * 2),
        ("domain_height", C.c_float),
        ("n_voxels", C.c_uint32 * 2),
        ("rng_seed", C.c_uint64),
        ("gel_pressure", C.c_float),
        ("surface_friction", C.c_float),
        ("surface_friction_distance", C.c_float),
        ("n_threads", C.c_uint32),
    ]


# numpy view of ``crm_rod_params`` (reference src/agent.rs:212-239).
ROD_PARAMS_DTYPE = np.dtype(
    [
        ("diffusion_constant", "<f4"),
        ("spring_tension", "<f4"),
        ("rigidity", "<f4"),
        ("spring_length", "<f4"),
        ("damping", "<f4"),
        ("potential", "<i4"),
        ("radius", "<f4"),
        ("strength", "<f4"),
        ("cutoff", "<f4"),
        ("potential_stiffness", "<f4"),
        ("bound", "<f4"),
        ("en", "<f4"),
        ("em", "<f4"),
        ("n_neighbors", "<u4"),
        ("growth_rate", "<f4"),
        ("growth_setter_kind", "<i4"),
        ("growth_setter_a", "<f4"),
        ("growth_setter_b", "<f4"),
        ("spring_length_threshold", "<f4"),
        ("threshold_setter_kind", "<i4"),
        ("threshold_setter_a", "<f4"),
        ("threshold_setter_b", "<f4"),
        ("has_neighbor_reduction", "<i4"),
        ("neighbor_reduction_max", "<u4"),
        ("neighbor_reduction_exp", "<f4"),
    ],
    align=False,
)
assert ROD_PARAMS_DTYPE.itemsize == 100

IDENT_DTYPE = np.dtype([("kind", "<i8"), ("voxel", "<u8"), ("counter", "<u8")])

POT_MORSE, POT_MIE = 0, 1
SETTER_NORMAL, SETTER_EXPLICIT = 0, 1
ID_DIVISION, ID_INITIAL, ID_INSERTED = 0, 1, 2
ERR_BOUNDARY, ERR_NUMERIC, ERR_DIVISION = -3, -4, -5  # crm_status values reported per simulation

STATUS_NAMES = {
    -1: "InvalidArgument",
    -2: "CudaError",
    -3: "BoundaryError",
    -4: "CalcError",
    -5: "DivisionError",
    -6: "TimeError",
    -7: "OutOfMemory",
}


class SimulationError(RuntimeError):
    """Mirror of cellular_raza's ``SimulationError`` raised by
    ``run_simulation_with_agents`` (reference src/simulation.rs:237)."""


def _fptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Library:
    """A loaded implementation of the C ABI (product: prefix ``crm_``)."""

    def __init__(self, path: os.PathLike | str, prefix: str = "crm_"):
        self.path = str(path)
        self.prefix = prefix
        if not os.path.exists(self.path):
            raise ImportError(
                f"{self.path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(cr_mech_coli_b200 has no CPU fallback)"
            )
        self.lib = C.CDLL(self.path, mode=C.RTLD_GLOBAL if prefix == "crm_" else C.RTLD_LOCAL)
        self._declare()

    def fn(self, name):
        return getattr(self.lib, self.prefix + name)

    def has(self, name) -> bool:
        return hasattr(self.lib, self.prefix + name)

    def _declare(self):
        vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
        P = C.POINTER
        sig = {
            "last_error": (C.c_char_p, []),
            "save_schedule": (C.c_int, [P(CrmConfig), vp, u32, P(u64)]),
            "sim_create": (C.c_int, [P(CrmConfig), C.c_int, P(vp)]),
            "sim_destroy": (None, [vp]),
            "sim_set_agents": (C.c_int, [vp, vp, vp, vp, u32, u32]),
            "sim_set_agents_batched": (C.c_int, [vp, vp, vp, vp, vp, u32, u32, u32]),
            "sim_get_sim_index": (C.c_int, [vp, vp]),
            "sim_snapshot_sim_index": (C.c_int, [vp, u32, vp]),
            "sim_run": (C.c_int, [vp]),
            "sim_step": (C.c_int, [vp, u32]),
            "sim_compute_forces": (C.c_int, [vp, vp, vp]),
            "sim_n_rods": (u32, [vp]),
            "sim_n_vertices": (u32, [vp]),
            "sim_iteration": (u64, [vp]),
            "sim_get_state": (C.c_int, [vp, vp, vp, vp, vp, vp]),
            "sim_n_snapshots": (u32, [vp]),
            "sim_snapshot_info": (C.c_int, [vp, u32, P(u64), P(u32)]),
            "sim_snapshot_copy": (C.c_int, [vp, u32, vp, vp, vp, vp, vp]),
            "sim_rod_steps": (u64, [vp]),
            # product only
            "abi_version": (C.c_int, []),
            "device_count": (C.c_int, []),
            "device_mem_info": (C.c_int, [C.c_int, P(C.c_size_t), P(C.c_size_t)]),
            "counter_to_color": (None, [u32, P(C.c_uint8 * 3)]),
            "color_to_counter": (u32, [P(C.c_uint8 * 3)]),
            "generate_positions": (
                C.c_int,
                [u32, C.c_float, P(CrmConfig), u64, P(C.c_float * 2), C.c_float, u32, vp],
            ),
            "sim_launch_count": (u64, [vp]),
            "sim_step_timed": (C.c_int, [vp, u32, P(C.c_double)]),
            "sim_timer_start": (C.c_int, [vp]),
            "sim_timer_stop": (C.c_int, [vp, P(C.c_double)]),
            "sim_set_exact_math": (C.c_int, [vp, C.c_int]),
            "sim_set_skin": (C.c_int, [vp, C.c_float]),
            "sim_set_batching": (C.c_int, [vp, C.c_int]),
            "sim_set_fused": (C.c_int, [vp, C.c_int]),
            "sim_fused_steps": (u64, [vp]),
            "sim_read_step_stats": (C.c_int, [vp, P(C.c_double), P(u64), P(u64), P(u32)]),
            "sim_set_solver_order": (C.c_int, [vp, C.c_int]),
            "sim_set_spring_end_weight": (C.c_int, [vp, C.c_float]),
            "sim_rebuild_reasons": (C.c_int, [vp, C.c_uint64 * 8]),
            "sim_list_totals": (C.c_int, [vp, C.c_uint64 * 6]),
            "sim_list_stats": (C.c_int, [vp, P(u64), P(u64)]),
            "sim_list_overflows": (C.c_int, [vp, P(u32), P(u64), P(u32)]),
            "microbench": (C.c_int, [C.c_int, P(C.c_double), P(C.c_double), P(C.c_double)]),
            "sim_enable_timers": (C.c_int, [vp, C.c_int]),
            "sim_read_timers": (C.c_int, [vp, P(C.c_double), P(u64), P(u64), P(u64)]),
            # sharding (product only)
            "sim_set_slab": (C.c_int, [vp, C.c_float, C.c_float]),
            "sim_set_global_slots": (C.c_int, [vp, vp, u32]),
            "sim_get_global_slots": (C.c_int, [vp, vp]),
            "sim_halo_record_bytes": (u32, [vp]),
            "sim_migrant_record_bytes": (u32, [vp]),
            "sim_max_extent": (C.c_float, [vp]),
            "sim_refresh_bounds": (C.c_int, [vp]),
            "sim_search_radius": (C.c_float, [vp]),
            "sim_halo_pack": (C.c_int, [vp, C.c_int, C.c_float, vp, u32, P(u32)]),
            "sim_set_ghosts": (C.c_int, [vp, vp, u32, vp, u32, C.c_float]),
            "sim_halo_pack_msgs": (C.c_int, [vp, C.c_float, vp, vp, u32, u32, C.c_int, P(u32), P(u32)]),
            "sim_wants_rebuild": (C.c_int, [vp]),
            "sim_keep_lists": (C.c_int, [vp]),
            "sim_skin": (C.c_float, [vp]),
            "sim_pack_leavers": (C.c_int, [vp, vp, u32, P(u32)]),
            "sim_set_ghosts_msgs": (C.c_int, [vp, vp, vp, vp, u32, vp, vp, P(C.c_float)]),
            "sim_isolate_sim_errors": (C.c_int, [vp, C.c_int]),
            "sim_get_sim_status": (C.c_int, [vp, vp]),
            "sim_run_costs": (C.c_int, [vp, vp, u32, u32, vp, vp, vp]),
            "sim_patch_count": (C.c_uint64, [vp]),
            "sim_last_forces": (C.c_int, [vp, vp]),
            "sim_set_patching": (C.c_int, [vp, C.c_int]),
            "sim_stream": (vp, [vp]),
            "sim_set_stream": (C.c_int, [vp, vp]),
            "sim_halo_repack": (C.c_int, [vp, vp, vp]),
            "sim_ghosts_refresh": (C.c_int, [vp, vp, vp]),
            "sim_step_begin": (C.c_int, [vp, vp]),
            "sim_step_end": (C.c_int, [vp]),
            "sim_migrate_out": (C.c_int, [vp, vp, u32, P(u32), vp, u32, P(u32)]),
            "sim_migrate_in": (C.c_int, [vp, vp, u32]),
            "sim_n_leaving": (u32, [vp]),
            "sim_pending_divisions": (u32, [vp, vp, vp, u32]),
            "sim_apply_divisions": (C.c_int, [vp, vp, vp, vp, u32]),
            "sim_take_snapshot": (C.c_int, [vp]),
            "sim_snapshot_global_slots": (C.c_int, [vp, u32, vp]),
            "polygon_hulls": (C.c_int, [vp, vp, u32, u32, C.c_float, C.c_float, vp, vp, u32]),
            "render_mask_2d": (C.c_int, [vp, vp, vp, u32, u32, P(C.c_float * 2), P(u32 * 2), C.c_float, C.c_float, vp,
                                         vp, P(C.c_double)]),
            "overlap": (C.c_int, [vp, vp, u32, u32, C.c_float, C.c_float, P(C.c_float)]),
            "parents_diff_mask": (C.c_int, [vp, vp, u64, vp, u32, vp, u32, C.c_float, vp, P(C.c_int64)]),
            "overlap_noalloc": (C.c_int, [vp, vp, u32, u32, C.c_float, C.c_float, vp, vp, vp]),
            "comm_unique_id": (C.c_int, [vp]),
            "comm_create": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, P(vp)]),
            "comm_destroy": (None, [vp]),
            "sim_set_comm": (C.c_int, [vp, vp]),
            "sim_steps_sharded": (C.c_int, [vp, u32, vp, vp, vp, vp, P(u32), vp]),
            "host_alloc": (vp, [C.c_size_t]),
            "host_free": (None, [vp]),
            # oracle only
            "sim_pairs_in_cutoff": (u64, [vp]),
            "num_threads": (C.c_int, []),
            "set_num_threads": (None, [C.c_int]),
            "set_mie_textbook": (None, [C.c_int]),
            "sim_set_pair_once": (C.c_int, [vp, C.c_int]),
            "sim_force_abs_sum": (C.c_int, [vp, vp]),
        }
        for name, (res, args) in sig.items():
            if self.has(name):
                f = self.fn(name)
                f.restype = res
                f.argtypes = args

    def last_error(self) -> str:
        msg = self.fn("last_error")()
        return msg.decode("utf-8", "replace") if msg else ""

    def check(self, rc: int):
        if rc < 0:
            kind = STATUS_NAMES.get(rc, f"error {rc}")
            raise SimulationError(f"{kind}: {self.last_error()}")
        return rc

    def pinned_empty(self, shape, dtype) -> np.ndarray:
        """Uninitialised array in page-locked host memory (``crm_host_alloc``): buffers for
        ``Sim.set_agents`` / ``Sim.get_state(out=...)`` that cross the bus at link speed."""
        dtype = np.dtype(dtype)
        shape = (shape,) if np.isscalar(shape) else tuple(shape)
        nbytes = int(np.prod(shape, dtype=np.int64)) * dtype.itemsize
        ptr = self.fn("host_alloc")(nbytes)
        if not ptr:
            raise MemoryError(self.last_error())
        buf = (C.c_char * max(nbytes, 1)).from_address(ptr)
        arr = np.frombuffer(buf, dtype=np.uint8, count=nbytes).view(dtype).reshape(shape)
        weakref.finalize(buf, self.fn("host_free"), ptr)  # freed when the last view of the block dies
        return arr

    def save_schedule(self, cfg: CrmConfig):
        out = np.zeros(int(cfg.n_saves) + 2, dtype=np.uint64)
        n_it = C.c_uint64(0)
        n = self.check(self.fn("save_schedule")(C.byref(cfg), _fptr(out), out.size, C.byref(n_it)))
        return out[:n].copy(), int(n_it.value)


class Sim:
    """Thin RAII wrapper around ``crm_sim`` (or the oracle's equivalent)."""

    def __init__(self, lib: Library, cfg: CrmConfig, device: int = 0):
        self.L = lib
        self.h = C.c_void_p()
        self.cfg = cfg
        lib.check(lib.fn("sim_create")(C.byref(cfg), int(device), C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.L.fn("sim_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- state ---------------------------------------------------------------
    def set_agents(self, pos, vel, params):
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        vel = np.ascontiguousarray(vel, dtype=np.float32)
        params = np.ascontiguousarray(params, dtype=ROD_PARAMS_DTYPE)
        if pos.ndim != 3 or pos.shape[2] != 3 or vel.shape != pos.shape or params.shape[0] != pos.shape[0]:
            raise ValueError("pos/vel must be (n_rods, n_vertices, 3) and params (n_rods,)")
        n, nv = pos.shape[0], pos.shape[1]
        self.L.check(self.L.fn("sim_set_agents")(self.h, _fptr(pos), _fptr(vel), _fptr(params), n, nv))

    def set_agents_batched(self, pos, vel, params, sim_of, n_sims: int):
        """Independent simulations in one device state: ``sim_of[a]`` = simulation of rod a."""
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        vel = np.ascontiguousarray(vel, dtype=np.float32)
        params = np.ascontiguousarray(params, dtype=ROD_PARAMS_DTYPE)
        sim_of = np.ascontiguousarray(sim_of, dtype=np.uint32)
        if pos.ndim != 3 or pos.shape[2] != 3 or vel.shape != pos.shape or params.shape[0] != pos.shape[0] or \
                sim_of.shape[0] != pos.shape[0]:
            raise ValueError("pos/vel must be (n_rods, n_vertices, 3), params and sim_of (n_rods,)")
        self.L.check(self.L.fn("sim_set_agents_batched")(self.h, _fptr(pos), _fptr(vel), _fptr(params), _fptr(sim_of),
                                                         pos.shape[0], pos.shape[1], int(n_sims)))

    def snapshots_by_sim(self, n_sims: int):
        """Per simulation: list of (iteration, pos, vel, params, ids, parents), rods in slot order."""
        out = [[] for _ in range(n_sims)]
        for k, (it, pos, vel, par, ids, parents) in enumerate(self.snapshots()):
            so = np.zeros(len(pos), dtype=np.uint32)
            self.L.check(self.L.fn("sim_snapshot_sim_index")(self.h, k, _fptr(so)))
            order = np.argsort(so, kind="stable")
            cuts = np.searchsorted(so[order], np.arange(n_sims + 1))
            for m in range(n_sims):
                sel = order[cuts[m]:cuts[m + 1]]
                out[m].append((it, pos[sel], vel[sel], par[sel], ids[sel], parents[sel]))
        return out

    @property
    def n_rods(self) -> int:
        return int(self.L.fn("sim_n_rods")(self.h))

    @property
    def n_vertices(self) -> int:
        return int(self.L.fn("sim_n_vertices")(self.h))

    @property
    def iteration(self) -> int:
        return int(self.L.fn("sim_iteration")(self.h))

    @property
    def rod_steps(self) -> int:
        return int(self.L.fn("sim_rod_steps")(self.h))

    def run(self):
        self.L.check(self.L.fn("sim_run")(self.h))

    def step(self, n_steps: int = 1):
        self.L.check(self.L.fn("sim_step")(self.h, int(n_steps)))

    def step_timed(self, n_steps: int) -> float:
        """Advance n_steps; returns the device time in ms (CUDA events on the sim's stream)."""
        ms = C.c_double(0)
        self.L.check(self.L.fn("sim_step_timed")(self.h, int(n_steps), C.byref(ms)))
        return float(ms.value)

    def set_exact_math(self, exact: bool):
        self.L.check(self.L.fn("sim_set_exact_math")(self.h, int(exact)))

    def set_solver_order(self, stored_increments: int):
        """History depth of the Adams-Bashforth solver: 0 Euler, 1 AB2, 2 AB3 (before set_agents)."""
        self.L.check(self.L.fn("sim_set_solver_order")(self.h, int(stored_increments)))

    def set_spring_end_weight(self, w: float):
        self.L.check(self.L.fn("sim_set_spring_end_weight")(self.h, float(w)))

    def set_fused(self, on: bool):
        """List steps as one kernel (forces + mechanics, default) or as two launches."""
        self.L.check(self.L.fn("sim_set_fused")(self.h, int(on)))

    def fused_steps(self) -> int:
        return int(self.L.fn("sim_fused_steps")(self.h))

    def set_batching(self, on: bool):
        """List steps in batches with a device-side latch (default) or one host round trip per step."""
        self.L.check(self.L.fn("sim_set_batching")(self.h, int(on)))

    def step_stats(self):
        ms, nl, nb, bl = C.c_double(0), C.c_uint64(0), C.c_uint64(0), C.c_uint32(0)
        self.L.check(self.L.fn("sim_read_step_stats")(self.h, C.byref(ms), C.byref(nl), C.byref(nb), C.byref(bl)))
        return {"integrate_ms": float(ms.value), "integrate_launches": int(nl.value), "batches": int(nb.value),
                "batch_len": int(bl.value)}

    def set_skin(self, skin: float):
        self.L.check(self.L.fn("sim_set_skin")(self.h, float(skin)))

    def list_stats(self):
        b, e = C.c_uint64(0), C.c_uint64(0)
        self.L.check(self.L.fn("sim_list_stats")(self.h, C.byref(b), C.byref(e)))
        out = {"builds": int(b.value), "eval_steps": int(e.value)}
        if self.L.has("sim_list_overflows"):
            r, c, m = C.c_uint32(0), C.c_uint64(0), C.c_uint32(0)
            self.L.fn("sim_list_overflows")(self.h, C.byref(r), C.byref(c), C.byref(m))
            out.update(overflows=int(c.value), overflow_reasons=int(r.value), longest_list=int(m.value))
        if self.L.has("sim_rebuild_reasons"):
            w = (C.c_uint64 * 8)()
            self.L.fn("sim_rebuild_reasons")(self.h, w)
            out["rebuild_reasons"] = dict(zip(("displacement", "left_slab", "list_overflow", "patch_overflow",
                                               "unpatched_divisions", "no_lists", "grid_redo", "steps"), map(int, w)))
        return out

    def list_totals(self):
        """Queued evaluations and listed rods of the live neighbour structure (zeros if there is none)."""
        w = (C.c_uint64 * 6)()
        self.L.check(self.L.fn("sim_list_totals")(self.h, w))
        return dict(zip(("own_evaluations", "ext_evaluations", "sensing_candidates", "listed_rods", "long_rows",
                         "two_launch_steps"), map(int, w)))

    def last_forces(self):
        f = np.zeros((self.n_rods, self.n_vertices, 3), dtype=np.float32)
        self.L.check(self.L.fn("sim_last_forces")(self.h, _fptr(f)))
        return f

    def patch_count(self) -> int:
        """Division steps absorbed by patching the live neighbour structure (no new search)."""
        return int(self.L.fn("sim_patch_count")(self.h))

    def compute_forces(self):
        n, nv = self.n_rods, self.n_vertices
        force = np.zeros((n, nv, 3), dtype=np.float32)
        nn = np.zeros(n, dtype=np.uint32)
        self.L.check(self.L.fn("sim_compute_forces")(self.h, _fptr(force), _fptr(nn)))
        return force, nn

    def get_state(self, out=None):
        """(pos, vel, params, ids, parents) in slot order.  ``out`` = the same five arrays to fill
        (e.g. page-locked ones from ``Library.pinned_empty``), sized for at least ``n_rods``."""
        n, nv = self.n_rods, self.n_vertices
        if out is not None:
            want = [((n, nv, 3), np.float32), ((n, nv, 3), np.float32), ((n,), ROD_PARAMS_DTYPE), ((n,), IDENT_DTYPE),
                    ((n,), IDENT_DTYPE)]
            got = []
            for a, (shape, dt) in zip(out, want):
                if a.dtype != dt or not a.flags.c_contiguous or a.shape[1:] != shape[1:] or a.shape[0] < n:
                    raise ValueError("out arrays must be C-contiguous, typed like get_state()'s and hold n_rods rows")
                got.append(a[:n])
            pos, vel, par, ids, parents = got
        else:
            pos = np.empty((n, nv, 3), dtype=np.float32)
            vel = np.empty((n, nv, 3), dtype=np.float32)
            par = np.empty(n, dtype=ROD_PARAMS_DTYPE)
            ids = np.empty(n, dtype=IDENT_DTYPE)
            parents = np.empty(n, dtype=IDENT_DTYPE)
        self.L.check(
            self.L.fn("sim_get_state")(self.h, _fptr(pos), _fptr(vel), _fptr(par), _fptr(ids), _fptr(parents))
        )
        return pos, vel, par, ids, parents

    def snapshots(self):
        """List of (iteration, pos, vel, params, ids, parents), one per save point."""
        out = []
        nv = self.n_vertices
        for k in range(int(self.L.fn("sim_n_snapshots")(self.h))):
            it, n = C.c_uint64(0), C.c_uint32(0)
            self.L.check(self.L.fn("sim_snapshot_info")(self.h, k, C.byref(it), C.byref(n)))
            n = int(n.value)
            pos = np.zeros((n, nv, 3), dtype=np.float32)
            vel = np.zeros((n, nv, 3), dtype=np.float32)
            par = np.zeros(n, dtype=ROD_PARAMS_DTYPE)
            ids = np.zeros(n, dtype=IDENT_DTYPE)
            parents = np.zeros(n, dtype=IDENT_DTYPE)
            self.L.check(
                self.L.fn("sim_snapshot_copy")(
                    self.h, k, _fptr(pos), _fptr(vel), _fptr(par), _fptr(ids), _fptr(parents)
                )
            )
            out.append((int(it.value), pos, vel, par, ids, parents))
        return out

    # -- product-only instrumentation -----------------------------------------
    def launch_count(self) -> int:
        return int(self.L.fn("sim_launch_count")(self.h))

    def enable_timers(self, enable: bool = True):
        self.L.check(self.L.fn("sim_enable_timers")(self.h, int(enable)))

    def read_timers(self):
        ms, nl, pc, pt = C.c_double(0), C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        self.L.check(self.L.fn("sim_read_timers")(self.h, C.byref(ms), C.byref(nl), C.byref(pc), C.byref(pt)))
        return {
            "force_ms": float(ms.value),
            "force_launches": int(nl.value),
            "pairs_in_cutoff": int(pc.value),
            "pairs_tested": int(pt.value),
        }


_PRODUCT: Library | None = None


def product() -> Library:
    """The CUDA library.  Raises ImportError when it has not been built."""
    global _PRODUCT
    if _PRODUCT is None:
        _PRODUCT = Library(_PKG_DIR / LIB_NAME, "crm_")
        if _PRODUCT.fn("abi_version")() != 3:
            raise ImportError("libcrm_b200.so: ABI version mismatch")
    return _PRODUCT


def make_config(**kw) -> CrmConfig:
    c = CrmConfig()
    c.t0, c.dt, c.t_max, c.n_saves = 0.0, 0.1, 100.0, 10
    c.domain_size[0] = c.domain_size[1] = 100.0
    c.domain_height = 2.5
    c.n_voxels[0] = c.n_voxels[1] = 1
    c.rng_seed = 0
    c.gel_pressure = 0.0
    c.surface_friction = 0.0
    c.surface_friction_distance = 1.0
    c.n_threads = 1
    for k, v in kw.items():
        if k in ("domain_size", "n_voxels"):
            arr = getattr(c, k)
            arr[0], arr[1] = v[0], v[1]
        else:
            setattr(c, k, v)
    return c
