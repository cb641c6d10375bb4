"""Ensemble cost evaluation for parameter fitting: the unit the reference batches is
``predict_calculate_cost_rs`` (reference src/crm_fit/predict.rs:231-287) — run one simulation, sum the
squared xy distance of every initial rod's vertices to the data at the chosen save iterations —
fanned out per parameter point by rayon / scipy workers (src/crm_fit/optimize.rs:75-95,235-252).
Here a whole population runs as ONE batched device state (``crm_sim_set_agents_batched``) and the
costs are reduced from the packed snapshots."""

from __future__ import annotations

import numpy as np

from . import _capi
from .simulation import pack_agents

ERROR_COST_DEFAULT = 1e10


def costs_from_snapshots(per_sim, positions_all, iterations, error_cost: float = ERROR_COST_DEFAULT) -> np.ndarray:
    """``per_sim[m]`` = snapshots of simulation m as (iteration, pos, vel, par, ids, parents);
    ``positions_all[n_result, n_agent, n_vertices, 2+]`` = data; ``iterations`` = indices into the
    sorted save iterations (reference src/crm_fit/predict.rs:258-285).  Non-finite cost ->
    ``error_cost`` (:224-228); a divided cell raises like the reference's IndexError."""
    positions_all = np.asarray(positions_all, dtype=np.float32)
    costs = np.zeros(len(per_sim), dtype=np.float32)
    for m, snaps in enumerate(per_sim):
        snaps = sorted(snaps, key=lambda s: s[0])
        total = np.float32(0.0)
        for n_result, data_index in enumerate(iterations):
            _, pos, _, _, ids, _ = snaps[data_index]
            if np.any(ids["kind"] != _capi.ID_INITIAL):
                raise IndexError("Cell division is currently not supported by this optimization algorithm")
            n_agent = ids["counter"].astype(np.int64)
            diff = pos[:, :, :2] - positions_all[n_result, n_agent, :, :2]
            total = np.float32(total + np.float32((diff.astype(np.float32) ** 2).sum(dtype=np.float32)))
        costs[m] = total if np.isfinite(total) else error_cost
    return costs


def predict_calculate_costs(config, agent_lists, positions_all, iterations, device: int = 0,
                            error_cost: float = ERROR_COST_DEFAULT, on_device: bool = True) -> np.ndarray:
    """Costs of a population: ``agent_lists[m]`` = the agents of parameter point m (what
    ``define_initial_agents`` builds, reference src/crm_fit/predict.rs:34-200).  One batched run on
    the device.  ``on_device`` (default): no snapshots, the costs are reduced by a kernel at the result
    iterations, and an error (BoundaryError, CalcError) or a division costs only the member it happens
    in ``error_cost`` — the reference's per-member rule (:224-228,249-256).  ``on_device=False`` keeps
    the snapshot path (every member gets ``error_cost`` if any of them fails)."""
    agent_lists = [list(a) for a in agent_lists]
    n_sims = len(agent_lists)
    flat = [a for lst in agent_lists for a in lst]
    sim_of = np.concatenate([np.full(len(lst), m, np.uint32) for m, lst in enumerate(agent_lists)])
    pos, vel, par = pack_agents(flat)
    lib = _capi.product()
    if on_device:
        return costs_on_device(lib, config._to_c(), pos, vel, par, sim_of, n_sims, positions_all, iterations, device,
                               error_cost)[0]
    try:
        with _capi.Sim(lib, config._to_c(), device) as sim:
            sim.set_agents_batched(pos, vel, par, sim_of, n_sims)
            sim.run()
            per_sim = sim.snapshots_by_sim(n_sims)
    except _capi.SimulationError:
        return np.full(n_sims, error_cost, dtype=np.float32)
    return costs_from_snapshots(per_sim, positions_all, iterations, error_cost)


def costs_on_device(lib, cfg, pos, vel, par, sim_of, n_sims: int, positions_all, iterations, device: int = 0,
                    error_cost: float = ERROR_COST_DEFAULT):
    """Packed-array form of ``predict_calculate_costs``: returns (costs, status per simulation).
    ``iterations`` index the sorted save iterations of ``cfg`` like the reference's."""
    import ctypes as C

    positions_all = np.ascontiguousarray(np.asarray(positions_all, dtype=np.float32)[..., :2])
    n_results, n_agents, nv = positions_all.shape[:3]
    if nv != pos.shape[1]:
        raise ValueError("positions_all and the agents disagree on n_vertices")
    its, _ = lib.save_schedule(cfg)
    its = np.sort(np.asarray(its, dtype=np.uint64))
    idx = np.asarray(list(iterations), dtype=np.int64)
    if len(idx) != n_results:
        raise ValueError("one iteration index per entry of positions_all")
    result_its = np.ascontiguousarray(its[idx])
    costs = np.zeros(n_sims, dtype=np.float32)
    status = np.zeros(n_sims, dtype=np.int32)
    with _capi.Sim(lib, cfg, device) as sim:
        sim.set_agents_batched(pos, vel, par, sim_of, n_sims)
        lib.check(lib.fn("sim_run_costs")(sim.h, positions_all.ctypes.data_as(C.c_void_p), n_results, n_agents,
                                          result_its.ctypes.data_as(C.c_void_p), costs.ctypes.data_as(C.c_void_p),
                                          status.ctypes.data_as(C.c_void_p)))
    bad = (status != 0) | ~np.isfinite(costs)
    costs[bad] = error_cost
    return costs, status
