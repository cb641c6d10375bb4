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
                            error_cost: float = ERROR_COST_DEFAULT) -> np.ndarray:
    """Costs of a population: ``agent_lists[m]`` = the agents of parameter point m (what
    ``define_initial_agents`` builds, reference src/crm_fit/predict.rs:34-200).  One batched run on
    the device; a simulation error (BoundaryError, CalcError) gives every member ``error_cost`` like
    the reference's per-member NaN -> error_cost rule, because the members share one device state."""
    agent_lists = [list(a) for a in agent_lists]
    n_sims = len(agent_lists)
    flat = [a for lst in agent_lists for a in lst]
    sim_of = np.concatenate([np.full(len(lst), m, np.uint32) for m, lst in enumerate(agent_lists)])
    pos, vel, par = pack_agents(flat)
    lib = _capi.product()
    try:
        with _capi.Sim(lib, config._to_c(), device) as sim:
            sim.set_agents_batched(pos, vel, par, sim_of, n_sims)
            sim.run()
            per_sim = sim.snapshots_by_sim(n_sims)
    except _capi.SimulationError:
        return np.full(n_sims, error_cost, dtype=np.float32)
    return costs_from_snapshots(per_sim, positions_all, iterations, error_cost)
