"""Entry points of the drop-in boundary (reference src/simulation.rs): ``generate_positions``,
``generate_agents``, ``run_simulation_with_agents``, ``sort_cellular_identifiers``.  All compute goes
through the C ABI of ``libcrm_b200.so``; there is no CPU fallback."""

from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi, storage
from .cell_container import CellContainer, _Snapshot
from .datatypes import AgentSettings, Configuration, PhysicalInteraction, RodAgent, StorageOption

SimulationError = _capi.SimulationError


def generate_positions(n_agents, agent_settings: AgentSettings, config: Configuration, rng_seed=0, dx=(0.0, 0.0),
                       randomize_positions=0.0, n_vertices=8):
    """reference src/simulation.rs:17-51 -> ``_generate_positions`` (:107-166), ChaCha8-stream exact."""
    lib = _capi.product()
    out = np.zeros((int(n_agents), int(n_vertices), 3), dtype=np.float32)
    cfg = config._to_c()
    dxa = (C.c_float * 2)(float(dx[0]), float(dx[1]))
    rc = lib.fn("generate_positions")(int(n_agents), float(agent_settings.mechanics.spring_length), C.byref(cfg),
                                      int(rng_seed), C.byref(dxa), float(randomize_positions), int(n_vertices),
                                      out.ctypes.data_as(C.c_void_p))
    if rc != 0:
        # rand panics on an empty sampling range (rods do not fit into their chunk)
        raise ValueError("cannot sample empty range: the domain is too small for the requested agents")
    return [out[k].copy() for k in range(int(n_agents))]


def generate_agents(n_agents, agent_settings: AgentSettings, config: Configuration, rng_seed=0, dx=(0.0, 0.0),
                    randomize_positions=0.0, n_vertices=8):
    """reference src/simulation.rs:56-104"""
    positions = generate_positions(n_agents, agent_settings, config, rng_seed, dx, randomize_positions, n_vertices)
    m = agent_settings.mechanics
    vel0 = m.vel
    agents = []
    for pos in positions:
        vel = vel0 if vel0.shape == pos.shape else np.zeros_like(pos)
        agents.append(RodAgent(
            pos, vel, PhysicalInteraction(agent_settings.interaction), m.diffusion_constant, m.spring_tension,
            m.rigidity, m.spring_length, m.damping, agent_settings.growth_rate, agent_settings.growth_rate_setter,
            agent_settings.spring_length_threshold, agent_settings.spring_length_threshold_setter,
            agent_settings.neighbor_reduction))
    return agents


def pack_agents(agents):
    """list[RodAgent] -> (pos, vel, params) arrays of the C ABI."""
    n = len(agents)
    nv = agents[0]._pos.shape[0] if n else 8
    pos = np.zeros((n, nv, 3), np.float32)
    vel = np.zeros((n, nv, 3), np.float32)
    par = np.zeros(n, _capi.ROD_PARAMS_DTYPE)
    for k, a in enumerate(agents):
        if a._pos.shape[0] != nv:
            raise ValueError("all agents must have the same number of vertices")
        pos[k], vel[k] = a._pos, a._vel
        a._to_record(par[k])
    return pos, vel, par


def run_simulation_with_agents(config: Configuration, agents, device: int = 0) -> CellContainer:
    """reference src/simulation.rs:232-319.  Agents are copied in, the whole run executes on the GPU
    (``crm_sim_run``), snapshots come back at the save points as packed arrays."""
    storage.check_options(config.storage_options)
    agents = list(agents)
    lib = _capi.product()
    pos, vel, par = pack_agents(agents)
    with _capi.Sim(lib, config._to_c(), device) as sim:
        if len(agents):
            sim.set_agents(pos, vel, par)
            sim.run()
            snaps = [_Snapshot(*s) for s in sim.snapshots()]
        else:
            its, _ = lib.save_schedule(sim.cfg)
            e = np.zeros((0, 8, 3), np.float32)
            snaps = [_Snapshot(int(i), e, e, np.zeros(0, _capi.ROD_PARAMS_DTYPE), np.zeros(0, _capi.IDENT_DTYPE),
                               np.zeros(0, _capi.IDENT_DTYPE)) for i in its]
    # StorageOption.SerdeJson: every save point also goes to disk; the container remembers where
    # (reference src/simulation.rs:221-230,307-316)
    path = None
    if StorageOption.SerdeJson in config.storage_options:
        path = storage.new_run_directory(config)
        for s in snaps:
            storage.write_snapshot(path, s.iteration, s.pos, s.vel, s.par, s.ids, s.parents)
    return CellContainer(path=path, _snapshots=snaps)


def params_from_settings(agent_settings: AgentSettings, n: int) -> np.ndarray:
    """``n`` parameter records (``crm_rod_params``) filled from one ``AgentSettings`` template: the array form of
    ``[RodAgent(pos=p, vel=v, **agent_settings.to_rod_agent_dict()) for ...]`` without building the objects."""
    m = agent_settings.mechanics
    template = RodAgent(np.zeros((2, 3), np.float32), np.zeros((2, 3), np.float32),
                        PhysicalInteraction(agent_settings.interaction), m.diffusion_constant, m.spring_tension, m.rigidity,
                        m.spring_length, m.damping, agent_settings.growth_rate, agent_settings.growth_rate_setter,
                        agent_settings.spring_length_threshold, agent_settings.spring_length_threshold_setter,
                        agent_settings.neighbor_reduction)
    rec = np.zeros(1, _capi.ROD_PARAMS_DTYPE)
    template._to_record(rec[0])
    return np.repeat(rec, int(n))


def run_simulation_with_arrays(config: Configuration, pos, vel, params, device: int = 0) -> CellContainer:
    """``run_simulation_with_agents`` for colonies too large to hold as Python objects (extension; no reference
    counterpart): positions / velocities ``(n_rods, n_vertices, 3)`` f32 and one ``crm_rod_params`` record per rod
    (``params_from_settings``, or ``_capi.ROD_PARAMS_DTYPE`` filled by hand) in, the same ``CellContainer`` out.
    The container keeps its snapshots packed, so a 1e6-rod run never materialises a million ``RodAgent`` objects
    unless a caller asks for them (``positions_at_iteration`` returns the arrays)."""
    storage.check_options(config.storage_options)
    lib = _capi.product()
    pos = np.ascontiguousarray(pos, dtype=np.float32)
    vel = np.ascontiguousarray(vel, dtype=np.float32)
    params = np.ascontiguousarray(params, dtype=_capi.ROD_PARAMS_DTYPE)
    with _capi.Sim(lib, config._to_c(), device) as sim:
        sim.set_agents(pos, vel, params)
        sim.run()
        snaps = [_Snapshot(*s) for s in sim.snapshots()]
    path = None
    if StorageOption.SerdeJson in config.storage_options:
        path = storage.new_run_directory(config)
        for s in snaps:
            storage.write_snapshot(path, s.iteration, s.pos, s.vel, s.par, s.ids, s.parents)
    return CellContainer(path=path, _snapshots=snaps)


def run_simulations_batched(config: Configuration, agent_lists, device: int = 0):
    """Many independent ``run_simulation_with_agents(config, agents)`` calls in ONE device state
    (parameter-fitting ensembles: reference src/crm_fit/predict.rs:231-287 run once per parameter
    point by src/crm_fit/optimize.rs:75-95).  Returns one ``CellContainer`` per agent list, equal to
    what the individual calls return (deterministic settings)."""
    storage.check_options(config.storage_options)
    agent_lists = [list(a) for a in agent_lists]
    n_sims = len(agent_lists)
    flat = [a for lst in agent_lists for a in lst]
    if not flat:
        return [run_simulation_with_agents(config, [], device) for _ in agent_lists]
    sim_of = np.concatenate([np.full(len(lst), m, np.uint32) for m, lst in enumerate(agent_lists)])
    lib = _capi.product()
    pos, vel, par = pack_agents(flat)
    with _capi.Sim(lib, config._to_c(), device) as sim:
        sim.set_agents_batched(pos, vel, par, sim_of, n_sims)
        sim.run()
        per_sim = sim.snapshots_by_sim(n_sims)
    return [CellContainer(path=None, _snapshots=[_Snapshot(*s) for s in snaps]) for snaps in per_sim]


def sort_cellular_identifiers(identifiers):
    """reference src/simulation.rs:332-339"""
    return sorted(identifiers)
