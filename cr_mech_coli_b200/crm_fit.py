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


# ---------------------------------------------------------------------------------------------
# Settings of a fit and the map parameter vector -> agents (reference src/crm_fit/settings.rs,
# src/crm_fit/predict.rs:34-229): the caller side of the ensemble path above.
# ---------------------------------------------------------------------------------------------


class SampledFloat:
    """A parameter the optimiser samples in [min, max] (reference src/crm_fit/settings.rs:13-42);
    ``individual`` = one value per rod instead of one for the colony."""

    def __init__(self, min, max, initial, individual=False):
        self.min, self.max, self.initial, self.individual = float(min), float(max), float(initial), individual

    def __repr__(self):
        return f"SampledFloat(min={self.min}, max={self.max}, initial={self.initial}, individual={self.individual})"

    def __eq__(self, o):
        return isinstance(o, SampledFloat) and vars(self) == vars(o)


def _parameter(v):
    """``Parameter`` (settings.rs:45-59, untagged): a SampledFloat table, a float, or a per-rod list."""
    if isinstance(v, SampledFloat):
        return v
    if isinstance(v, dict):
        unknown = set(v) - {"min", "max", "initial", "individual"}
        if unknown or not {"min", "max", "initial"} <= set(v):
            raise ValueError(f"not a SampledFloat table: {v}")
        return SampledFloat(v["min"], v["max"], v["initial"], v.get("individual", None))
    if isinstance(v, (list, tuple, np.ndarray)):
        return [float(x) for x in v]
    return float(v)


class Morse:
    """settings.rs:224-230"""

    def __init__(self, potential_stiffness):
        self.potential_stiffness = _parameter(potential_stiffness)


class Mie:
    """settings.rs:233-243"""

    def __init__(self, en, em, bound):
        self.en, self.em, self.bound = _parameter(en), _parameter(em), float(bound)


class PotentialType:
    """Mie | Morse (settings.rs:248-306)."""

    def __init__(self, inner):
        if not isinstance(inner, (Mie, Morse)):
            raise TypeError("PotentialType holds a Mie or a Morse")
        self.inner = inner

    Mie = staticmethod(lambda mie: PotentialType(mie))
    Morse = staticmethod(lambda morse: PotentialType(morse))

    def to_short_string(self):
        return "mie" if isinstance(self.inner, Mie) else "morse"

    def clone_inner(self):
        import copy
        return copy.deepcopy(self.inner)


class Parameters:
    """settings.rs:100-118; field order = order in the parameter vector."""

    FIELDS = ("radius", "rigidity", "spring_tension", "damping", "strength", "growth_rate")

    def __init__(self, radius, rigidity, spring_tension, damping, strength, potential_type, growth_rate):
        self.radius, self.rigidity, self.spring_tension = _parameter(radius), _parameter(rigidity), _parameter(spring_tension)
        self.damping, self.strength, self.growth_rate = _parameter(damping), _parameter(strength), _parameter(growth_rate)
        self.potential_type = potential_type


class Constants:
    """settings.rs:433-466 (defaults :468-478)."""

    def __init__(self, t_max, dt, domain_size, rng_seed, cutoff, n_vertices, displacement_error, n_voxels=(1, 1),
                 n_saves=0, error_cost=1e9):
        self.t_max, self.dt = float(t_max), float(dt)
        self.domain_size = [float(domain_size[0]), float(domain_size[1])]
        self.n_voxels = [int(n_voxels[0]), int(n_voxels[1])]
        self.rng_seed, self.cutoff, self.n_vertices = int(rng_seed), float(cutoff), int(n_vertices)
        self.n_saves, self.displacement_error, self.error_cost = int(n_saves), float(displacement_error), float(error_cost)
        if self.n_vertices < 1 or min(self.n_voxels) < 1:
            raise ValueError("n_vertices and n_voxels must be non-zero")


class OptimizationInfos:
    """settings.rs:502-517"""

    def __init__(self):
        self.bounds_lower, self.bounds_upper, self.initial_values = [], [], []
        self.parameter_infos, self.constants, self.constant_infos = [], [], []


class Settings:
    """settings.rs:520-536.  ``optimization`` / ``others`` are kept as the parsed tables (the optimiser
    drivers are out of scope)."""

    def __init__(self, constants, parameters, optimization=None, others=None):
        self.constants, self.parameters, self.optimization, self.others = constants, parameters, optimization, others

    @staticmethod
    def from_toml_string(toml_string: str) -> "Settings":
        import tomllib
        try:
            t = tomllib.loads(toml_string)
            p = dict(t["parameters"])
            pt = p.pop("potential_type")
            if set(pt) == {"Mie"}:
                pot = PotentialType(Mie(**pt["Mie"]))
            elif set(pt) == {"Morse"}:
                pot = PotentialType(Morse(**pt["Morse"]))
            else:
                raise ValueError("parameters.potential_type must hold exactly one of Mie, Morse")
            return Settings(Constants(**t["constants"]), Parameters(potential_type=pot, **p), t.get("optimization"),
                            t.get("others"))
        except (tomllib.TOMLDecodeError, KeyError, TypeError) as e:  # reference: toml error -> ValueError
            raise ValueError(f"invalid settings: {e}") from e

    @staticmethod
    def from_toml(toml_filename) -> "Settings":
        with open(toml_filename, "r") as f:
            return Settings.from_toml_string(f.read())

    @property
    def domain_height(self) -> float:
        return 2.5  # settings.rs:583-586

    def to_config(self):
        """settings.rs:654-698"""
        from .datatypes import Configuration
        c = self.constants
        progressbar = (self.others or {}).get("progressbar") if isinstance(self.others, dict) else None
        return Configuration(domain_height=self.domain_height, n_threads=1, t0=0.0, dt=c.dt, t_max=c.t_max,
                             n_saves=c.n_saves, progressbar=progressbar, domain_size=list(c.domain_size),
                             n_voxels=list(c.n_voxels), rng_seed=c.rng_seed, gel_pressure=0.0, surface_friction=0.0,
                             surface_friction_distance=1.0)

    def _slots(self):
        p = self.parameters
        names = [("radius", "Radius", "µm", "r"), ("rigidity", "Rigidity", "µm/min", "κ"),
                 ("spring_tension", "Spring Tension", "1/min²", "γ"), ("damping", "Damping", "1/min", "λ"),
                 ("strength", "Strength", "µm^2/min^2", "V$_0$"), ("growth_rate", "Growth Rate", "1/min", "µ")]
        out = [(getattr(p, f), n, u, s) for f, n, u, s in names]
        inner = p.potential_type.inner
        if isinstance(inner, Mie):
            out += [(inner.en, "Exponent n", "1", "n"), (inner.em, "Exponent m", "1", "m")]
        else:
            out += [(inner.potential_stiffness, "Potential Stiffness", "µm", "λ")]
        return out

    def generate_optimization_infos(self, n_agents: int) -> OptimizationInfos:
        """settings.rs:701-799"""
        o = OptimizationInfos()
        for v, name, units, sym in self._slots():
            if isinstance(v, SampledFloat):
                if not v.individual:
                    o.bounds_lower.append(v.min); o.bounds_upper.append(v.max); o.initial_values.append(v.initial)
                    o.parameter_infos.append((name, units, sym))
                else:
                    o.bounds_lower += [v.min] * n_agents; o.bounds_upper += [v.max] * n_agents
                    o.initial_values += [v.initial] * n_agents
                    o.parameter_infos += [(f"{name} {i}", units, sym) for i in range(n_agents)]
            elif isinstance(v, list):
                o.constants += v
                o.constant_infos.append((name, units, sym))
            else:
                o.constants.append(v)
                o.constant_infos.append((name, units, sym))
        return o


def _expand_parameters(parameters, n_agents: int, settings_parameters: Parameters):
    """Per-rod values of every model parameter for a population of parameter vectors
    (``parameters[M, D]``), in the reference's consumption order (predict.rs:58-151).  Returns a dict
    name -> f32[M, n_agents]."""
    P = np.atleast_2d(np.asarray(parameters, dtype=np.float32))
    M = P.shape[0]
    counter = 0

    def take(v):
        nonlocal counter
        if isinstance(v, SampledFloat):
            if v.individual is True:
                if counter + n_agents > P.shape[1]:
                    raise IndexError("parameter vector too short")
                res = P[:, counter:counter + n_agents]
                counter += n_agents
                return res
            if counter >= P.shape[1]:
                raise IndexError("parameter vector too short")
            res = np.repeat(P[:, counter:counter + 1], n_agents, axis=1)
            counter += 1
            return res
        if isinstance(v, list):
            if len(v) != n_agents:
                raise IndexError("per-rod parameter list does not match the number of rods")
            return np.broadcast_to(np.asarray(v, np.float32), (M, n_agents))
        return np.full((M, n_agents), v, np.float32)

    sp = settings_parameters
    out = {f: take(getattr(sp, f)) for f in Parameters.FIELDS}
    inner = sp.potential_type.inner
    if isinstance(inner, Mie):
        out["en"], out["em"] = take(inner.en), take(inner.em)
        out["bound"] = np.full((M, n_agents), inner.bound, np.float32)
    else:
        out["potential_stiffness"] = take(inner.potential_stiffness)
    return out


def define_initial_agents_packed(parameters, initial_positions, domain_height: float, parameter_defs: Parameters,
                                 constants: Constants):
    """Array form of ``define_initial_agents`` (predict.rs:34-200) for a POPULATION of parameter vectors
    ``parameters[M, D]``: returns (pos[M*n, nv, 3], vel, crm_rod_params[M*n], sim_of[M*n]).  Rods get z =
    domain_height/2 when the data are 2D, spring_length = mean xy segment length, no division
    (threshold inf), growth setter Normal(rate, 0)."""
    from .testing import base_params

    ip = np.asarray(initial_positions, dtype=np.float32)
    if ip.ndim != 3:
        raise ValueError("initial_positions must be (n_agents, n_vertices, 2 or 3)")
    n, nv = ip.shape[:2]
    if nv != constants.n_vertices:
        raise ValueError("initial_positions and constants.n_vertices disagree")
    if ip.shape[2] != 3:
        pos1 = np.zeros((n, nv, 3), np.float32)
        pos1[:, :, :2] = ip[:, :, :2]
        pos1[:, :, 2] += np.float32(domain_height / 2.0)
    else:
        pos1 = ip.copy()
    vals = _expand_parameters(parameters, n, parameter_defs)
    M = vals["radius"].shape[0]
    seg = np.zeros(n, np.float32)
    for i in range(nv - 1):  # sequential f32 sum like the reference's loop
        dx = pos1[:, i + 1, 0] - pos1[:, i, 0]
        dy = pos1[:, i + 1, 1] - pos1[:, i, 1]
        seg = seg + np.sqrt(dx * dx + dy * dy, dtype=np.float32)
    spring_length = seg / np.float32(nv - 1)
    par = base_params(M * n)
    is_mie = "en" in vals
    par["potential"] = _capi.POT_MIE if is_mie else _capi.POT_MORSE
    for k in ("radius", "strength", "rigidity", "spring_tension", "damping", "growth_rate") + (
            ("en", "em", "bound") if is_mie else ("potential_stiffness",)):
        par[k] = np.ascontiguousarray(vals[k]).reshape(-1)
    par["cutoff"] = constants.cutoff
    par["diffusion_constant"] = 0.0
    par["spring_length"] = np.tile(spring_length, M)
    par["growth_setter_kind"] = _capi.SETTER_NORMAL
    par["growth_setter_a"] = par["growth_rate"]
    par["growth_setter_b"] = 0.0
    par["spring_length_threshold"] = np.inf
    par["threshold_setter_kind"] = _capi.SETTER_EXPLICIT
    par["threshold_setter_a"] = np.inf
    par["threshold_setter_b"] = np.inf
    pos = np.tile(pos1, (M, 1, 1))
    sim_of = np.repeat(np.arange(M, dtype=np.uint32), n)
    return pos, np.zeros_like(pos), par, sim_of


def define_initial_agents(parameters, initial_positions, domain_height: float, parameter_defs: Parameters,
                          constants: Constants):
    """``define_initial_agents`` (predict.rs:34-200) -> list[RodAgent] for ONE parameter vector."""
    from .datatypes import MiePotentialF32, MorsePotentialF32, RodAgent

    pos, vel, par, _ = define_initial_agents_packed(np.asarray(parameters, np.float32)[None, :], initial_positions,
                                                    domain_height, parameter_defs, constants)
    agents = []
    for a in range(len(pos)):
        q = par[a]
        if q["potential"] == _capi.POT_MIE:
            inter = MiePotentialF32(float(q["radius"]), float(q["strength"]), float(q["bound"]), float(q["cutoff"]),
                                    float(q["en"]), float(q["em"]))
        else:
            inter = MorsePotentialF32(float(q["radius"]), float(q["potential_stiffness"]), float(q["cutoff"]),
                                      float(q["strength"]))
        agents.append(RodAgent(pos[a], vel[a], inter, diffusion_constant=0.0, spring_tension=float(q["spring_tension"]),
                               rigidity=float(q["rigidity"]), spring_length=float(q["spring_length"]),
                               damping=float(q["damping"]), growth_rate=float(q["growth_rate"]),
                               growth_rate_setter={"mean": float(q["growth_rate"]), "std": 0.0},
                               spring_length_threshold=float("inf"),
                               spring_length_threshold_setter={"l1": float("inf"), "l2": float("inf")},
                               neighbor_reduction=None))
    return agents


def run_simulation(parameters, initial_positions, settings: Settings):
    """``crm_fit.run_simulation`` (predict.rs:10-30) -> CellContainer."""
    from .simulation import run_simulation_with_agents
    agents = define_initial_agents(parameters, initial_positions, settings.domain_height, settings.parameters,
                                   settings.constants)
    return run_simulation_with_agents(settings.to_config(), agents)


def _is_normal(x) -> np.ndarray:
    x = np.asarray(x, np.float32)
    return np.isfinite(x) & (np.abs(x) >= np.finfo(np.float32).tiny)


def predict_calculate_cost_population(parameters, positions_all, iterations, settings: Settings, device: int = 0,
                                      return_status: bool = False):
    """Costs of a whole population ``parameters[M, D]`` in ONE batched device run (the unit scipy's
    differential evolution evaluates per generation with ``updating='deferred'``; the reference fans
    ``predict_calculate_cost`` out over processes, src/crm_fit/optimize.rs:75-95,235-252).  Per member:
    simulation error or non-normal cost -> ``constants.error_cost`` (predict.rs:224-228,249-256).  A member
    whose cells divide also gets ``error_cost`` here (population extension: the other members' costs stay
    usable); the reference's single-vector call raises for it, and so does ``predict_calculate_cost`` below.
    ``return_status=True`` adds the per-member status (0 | ERR_BOUNDARY | ERR_NUMERIC | ERR_DIVISION)."""
    positions_all = np.asarray(positions_all, dtype=np.float32)
    if positions_all.ndim != 4:
        raise ValueError("positions_all must be (n_results, n_agents, n_vertices, 2 or 3)")
    c = settings.constants
    pos, vel, par, sim_of = define_initial_agents_packed(parameters, positions_all[0], settings.domain_height,
                                                         settings.parameters, c)
    M = int(sim_of[-1]) + 1 if len(sim_of) else 0
    costs, status = costs_on_device(_capi.product(), settings.to_config()._to_c(), pos, vel, par, sim_of, M,
                                    positions_all, iterations, device, error_cost=c.error_cost)
    costs[~_is_normal(costs)] = c.error_cost
    return (costs, status) if return_status else costs


def predict_calculate_cost(parameters, positions_all, iterations, settings: Settings, device: int = 0) -> float:
    """``crm_fit.predict_calculate_cost`` (predict.rs:202-229) for one parameter vector.  A division raises
    IndexError like the reference (predict.rs:271-280 propagates it through ``?``); BoundaryError / CalcError and
    non-normal costs give ``constants.error_cost``."""
    costs, status = predict_calculate_cost_population(np.asarray(parameters, np.float32)[None, :], positions_all,
                                                      iterations, settings, device, return_status=True)
    if int(status[0]) == _capi.ERR_DIVISION:
        raise IndexError("Cell division is currently not supported by this optimization algorithm")
    return float(costs[0])
