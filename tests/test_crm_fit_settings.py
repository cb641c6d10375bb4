"""crm_fit caller side: Settings (reference src/crm_fit/settings.rs), the parameter-vector -> agents map
(src/crm_fit/predict.rs:34-200; the reference tests this surface in tests/test_fit.py,
tests/test_crm_fit_settings.py) and, on the GPU, the population cost against stand-alone oracle runs."""
import numpy as np
import pytest

from cr_mech_coli_b200 import _capi, crm_fit

TOML = """
[constants]
t_max = 1.0
dt = 0.01
domain_size = [120.0, 90.0]
rng_seed = 0
cutoff = 12.0
n_vertices = 6
displacement_error = 0.3
n_saves = 3

[parameters]
radius = { min = 0.5, max = 2.0, initial = 1.0, individual = false }
rigidity = 8.0
spring_tension = 3.0
damping = { min = 0.0, max = 5.0, initial = 1.0 }
strength = { min = 0.0, max = 0.3, initial = 0.05 }
growth_rate = { min = 0.0, max = 0.08, initial = 0.016, individual = true }

[parameters.potential_type.Mie]
en = { min = 0.5, max = 10.0, initial = 4.0 }
em = 2.0
bound = 50.0

[optimization.differential_evolution]
seed = 0
"""


def _rods(n, nv=6, seed=0):
    rng = np.random.default_rng(seed)
    c = np.stack([20 + 16.0 * (np.arange(n) % 5), 20 + 9.0 * (np.arange(n) // 5)], 1)
    ang = rng.normal(0, 0.2, n)
    t = (np.arange(nv) - (nv - 1) / 2) * 2.0
    p = np.zeros((n, nv, 2), np.float32)
    p[:, :, 0] = c[:, :1] + t * np.cos(ang)[:, None]
    p[:, :, 1] = c[:, 1:] + t * np.sin(ang)[:, None]
    return p


def test_settings_from_toml_and_config():
    s = crm_fit.Settings.from_toml_string(TOML)
    assert s.constants.n_voxels == [1, 1] and s.constants.error_cost == 1e9 and s.constants.n_saves == 3
    assert s.parameters.radius == crm_fit.SampledFloat(0.5, 2.0, 1.0, False)
    assert s.parameters.damping.individual is None and s.parameters.rigidity == 8.0
    assert s.parameters.potential_type.to_short_string() == "mie"
    cfg = s.to_config()
    assert (cfg.dt, cfg.t_max, cfg.n_saves, cfg.domain_height) == (np.float32(0.01), 1.0, 3, 2.5)
    assert list(cfg.domain_size) == [120.0, 90.0] and cfg.gel_pressure == 0.0
    with pytest.raises(ValueError):
        crm_fit.Settings.from_toml_string(TOML.replace("cutoff = 12.0", ""))
    with pytest.raises(ValueError):
        crm_fit.Settings.from_toml_string("constants = 3 =")


def test_optimization_infos_order_and_sizes():
    s = crm_fit.Settings.from_toml_string(TOML)
    o = s.generate_optimization_infos(4)
    # radius, damping, strength, growth x4, en  (rigidity, tension, em are constants)
    assert len(o.bounds_lower) == len(o.bounds_upper) == len(o.initial_values) == 8
    assert [i[0] for i in o.parameter_infos] == ["Radius", "Damping", "Strength", "Growth Rate 0", "Growth Rate 1",
                                                 "Growth Rate 2", "Growth Rate 3", "Exponent n"]
    assert o.constants == [8.0, 3.0, 2.0] and [c[0] for c in o.constant_infos] == ["Rigidity", "Spring Tension",
                                                                                  "Exponent m"]
    assert o.bounds_upper[3:7] == [pytest.approx(0.08)] * 4


def test_define_initial_agents_maps_the_vector():
    s = crm_fit.Settings.from_toml_string(TOML)
    p2d = _rods(4)
    x = [1.25, 0.7, 0.11, 0.01, 0.02, 0.03, 0.04, 5.0]
    agents = crm_fit.define_initial_agents(x, p2d, s.domain_height, s.parameters, s.constants)
    assert len(agents) == 4
    for i, a in enumerate(agents):
        assert a.pos.shape == (6, 3) and np.all(a.pos[:, 2] == np.float32(1.25)) and np.all(a.vel == 0)
        assert np.array_equal(a.pos[:, :2], p2d[i])
        assert a.radius == np.float32(1.25) and a.growth_rate == np.float32(x[3 + i])
        assert a.growth_rate_setter.to_pydict() == {"mean": np.float32(x[3 + i]), "std": 0.0}
        assert a.spring_length_threshold == np.inf
        assert a.spring_length_threshold_setter.to_pydict() == {"l1": np.inf, "l2": np.inf}
        d = a.interaction.inner if hasattr(a.interaction, "inner") else a.interaction
        assert abs(float(a.mechanics.spring_length if hasattr(a, "mechanics") else a.spring_length) - 2.0) < 1e-5
    pos, vel, par, sim_of = crm_fit.define_initial_agents_packed(np.array([x, x]), p2d, 2.5, s.parameters, s.constants)
    assert pos.shape == (8, 6, 3) and list(sim_of) == [0] * 4 + [1] * 4
    assert np.all(par["potential"] == _capi.POT_MIE) and np.all(par["en"] == 5.0) and np.all(par["em"] == 2.0)
    assert np.all(par["bound"] == 50.0) and np.all(par["cutoff"] == 12.0) and np.all(par["damping"] == np.float32(0.7))
    assert np.allclose(par["growth_rate"], [0.01, 0.02, 0.03, 0.04] * 2) and np.all(par["rigidity"] == 8.0)
    with pytest.raises(IndexError):
        crm_fit.define_initial_agents(x[:-1], p2d, 2.5, s.parameters, s.constants)
    p3d = np.concatenate([p2d, np.full((4, 6, 1), 0.5, np.float32)], axis=2)
    assert np.all(crm_fit.define_initial_agents_packed([x], p3d, 2.5, s.parameters, s.constants)[0][:, :, 2] == 0.5)


@pytest.mark.gpu
def test_population_costs_equal_standalone_oracle_runs(product_lib):
    from oracle import oracle

    s = crm_fit.Settings.from_toml_string(TOML)
    n = 10
    p0 = _rods(n)
    rng = np.random.default_rng(1)
    data = np.stack([p0, p0 + rng.normal(0, 0.05, p0.shape), p0 + rng.normal(0, 0.1, p0.shape)]).astype(np.float32)
    o = s.generate_optimization_infos(n)
    lo, hi = np.array(o.bounds_lower), np.array(o.bounds_upper)
    X = (lo + rng.uniform(0.2, 0.8, (6, len(lo))) * (hi - lo)).astype(np.float32)
    X[3, 1] = np.nan  # damping: this member's run fails (CalcError) and costs error_cost, the others are untouched
    iterations = [0, 2, 4]
    got = crm_fit.predict_calculate_cost_population(X, data, iterations, s)
    assert got[3] == np.float32(s.constants.error_cost)
    assert crm_fit.predict_calculate_cost(X[1], data, iterations, s) == pytest.approx(float(got[1]), rel=1e-6)
    cfg = s.to_config()._to_c()
    for m in (0, 1, 2, 4, 5):
        pos, vel, par, _ = crm_fit.define_initial_agents_packed(X[m:m + 1], p0, 2.5, s.parameters, s.constants)
        sim = oracle.sim(cfg)
        sim.set_agents(pos, vel, par)
        sim.run()
        snaps = sorted(sim.snapshots(), key=lambda t: t[0])
        want = sum(float(((snaps[k][1][:, :, :2] - data[r]) ** 2).sum()) for r, k in enumerate(iterations))
        assert float(got[m]) == pytest.approx(want, rel=2e-4), m
