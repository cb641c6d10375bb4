"""Boundary classes: the reference's own API tests (tests/test_config.py, tests/test_create_potentials.py,
tests/test_counter_color_conversion.py) restated against cr_mech_coli_b200, plus the container maps."""
import copy
import pickle

import numpy as np
import pytest

import cr_mech_coli_b200 as crm


def test_config_set_attributes():  # reference tests/test_config.py:4-26
    assert crm.Configuration(domain_size=[1000, 900]).domain_size == [1000.0, 900.0]
    assert crm.Configuration(n_threads=2).n_threads == 2
    assert abs(crm.Configuration(domain_height=10.0).domain_height - 10.0) < 1e-8
    assert crm.Configuration(n_voxels=[3, 2]).n_voxels == [3, 2]
    with pytest.raises(AttributeError):
        crm.Configuration(not_an_attribute=1)


def test_config_defaults():  # reference src/config.rs:444-465
    c = crm.Configuration()
    assert (c.n_threads, c.t0, c.n_saves, c.rng_seed, c.n_voxels) == (1, 0.0, 10, 0, [1, 1])
    assert c.dt == np.float32(0.1) and c.t_max == 100.0 and c.domain_height == 2.5
    assert c.storage_options == [crm.StorageOption.Memory] and c.progressbar is None


def test_config_roundtrips():  # reference src/config.rs:566-606
    c1, c2 = crm.Configuration(), crm.Configuration(n_saves=100)
    assert c1.to_hash() != c2.to_hash()
    assert crm.Configuration.from_json(c2.to_json()) == c2
    assert pickle.loads(pickle.dumps(c2)) == c2
    assert copy.deepcopy(c2) == c2
    toml = ("n_threads=1\nt0=0.0\ndt=0.1\nt_max=100.0\nn_saves=10\ndomain_size=[100.0, 100.0]\ndomain_height=2.5\n"
            "n_voxels=[1, 1]\nrng_seed=0\ngel_pressure=0\nsurface_friction=0\nsurface_friction_distance=1\n"
            "storage_options=['Memory']\nstorage_location='out'\n")
    c = crm.Configuration.from_toml_string(toml)
    assert c.dt == np.float32(0.1) and c.t_max == 100.0
    with pytest.raises(ValueError):
        crm.Configuration.from_toml_string("dt = [")


def test_agent_settings():  # reference tests/test_config.py:29-38
    s = crm.AgentSettings(growth_rate_setter={"mean": 0.003, "std": 937.85})
    assert abs(s.growth_rate_setter.mean - 0.003) < 1e-3 and abs(s.growth_rate_setter.std - 937.85) < 1e-3
    s = crm.AgentSettings(growth_rate_setter={"g1": 0.03, "g2": -0.1})
    assert abs(s.growth_rate_setter.g1 - 0.03) < 1e-5 and abs(s.growth_rate_setter.g2 + 0.1) < 1e-5
    with pytest.raises(KeyError):
        crm.AgentSettings(growth_rate_setter={"mean": 1.0, "g1": 2.0})


def test_agent_settings_defaults_and_fallthrough():  # reference src/config.rs:243-301
    s = crm.AgentSettings()
    assert (s.interaction.radius, s.interaction.potential_stiffness, s.interaction.cutoff) == (3.0, 0.5, 10.0)
    assert s.growth_rate == np.float32(0.01) and s.spring_length_threshold == 6.0 and s.neighbor_reduction is None
    assert s.spring_length_threshold_setter.to_pydict() == {"l1": 6.0, "l2": 6.0}
    m = s.mechanics
    assert (m.diffusion_constant, m.spring_tension, m.rigidity, m.spring_length, m.damping) == (0.0, 1.0, 2.0, 3.0, 1.0)
    s = crm.AgentSettings(radius=1.5, rigidity=7.0, growth_rate=0.2, neighbor_reduction=(200, 0.5))
    assert s.interaction.radius == 1.5 and s.mechanics.rigidity == 7.0 and s.neighbor_reduction == (200, 0.5)
    with pytest.raises(AttributeError):
        crm.AgentSettings(no_such_thing=1)


def test_interaction_assign():  # reference tests/test_config.py:41-51
    s = crm.AgentSettings()
    s.interaction.strength = 0.1091
    assert abs(s.interaction.strength - 0.1091) < 1e-4
    s.interaction.potential_stiffness = 11.1
    assert abs(s.interaction.potential_stiffness - 11.1) < 1e-3
    s.interaction.cutoff = 33.2391
    assert abs(s.interaction.cutoff - 33.2391) < 1e-5
    s.interaction.radius = 1.93e-8
    assert abs(s.interaction.radius - 1.93e-8) < 1e-10
    with pytest.raises(AttributeError):
        s.interaction.en = 3.0  # Morse has no exponent


def test_create_potentials():  # reference tests/test_create_potentials.py:4-28
    p1 = crm.MorsePotentialF32(radius=1.0, potential_stiffness=0.5, cutoff=3.0, strength=0.1)
    p2 = crm.MiePotentialF32(radius=1.0, strength=0.1, bound=3.0, cutoff=3.0, en=2.0, em=1.0)
    pot1, pot2 = crm.PhysicalInteraction(p1), crm.PhysicalInteraction(p2)
    assert (pot1.radius, pot1.cutoff, pot1.potential_stiffness) == (1.0, 3.0, 0.5)
    assert (pot2.radius, pot2.cutoff, pot2.en, pot2.em) == (1.0, 3.0, 2.0, 1.0)
    assert crm.PhysicalInteraction(pot2) == pot2
    with pytest.raises(ValueError):
        crm.PhysicalInteraction("morse")


def test_counter_color_conversion():  # reference tests/test_counter_color_conversion.py:5-20
    for counter in range(1, 251**3, 1301):
        assert crm.color_to_counter(crm.counter_to_color(counter)) == counter
    for i in range(1, 251, 17):
        for j in range(1, 251, 19):
            for k in range(1, 251, 23):
                assert crm.counter_to_color(crm.color_to_counter((i, j, k))) == (i, j, k)
    assert crm.counter_to_color(1) == ((157 * 163 * 173) // 251**2, ((157 * 163 * 173) % 251**2) // 251,
                                       (157 * 163 * 173) % 251)


def test_rod_agent_ctor_and_record_roundtrip():
    s = crm.AgentSettings(neighbor_reduction=(16, 0.5), growth_rate_setter={"g1": 0.1, "g2": 0.2})
    s.interaction = crm.MiePotentialF32(1.5, 0.1, 100.0, 6.0, 6.0, 3.0)
    p = np.arange(24, dtype=np.float32).reshape(8, 3)
    a = crm.RodAgent(pos=p, vel=p * 0.0, **s.to_rod_agent_dict())
    assert a.radius == 1.5 and np.array_equal(a.pos, p)
    a.radius = 2.0
    assert a.interaction.radius == 2.0
    from cr_mech_coli_b200 import _capi
    rec = np.zeros(1, _capi.ROD_PARAMS_DTYPE)
    a._to_record(rec[0])
    assert crm.RodAgent._from_record(a.pos, a.vel, rec[0]) == a
    # defaults of the constructor: both setters default to {mean: 0.01, std: 0} (reference src/agent.rs:28-35,465)
    b = crm.RodAgent(p, p * 0, crm.MorsePotentialF32(3, 0.5, 10, 0.1))
    assert b.spring_length_threshold_setter.to_pydict() == {"mean": pytest.approx(0.01), "std": 0.0}


def test_cell_identifier_order_and_container_maps():
    I, D = crm.CellIdentifier.new_initial, crm.CellIdentifier
    assert D(5, 9) < I(0) < crm.CellIdentifier.new_inserted(0, 0)  # Division < Initial < Inserted
    assert D(0, 3) < D(1, 1) and I(2)[0] == 2 and D(4, 7)[0] == 4 and D(4, 7)[1] == 7
    assert crm.sort_cellular_identifiers([I(1), D(0, 2), I(0), D(0, 1)]) == [D(0, 1), D(0, 2), I(0), I(1)]
    s = crm.AgentSettings()
    p = np.zeros((8, 3), np.float32)
    ag = crm.RodAgent(pos=p, vel=p, **s.to_rod_agent_dict())
    cells = {0: {I(0): (ag, None), I(1): (ag, None)}, 7: {D(0, 1): (ag, I(0)), D(0, 2): (ag, I(0)), I(1): (ag, None)}}
    cc = crm.CellContainer(cells)
    assert cc.get_all_iterations() == [0, 7]
    assert cc.get_parent(D(0, 2)) == I(0) and cc.get_parent(I(1)) is None
    assert cc.get_daughter_map() == {I(0): [D(0, 1), D(0, 2)]}
    assert cc.get_children(I(0)) == [D(0, 1), D(0, 1), D(0, 2)]  # reference quirk, src/cell_container.rs:86-92
    assert cc.cells_are_siblings(D(0, 1), D(0, 2)) and not cc.cells_are_siblings(D(0, 1), I(1))
    ids = cc.get_all_identifiers()
    assert ids == [D(0, 1), D(0, 2), I(0), I(1)]
    assert [cc.get_color(i) for i in ids] == [crm.counter_to_color(n + 1) for n in range(4)]
    assert cc.get_cell_from_color(crm.counter_to_color(3)) == I(0)
    assert cc.cell_identifier_to_counter(I(1)) == 3 and cc.counter_to_cell_identifier(0) == D(0, 1)
    hist, parent = cc.get_cell_history(I(1))
    assert sorted(hist) == [0, 7] and parent is None
    with pytest.raises(KeyError):
        cc.get_parent(D(9, 9))
    new = cc.add_ident_divided(I(1))
    assert cc.get_parent(new) == I(1) and cc.get_color(new) == crm.counter_to_color(5)
    cc2 = crm.CellContainer.deserialize(cc.serialize())
    assert cc2.get_cells()[7][D(0, 2)][0] == ag and cc2.get_cells()[7][D(0, 2)][1] == I(0)
