"""The public Python API on the GPU: README example (reference README.md:24-38, tests/test_getting_started.py)
and the colour assignment after a simulation (reference tests/test_counter_color_conversion.py:23-49)."""
import numpy as np
import pytest

import cr_mech_coli_b200 as crm
from cr_mech_coli_b200 import _capi
from oracle import oracle

pytestmark = pytest.mark.gpu


def test_getting_started(product_lib):
    config = crm.Configuration()
    agents = crm.generate_agents(4, crm.AgentSettings(), config)
    cc = crm.run_simulation_with_agents(config, agents)
    its = cc.get_all_iterations()
    assert its == [0, 91, 182, 273, 364, 455, 545, 636, 727, 818, 909, 1000]  # n_saves + 2 snapshots
    first, last = cc.get_cells_at_iteration(0), cc.get_cells_at_iteration(1000)
    assert len(first) == 4 and len(last) == 8  # every rod divided once (growth 0.01, threshold 6)
    I = crm.CellIdentifier.new_initial
    assert sorted(first) == [I(k) for k in range(4)]
    assert all(np.array_equal(first[I(k)][0].pos, agents[k].pos) for k in range(4))
    dm = cc.get_daughter_map()
    assert sorted(dm) == [I(k) for k in range(4)] and all(len(v) == 2 for v in dm.values())
    assert all(cc.get_parent(d) == p for p, ds in dm.items() for d in ds)
    # the same run through the oracle gives the same lineage
    pos = np.stack([a.pos for a in agents])
    from cr_mech_coli_b200.simulation import pack_agents
    with oracle.sim(config._to_c()) as o:
        o.set_agents(*pack_agents(agents))
        o.run()
        ids = o.snapshots()[-1][4]
    assert sorted(last) == sorted(crm.CellIdentifier._from_record(r) for r in ids)


def test_assign_colors_after_simulation(product_lib):
    config = crm.Configuration(t_max=20.0)
    s = crm.AgentSettings()
    positions = crm.generate_positions(8, s, config)
    agents = [crm.RodAgent(pos=p, vel=p * 0.0, **s.to_rod_agent_dict()) for p in positions]
    cc = crm.run_simulation_with_agents(config, agents)
    cells = cc.get_cells()[cc.get_all_iterations()[0]]
    for n, ident in enumerate(sorted(cells)):
        assert cc.cell_to_color[ident] == crm.counter_to_color(n + 1)
        assert cc.get_cell_from_color(cc.cell_to_color[ident]) == ident


def test_simulation_error_is_raised(product_lib):
    config = crm.Configuration(domain_size=(60.0, 60.0), t_max=5.0)
    s = crm.AgentSettings()
    p = np.zeros((8, 3), np.float32)
    p[:, 0] = 10 + 3 * np.arange(8)
    p[:, 1:] = (30.0, 1.25)
    v = np.zeros_like(p)
    v[:, 1] = 5e3
    with pytest.raises(crm.SimulationError, match="BoundaryError"):
        crm.run_simulation_with_agents(config, [crm.RodAgent(pos=p, vel=v, **s.to_rod_agent_dict())])


def test_empty_agent_list(product_lib):
    cc = crm.run_simulation_with_agents(crm.Configuration(t_max=1.0, n_saves=1), [])
    assert cc.get_all_iterations() == [0, 5, 10] and cc.get_cells_at_iteration(5) == {}


def test_pinned_buffers_round_trip(product_lib):
    """crm_host_alloc buffers in, get_state(out=...) buffers out: same values as the pageable path."""
    import numpy as np
    from cr_mech_coli_b200 import _capi
    from cr_mech_coli_b200.testing import lattice_colony, lattice_domain, mie_params

    n = 40_000  # above the host-thread split of the parameter conversion
    L = lattice_domain(n)
    cfg = _capi.make_config(domain_size=(L, L), dt=0.01, t_max=1e6, n_saves=0)
    pos = lattice_colony(n, seed=3)
    par = mie_params(n, growth_rate=0.0)
    par["en"][::7] = 5.0  # the constant cache must follow every change of (en, em, cutoff)
    par["cutoff"][::5] = 5.5
    a = _capi.Sim(product_lib, cfg)
    a.set_fused(False)  # (last_forces: the fused list step keeps its forces in registers)
    a.set_agents(pos, np.zeros_like(pos), par)
    a.step(3)
    ref = a.get_state()
    fa = a.last_forces()
    pp = [product_lib.pinned_empty(x.shape, x.dtype) for x in (pos, pos, par)]
    pp[0][...] = pos
    pp[1][...] = 0
    pp[2][...] = par
    out = (product_lib.pinned_empty((n + 5, 8, 3), np.float32), product_lib.pinned_empty((n, 8, 3), np.float32),
           product_lib.pinned_empty(n, _capi.ROD_PARAMS_DTYPE), product_lib.pinned_empty(n, _capi.IDENT_DTYPE),
           product_lib.pinned_empty(n, _capi.IDENT_DTYPE))
    b = _capi.Sim(product_lib, cfg)
    b.set_fused(False)
    b.set_agents(*pp)
    b.step(3)
    got = b.get_state(out=out)
    assert got[0].base is not None and len(got[0]) == n
    for x, y in zip(ref, got):
        assert x.tobytes() == y.tobytes()
    assert fa.tobytes() == b.last_forces().tobytes()
    with pytest.raises(ValueError):
        b.get_state(out=(out[0][: n - 1],) + out[1:])


def test_serde_json_storage_round_trip(product_lib, tmp_path):
    """StorageOption.SerdeJson: the run writes its save points below storage_location/<date>/cells/json/...,
    the container remembers the run directory, and load_from_storage returns the same container
    (reference src/simulation.rs:307-316, src/cell_container.rs:322-359; the multilayer driver's
    load_or_compute_container, cr_mech_coli/crm_multilayer/runner.py:130-150)."""
    import cr_mech_coli as crm_shim

    config = crm_shim.Configuration(t_max=80.0, n_saves=3, storage_location=str(tmp_path / "out"),
                                    storage_options=[crm_shim.StorageOption.Memory, crm_shim.StorageOption.SerdeJson])
    agents = crm_shim.generate_agents(4, crm_shim.AgentSettings(), config, rng_seed=0, randomize_positions=0.1)
    cc = crm_shim.run_simulation_with_agents(config, agents)
    assert cc.path is not None and (cc.path / "cells" / "json").is_dir()
    assert cc.path.parent == tmp_path / "out"
    back = crm_shim.CellContainer.load_from_storage(config, cc.path)
    assert back.get_all_iterations() == cc.get_all_iterations() and len(cc.get_all_iterations()) == 5
    assert back.get_parent_map() == cc.get_parent_map() and len(cc.get_cells_at_iteration(800)) == 8
    for it in cc.get_all_iterations():
        a, b = cc.get_cells_at_iteration(it), back.get_cells_at_iteration(it)
        assert list(a) == list(b)
        for ident in a:
            assert np.array_equal(a[ident][0].pos, b[ident][0].pos) and np.array_equal(a[ident][0].vel, b[ident][0].vel)
            assert a[ident][0].spring_length == b[ident][0].spring_length and a[ident][1] == b[ident][1]
    # a suffix becomes a sub-directory of the run directory
    config.storage_suffix = "trial-7"
    cc2 = crm_shim.run_simulation_with_agents(config, agents)
    assert cc2.path.name == "trial-7" and (cc2.path / "cells" / "json").is_dir()


def test_run_simulation_with_arrays_equals_the_object_api(product_lib):
    """The packed entry point (colonies too large for a list of RodAgent objects) returns the container the object
    API returns."""
    config = crm.Configuration(t_max=30.0, n_saves=2)
    s = crm.AgentSettings(growth_rate=0.03)
    positions = crm.generate_positions(6, s, config, rng_seed=2)
    agents = [crm.RodAgent(pos=p, vel=p * 0.0, **s.to_rod_agent_dict()) for p in positions]
    a = crm.run_simulation_with_agents(config, agents)
    pos = np.stack(positions)
    b = crm.run_simulation_with_arrays(config, pos, np.zeros_like(pos), crm.params_from_settings(s, len(pos)))
    assert a.get_all_iterations() == b.get_all_iterations()
    assert a.get_parent_map() == b.get_parent_map()
    for it in a.get_all_iterations():
        assert np.array_equal(a.positions_at_iteration(it), b.positions_at_iteration(it))
