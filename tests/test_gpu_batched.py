"""Batched independent simulations (BASELINE config 5: crm_fit ensembles, many sims per GPU, no
communication): one device state, one tile of cells per simulation.  Every simulation must equal its
stand-alone run — against the CPU oracle and against the stand-alone CUDA path."""
import numpy as np
import pytest

import cr_mech_coli_b200 as crm
from cr_mech_coli_b200 import _capi
from cr_mech_coli_b200.testing import lattice_colony, morse_params
from oracle import oracle

pytestmark = pytest.mark.gpu


def _ensemble(n_sims, seed=0):
    rng = np.random.default_rng(seed)
    sims = []
    for m in range(n_sims):
        n = int(rng.integers(3, 9))
        pos = lattice_colony(n, spacing=(40.0, 9.0), margin=30.0, seed=100 + m, angle_std=0.4, domain=200.0)
        par = morse_params(n, radius=float(rng.uniform(2.0, 3.5)), strength=float(rng.uniform(0.05, 0.2)),
                           potential_stiffness=float(rng.uniform(0.3, 0.8)), cutoff=10.0,
                           damping=float(rng.uniform(0.5, 1.5)))
        par["growth_rate"] = rng.uniform(0.04, 0.1, n)
        par["growth_setter_kind"] = _capi.SETTER_EXPLICIT
        par["growth_setter_a"] = 0.07
        par["growth_setter_b"] = 0.05
        par["spring_length"] = rng.uniform(2.6, 3.4, n)
        sims.append((pos, np.zeros_like(pos), par))
    return sims


def test_batched_sims_equal_standalone_oracle(product_lib):
    n_sims = 24
    cfg = _capi.make_config(domain_size=(200.0, 200.0), n_voxels=(2, 2), t_max=25.0, dt=0.1, n_saves=4)
    sims = _ensemble(n_sims)
    pos = np.concatenate([s[0] for s in sims])
    vel = np.concatenate([s[1] for s in sims])
    par = np.concatenate([s[2] for s in sims])
    sim_of = np.concatenate([np.full(len(s[0]), m, np.uint32) for m, s in enumerate(sims)])
    # interleave the rods of different sims: the grouping must not rely on the input order
    # (a shuffle of whole sims that keeps each sim's internal order)
    key = np.random.default_rng(2).random(n_sims)[sim_of] + np.arange(len(pos)) * 1e-9
    order = np.argsort(key, kind="stable")
    with _capi.Sim(product_lib, cfg) as g:
        g.set_agents_batched(pos[order], vel[order], par[order], sim_of[order], n_sims)
        g.run()
        per_sim = g.snapshots_by_sim(n_sims)
        assert g.list_stats()["builds"] >= 1
    divided = 0
    for m, (p0, v0, par0) in enumerate(sims):
        with oracle.sim(cfg) as o:
            o.set_agents(p0, v0, par0)
            o.run()
            so = o.snapshots()
        sg = per_sim[m]
        assert len(sg) == len(so) == 6
        for (it, pg, vg, parg, idg, prg), (_, po, vo, paro, ido, pro) in zip(sg, so):
            assert pg.shape == po.shape, f"sim {m}: rod count differs at iteration {it}"
            assert (idg == ido).all() and (prg == pro).all(), f"sim {m}: lineage differs at iteration {it}"
            assert (parg["spring_length"] == paro["spring_length"]).all()
            # short horizon; later the contact dynamics of freshly divided rods amplifies f32 summation-order
            # noise (the stand-alone CUDA run deviates from the oracle by exactly the same amount:
            # scripts/diag_batched.py), so only counts, lineage and growth state are compared
            if it <= 100:
                assert np.abs(pg - po).max() < 1e-3, f"sim {m} iteration {it}: {np.abs(pg - po).max()}"
        divided += len(sg[-1][1]) > len(p0)
    assert divided > n_sims // 2  # the ensemble does exercise divisions


def test_run_simulations_batched_api():
    config = crm.Configuration(t_max=30.0, n_saves=2, domain_size=[120.0, 120.0])
    lists = []
    for m in range(5):
        settings = crm.AgentSettings(growth_rate=0.03 + 0.01 * m)
        lists.append(crm.generate_agents(3 + m % 3, settings, config, rng_seed=m))
    batched = crm.run_simulations_batched(config, lists)
    assert len(batched) == 5
    for agents, cb in zip(lists, batched):
        single = crm.run_simulation_with_agents(config, agents)
        assert cb.get_all_iterations() == single.get_all_iterations()
        for it in single.get_all_iterations():
            a, b = single.get_cells_at_iteration(it), cb.get_cells_at_iteration(it)
            assert sorted(a.keys()) == sorted(b.keys())
            for ident in a:
                assert np.abs(a[ident][0].pos - b[ident][0].pos).max() < 1e-4
                assert a[ident][1] == b[ident][1]
        assert single.get_parent_map() == cb.get_parent_map()


def _fit_population(n_sims, n_agents=6, seed=5):
    """Members of a fit: the same initial rods, different interaction parameters per member."""
    rng = np.random.default_rng(seed)
    pos0 = lattice_colony(n_agents, spacing=(30.0, 8.0), margin=60.0, seed=77, angle_std=0.3, domain=200.0)
    members = []
    for m in range(n_sims):
        par = morse_params(n_agents, radius=float(rng.uniform(2.0, 3.5)), strength=float(rng.uniform(0.05, 0.3)),
                           potential_stiffness=float(rng.uniform(0.3, 0.8)), cutoff=10.0,
                           damping=float(rng.uniform(0.5, 1.5)))
        par["growth_rate"] = rng.uniform(0.0, 0.02, n_agents)
        par["spring_length_threshold"] = 1e30
        members.append(par)
    return pos0, members


def test_costs_on_device_equal_the_snapshot_path(product_lib):
    """crm_sim_run_costs (kernel reduction at the result iterations, no snapshots) against
    costs_from_snapshots on the snapshots of the same batched run and on stand-alone oracle runs."""
    from cr_mech_coli_b200.crm_fit import costs_from_snapshots, costs_on_device

    n_sims, n_agents = 40, 6
    cfg = _capi.make_config(domain_size=(200.0, 200.0), n_voxels=(1, 1), t_max=12.0, dt=0.05, n_saves=3)
    pos0, members = _fit_population(n_sims, n_agents)
    pos = np.concatenate([pos0] * n_sims)
    vel = np.zeros_like(pos)
    par = np.concatenate(members)
    sim_of = np.repeat(np.arange(n_sims, dtype=np.uint32), n_agents)
    rng = np.random.default_rng(1)
    data = (pos0[None, :, :, :2] + rng.normal(0, 0.5, (3, n_agents, 8, 2))).astype(np.float32)
    iterations = [1, 2, 4]  # indices into the sorted save iterations (5 here: 0, 3 saves, t_max)
    costs, status = costs_on_device(product_lib, cfg, pos, vel, par, sim_of, n_sims, data, iterations, error_cost=-1.0)
    # a few of the random members push crossing rods out of the 2.5 um high domain: exactly those are
    # flagged (the oracle raises BoundaryError for them), the others are unaffected
    ok = np.nonzero(status == 0)[0]
    failed = np.nonzero(status != 0)[0]
    assert len(ok) >= n_sims - 6 and (costs[failed] == -1.0).all()
    for m in failed[:2]:
        assert status[m] == _capi.ERR_BOUNDARY
        with pytest.raises(_capi.SimulationError):
            with oracle.sim(cfg) as o:
                o.set_agents(pos0, np.zeros_like(pos0), members[m])
                o.run()
    keep = np.isin(sim_of, ok)
    renum = np.cumsum(np.isin(np.arange(n_sims), ok)) - 1
    with _capi.Sim(product_lib, cfg) as g:  # the snapshot path on the members that survive
        g.set_agents_batched(pos[keep], vel[keep], par[keep], renum[sim_of[keep]].astype(np.uint32), len(ok))
        g.run()
        per_sim = g.snapshots_by_sim(len(ok))
    want = costs_from_snapshots(per_sim, data, iterations)
    assert np.allclose(costs[ok], want, rtol=2e-5), np.abs(costs[ok] / want - 1).max()
    for m in ok[[0, len(ok) // 2, -1]]:  # stand-alone oracle runs of a few members
        with oracle.sim(cfg) as o:
            o.set_agents(pos0, np.zeros_like(pos0), members[m])
            o.run()
            so = sorted(o.snapshots(), key=lambda s: s[0])
        c = sum(float(((so[i][1][:, :, :2] - data[r]) ** 2).sum()) for r, i in enumerate(iterations))
        assert abs(costs[m] - c) < 1e-3 * c


def test_a_failing_member_costs_only_itself(product_lib):
    """Error isolation: one member is thrown out of the domain (BoundaryError in a stand-alone run);
    it gets the error cost, every other member its usual cost."""
    from cr_mech_coli_b200.crm_fit import costs_on_device

    n_sims, n_agents = 8, 6
    cfg = _capi.make_config(domain_size=(200.0, 200.0), n_voxels=(1, 1), t_max=6.0, dt=0.05, n_saves=1)
    pos0, members = _fit_population(n_sims, n_agents)
    pos = np.concatenate([pos0] * n_sims)
    vel = np.zeros_like(pos)
    par = np.concatenate(members)
    sim_of = np.repeat(np.arange(n_sims, dtype=np.uint32), n_agents)
    data = np.repeat(pos0[None, :, :, :2], 2, axis=0).astype(np.float32)
    good, st0 = costs_on_device(product_lib, cfg, pos, vel, par, sim_of, n_sims, data, [1, 2])
    assert (st0 == 0).all()
    bad = 5
    vel_bad = vel.copy()
    vel_bad[sim_of == bad, :, 0] = 5000.0  # leaves the 200 um domain in the first step, reflection cannot save it
    with pytest.raises(_capi.SimulationError):  # without isolation the run stops
        with _capi.Sim(product_lib, cfg) as g:
            g.set_agents_batched(pos, vel_bad, par, sim_of, n_sims)
            g.run()
    costs, status = costs_on_device(product_lib, cfg, pos, vel_bad, par, sim_of, n_sims, data, [1, 2], error_cost=777.0)
    assert status[bad] == _capi.ERR_BOUNDARY and costs[bad] == 777.0
    others = np.arange(n_sims) != bad
    assert (status[others] == 0).all()
    assert np.array_equal(costs[others], good[others])
