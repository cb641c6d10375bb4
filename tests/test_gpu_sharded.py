"""Voxel-slab sharding on the device: several slabs of one colony, each a crm_sim with
halo exchange, migration and deferred divisions through the C ABI, against the unsharded CUDA path
and the CPU oracle.  LocalCluster drives all slabs in one process (one GPU is enough); the NCCL
transport between processes is covered by test_two_process_nccl when two GPUs are visible."""
import os
import subprocess
import sys

import numpy as np
import pytest

from cr_mech_coli_b200 import _capi, sharded
from cr_mech_coli_b200.testing import lattice_colony, lattice_domain, mie_params, morse_params, random_colony
from oracle import oracle

pytestmark = pytest.mark.gpu


def _rel(a, b):
    s = np.abs(b).max()
    return float(np.abs(a - b).max() / s) if s > 0 else float(np.abs(a).max())


@pytest.mark.parametrize("world", [2, 3])
def test_forces_match_unsharded_and_oracle(product_lib, world):
    n = 3000
    L = lattice_domain(n)
    nvx = max(1, int(L // 40))
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(nvx, nvx))
    pos = lattice_colony(n, seed=3)
    par = mie_params(n)
    vel = np.zeros_like(pos)
    part = sharded.SlabPartition.balanced(sharded.centroid_x(pos), world)
    cl = sharded.LocalCluster(product_lib, cfg, part)
    cl.set_agents(pos, vel, par)
    assert sum(s.n_rods for s in cl.shards) == n and min(s.n_rods for s in cl.shards) > n // (2 * world)
    fs, ns = cl.compute_forces()
    with _capi.Sim(product_lib, cfg) as g, oracle.sim(cfg) as o:
        g.set_agents(pos, vel, par)
        o.set_agents(pos, vel, par)
        fg, ng = g.compute_forces()
        fo, no = o.compute_forces()
    assert _rel(fs, fo) < 1e-5
    assert _rel(fs, fg) < 2e-6  # same arithmetic, only the summation order over neighbours differs
    assert (ns == no).all() and (ns == ng).all()
    cl.close()


def test_dense_random_colony_three_slabs(product_lib):
    n, L = 900, 300.0
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(2, 2))
    pos = random_colony(n, L, seed=2)
    par = morse_params(n, radius=1.0, cutoff=5.0, strength=0.2)
    par["radius"] = np.random.default_rng(1).uniform(0.8, 1.5, n)
    vel = np.zeros_like(pos)
    cl = sharded.LocalCluster(product_lib, cfg, sharded.SlabPartition.uniform(L, 3))
    cl.set_agents(pos, vel, par)
    fs, ns = cl.compute_forces()
    with oracle.sim(cfg) as o:
        o.set_agents(pos, vel, par)
        fo, no = o.compute_forces()
    assert _rel(fs, fo) < 1e-5
    assert (ns == no).all()
    cl.close()


def test_trajectory_with_migration(product_lib):
    """Rods drift across the slab edges: ownership changes hands, the trajectory does not."""
    n = 1200
    L = lattice_domain(n)
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(1, 1), dt=0.05)
    pos = lattice_colony(n, seed=5)
    par = mie_params(n, growth_rate=0.0, damping=0.0)
    par["spring_length_threshold"] = 1e30
    vel = np.zeros_like(pos)
    vel[:, :, 0] = 1.5  # the whole colony moves +x: 0.075 um per step
    part = sharded.SlabPartition.balanced(sharded.centroid_x(pos), 3)
    cl = sharded.LocalCluster(product_lib, cfg, part)
    cl.set_agents(pos, vel, par)
    n0 = [s.n_rods for s in cl.shards]
    steps = 200  # 15 um: more than half a lattice column changes owner at each edge
    cl.step(steps)
    n1 = [s.n_rods for s in cl.shards]
    assert sum(n1) == n and n1[0] < n0[0] and n1[2] > n0[2], (n0, n1)
    ps, vs, pars, ids, parents = cl.get_state()
    with _capi.Sim(product_lib, cfg) as g:
        g.set_agents(pos, vel, par)
        g.step(steps)
        pg, vg, parg, idg, prg = g.get_state()
    assert (ids == idg).all() and (parents == prg).all()
    assert np.abs(ps - pg).max() < 1e-3
    assert np.abs(vs - vg).max() < 2e-3  # (1.5 um/min drift, no damping: f32 summation-order noise of 200 steps)
    # every rod is owned by the slab that holds its centroid
    for r, s in enumerate(cl.shards):
        lo, hi = part.bounds(r)
        c = sharded.centroid_x(s.sim.get_state()[0])
        assert ((c >= lo) & (c < hi)).all()
    cl.close()


@pytest.mark.parametrize("world", [2, 4])
def test_growth_division_lineage_bit_exact(product_lib, world):
    """Growing colony with divisions in several slabs and user voxels: identifiers, parents, counts and
    growth state equal the unsharded run and the oracle bit for bit (fixed division schedule)."""
    n, L = 36, 400.0
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(3, 3), t_max=30.0, dt=0.1, n_saves=14)
    pos = lattice_colony(n, spacing=(60.0, 40.0), margin=40.0, seed=11, angle_std=0.5)
    par = morse_params(n, growth_rate=0.08)
    rng = np.random.default_rng(4)
    par["growth_rate"] = rng.uniform(0.05, 0.1, n)
    par["growth_setter_kind"] = _capi.SETTER_EXPLICIT
    par["growth_setter_a"] = 0.09
    par["growth_setter_b"] = 0.06
    par["spring_length"] = rng.uniform(2.5, 3.5, n)
    vel = np.zeros_like(pos)
    cl = sharded.LocalCluster(product_lib, cfg, sharded.SlabPartition.uniform(L, world))
    cl.set_agents(pos, vel, par)
    ss = cl.run()
    with oracle.sim(cfg) as o:
        o.set_agents(pos, vel, par)
        o.run()
        so = o.snapshots()
    assert len(ss) == len(so) == 16
    for (it, ps, vs, pars, ids, prs), (_, po, vo, paro, ido, pro) in zip(ss, so):
        assert ps.shape == po.shape, f"rod count differs at iteration {it}"
        assert (ids == ido).all() and (prs == pro).all(), f"lineage differs at iteration {it}"
        assert (pars["spring_length"] == paro["spring_length"]).all()
        assert (pars["growth_rate"] == paro["growth_rate"]).all()
        if it <= 140:
            assert np.abs(ps - po).max() < 1e-3, f"iteration {it}: {np.abs(ps - po).max()}"
    assert len(ss[-1][1]) > 4 * n
    cl.close()


def test_two_process_nccl(product_lib):
    """One process per GPU over NCCL (torchrun): merged snapshots equal the unsharded run."""
    if product_lib.fn("device_count")() < 2:
        pytest.skip("needs two GPUs")
    worker = os.path.join(os.path.dirname(__file__), "dist_worker.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29631", worker],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "dist_worker ok" in r.stdout


def test_sharded_persistent_lists(product_lib):
    """Collective rebuilds with frozen halos: the slabs reuse their persistent neighbour structure
    between rebuilds (most steps) and still follow the unsharded trajectory; rods that drift over
    the slab edges force rebuilds and change owner."""
    n = 1500
    L = lattice_domain(n)
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(1, 1), dt=0.01)
    pos = lattice_colony(n, seed=6)
    par = mie_params(n, growth_rate=0.0, damping=0.0)
    par["spring_length_threshold"] = 1e30
    vel = np.zeros_like(pos)
    vel[:, :, 0] = 1.0  # 0.01 um per step: a rebuild every ~25 steps, a few rods change owner
    # lattice columns sit at x = 14 + 24 k (+- 0.25 jitter): columns 5 and 10 cross these edges
    # between steps ~150 and ~200; outside that window nothing changes owner
    part = sharded.SlabPartition([14 + 5 * 24 + 1.75, 14 + 10 * 24 + 1.75])
    cl = sharded.LocalCluster(product_lib, cfg, part)
    cl.set_agents(pos, vel, par)
    n0 = [s.n_rods for s in cl.shards]
    steps = 400
    cl.step(steps)
    stats = [s.sim.list_stats() for s in cl.shards]
    assert all(st["eval_steps"] > steps // 2 for st in stats), stats
    assert all(3 <= st["builds"] < steps // 4 for st in stats), stats
    n1 = [s.n_rods for s in cl.shards]
    assert sum(n1) == n and n1 != n0
    ps, vs, pars, ids, parents = cl.get_state()
    with _capi.Sim(product_lib, cfg) as g:
        g.set_agents(pos, vel, par)
        g.step(steps)
        pg, vg, parg, idg, prg = g.get_state()
    assert (ids == idg).all()
    assert np.abs(ps - pg).max() < 1e-3
    cl.close()


@pytest.mark.parametrize("world", [2, 3])
def test_stochastic_run_is_partition_independent(product_lib, world):
    """Wiener term and Normal setters are keyed by the GLOBAL slot (the slot a rod has in the unsharded run), so a
    sharded stochastic run draws exactly the numbers of the oracle: lineage bit-exact, positions 1e-3 um."""
    n, L = 48, 420.0
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(3, 3), t_max=16.0, dt=0.1, n_saves=7, rng_seed=11)
    pos = lattice_colony(n, spacing=(50.0, 40.0), margin=40.0, seed=12, angle_std=0.5)
    par = morse_params(n, growth_rate=0.08, diffusion_constant=0.02)
    rng = np.random.default_rng(5)
    par["growth_rate"] = rng.uniform(0.06, 0.1, n)
    par["growth_setter_kind"] = _capi.SETTER_NORMAL
    par["growth_setter_a"] = 0.08
    par["growth_setter_b"] = 0.015
    par["threshold_setter_kind"] = _capi.SETTER_NORMAL
    par["threshold_setter_a"] = 6.0
    par["threshold_setter_b"] = 0.4
    par["spring_length"] = rng.uniform(2.5, 3.5, n)
    vel = np.zeros_like(pos)
    cl = sharded.LocalCluster(product_lib, cfg, sharded.SlabPartition.uniform(L, world))
    cl.set_agents(pos, vel, par)
    ss = cl.run()
    with oracle.sim(cfg) as o:
        o.set_agents(pos, vel, par)
        o.run()
        so = o.snapshots()
    assert len(ss) == len(so) == 9
    for (it, ps, vs, pars, ids, prs), (_, po, vo, paro, ido, pro) in zip(ss, so):
        assert ps.shape == po.shape, f"rod count differs at iteration {it}"
        assert (ids == ido).all() and (prs == pro).all(), f"lineage differs at iteration {it}"
        assert (pars["growth_rate"] == paro["growth_rate"]).all()
        assert (pars["spring_length_threshold"] == paro["spring_length_threshold"]).all()
        # (short horizon 1e-3 um; later the contact dynamics of the new daughters amplify f32 summation noise)
        assert np.abs(ps - po).max() < (1e-3 if it <= 100 else 1e-2), f"iteration {it}: {np.abs(ps - po).max()}"
    assert len(ss[-1][1]) > n
    cl.close()
