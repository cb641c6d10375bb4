"""BASELINE.json's full size (configs[2]: 1e6 rods x 8 vertices, Mie, the bench.py workload) through
properties that do not need the oracle to step a million rods: momentum balance, oracle spot checks of
windows cut out of the colony, independence of the rod order, list steps vs search steps."""
import numpy as np
import pytest

import bench
from cr_mech_coli_b200 import _capi
from oracle import oracle

pytestmark = pytest.mark.gpu

N = 1_000_000


@pytest.fixture(scope="module")
def colony(product_lib):
    cfg, pos, par = bench.workload(N, seed=12345)
    sim = _capi.Sim(product_lib, cfg)
    sim.set_agents(pos, np.zeros_like(pos), par)
    force, nn = sim.compute_forces()
    yield cfg, pos, par, sim, force, nn
    sim.close()


def test_momentum_balance(colony):
    """Every interaction is split between the two rods with opposite signs (reference pair return,
    src/agent.rs:387-394): the forces of the whole colony sum to zero up to f32 rounding."""
    _, _, _, _, force, _ = colony
    assert np.isfinite(force).all()
    total = force.astype(np.float64).sum(axis=(0, 1))
    scale = np.abs(force).astype(np.float64).sum(axis=(0, 1))
    assert scale[0] > 0 and scale[1] > 0
    assert np.all(np.abs(total[:2]) < 1e-6 * scale[:2]), (total, scale)
    # planar colony: z forces are rounding residue of `a(1-t) + bt` at z = 1.25 (reproduced from the reference)
    assert np.abs(force[:, :, 2]).max() < 1e-5 * np.abs(force).max()


@pytest.mark.parametrize("corner", [(0.0, 0.0), (0.47, 0.52), (0.93, 0.9)])
def test_oracle_spot_check(colony, corner):
    """Forces of the rods inside a 300 um window = forces the oracle computes for that window plus a
    margin of (rod length + cutoff) cut out of the colony (1e-5 relative, the north-star tolerance)."""
    cfg, pos, par, _, force, nn = colony
    L = cfg.domain_size[0]
    c = pos[:, :, :2].mean(axis=1)
    x0, y0 = corner[0] * (L - 300.0), corner[1] * (L - 300.0)
    margin = 30.0
    outer = (c[:, 0] > x0 - margin) & (c[:, 0] < x0 + 300 + margin) & (c[:, 1] > y0 - margin) & (c[:, 1] < y0 + 300 + margin)
    idx = np.nonzero(outer)[0]
    inner = (c[idx, 0] > x0) & (c[idx, 0] < x0 + 300) & (c[idx, 1] > y0) & (c[idx, 1] < y0 + 300)
    assert inner.sum() > 500
    o = oracle.sim(cfg, openmp=True)
    o.set_agents(pos[idx], np.zeros_like(pos[idx]), par[idx])
    fo, no = o.compute_forces()
    fg = force[idx]
    scale = np.abs(fo[inner]).max()
    assert scale > 0
    assert np.abs(fg[inner] - fo[inner]).max() / scale < 1e-5
    assert (nn[idx][inner] == no[inner]).all()


def test_rod_order_does_not_matter(colony, product_lib):
    """The same colony uploaded in a random rod order: every rod gets the same force (the neighbour
    search sums a rod's partners in slot order, so only rounding may differ)."""
    cfg, pos, par, _, force, nn = colony
    perm = np.random.default_rng(0).permutation(N)
    sim = _capi.Sim(product_lib, cfg)
    sim.set_agents(pos[perm], np.zeros_like(pos), par[perm])
    f2, n2 = sim.compute_forces()
    sim.close()
    scale = np.abs(force).max()
    assert np.abs(f2 - force[perm]).max() / scale < 1e-5
    assert (n2 == nn[perm]).all()


def test_list_steps_equal_search_steps(colony, product_lib):
    """20 steps with the persistent neighbour structure vs 20 steps searching every step."""
    cfg, pos, par, _, _, _ = colony
    out = []
    for skin in (0.5, 0.0):
        sim = _capi.Sim(product_lib, cfg)
        sim.set_skin(skin)
        sim.set_agents(pos, np.zeros_like(pos), par)
        sim.step(20)
        out.append((sim.get_state()[0], sim.list_stats()))
        sim.close()
    assert out[0][1]["eval_steps"] >= 15 and out[1][1]["eval_steps"] == 0
    assert np.isfinite(out[0][0]).all()
    # coordinates reach 1e4 um (one ulp = 1e-3 um): compare in units in the last place
    ulps = np.abs(out[0][0] - out[1][0]) / np.spacing(np.abs(out[1][0]))
    assert ulps.max() <= 4.0, ulps.max()
    assert np.abs(out[0][0] - pos).max() > 1e-4  # the colony did move
