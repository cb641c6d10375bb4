"""Known-answer tests of the CPU oracle (the reference pins no forces, so the restatement
is pinned on analytic cases of the published model, reference docs/source/model.rst:18-124)."""
import numpy as np
import pytest

from cr_mech_coli_b200 import _capi
from cr_mech_coli_b200.testing import mie_params, morse_params
from oracle import oracle


def straight_rod(x0, y0, angle=0.0, nv=8, sl=3.0, z=1.25):
    t = np.arange(nv) * sl
    p = np.zeros((nv, 3), np.float32)
    p[:, 0], p[:, 1], p[:, 2] = x0 + t * np.cos(angle), y0 + t * np.sin(angle), z
    return p


def forces(pos, par, **cfg):
    c = _capi.make_config(**cfg)
    with oracle.sim(c) as o:
        o.set_agents(pos, np.zeros_like(pos), par)
        return o.compute_forces()


def test_morse_parallel_rods_at_equilibrium_distance():
    # two parallel rods exactly r_own + r_ext apart: every nearest-point distance >= R, and the pairs
    # directly opposite sit at the potential minimum -> the lateral force of the interior vertices is the
    # (attractive) sum of the oblique neighbours only, and the x components cancel by symmetry
    R = 3.0
    pos = np.stack([straight_rod(20, 20), straight_rod(20, 20 + R)])
    f, nn = forces(pos, morse_params(2, radius=1.5, cutoff=100.0, potential_stiffness=0.5, strength=0.1))
    assert np.allclose(f[0, :, 1], -f[1, :, 1], atol=1e-7)          # action = reaction
    assert np.all(f[0, :, 1] >= -1e-7)                              # only attraction towards the partner
    assert np.allclose(f[0, :, 0], -f[0, ::-1, 0], atol=1e-6)       # mirror symmetry along the rod
    assert np.allclose(f[:, :, 2], 0.0)
    assert list(nn) == [0, 0]  # sensing threshold is (r+r)/sqrt(2) < R


def test_morse_single_pair_magnitude():
    # two 2-vertex rods far longer than the range of interest reduce to the analytic point force
    nv = 2
    a = np.array([[0, 10, 1], [40, 10, 1]], np.float32)
    d = 2.0  # < R: repulsion
    b = np.array([[20, 10 + d, 1], [20, 60, 1]], np.float32)  # perpendicular rod whose end faces the middle of a
    par = morse_params(2, radius=1.5, cutoff=2.5, potential_stiffness=0.5, strength=0.1)
    f, _ = forces(np.stack([a, b]), par, domain_size=(100.0, 100.0))
    lam, V0, R = 0.5, 0.1, 3.0
    e = np.exp(-lam * (d - R))
    mag = 2 * lam * V0 * e * (e - 1)  # repulsive magnitude for d < R
    # only (b_0 vs segment a) and (a-vertices vs b's segment: beyond cutoff) contribute; both calls see the
    # same pair (b_0, midpoint of a) once: own share of B's call and ext share of ... A's call is zero
    # (A's vertices are 20 um away) -> force on b_0 = 1/2 * mag, reaction split 1/2:1/2 onto a's vertices
    assert f[1, 0, 1] == pytest.approx(0.5 * mag, rel=1e-5)
    assert f[0, 0, 1] == pytest.approx(-0.25 * mag, rel=1e-5)
    assert f[0, 1, 1] == pytest.approx(-0.25 * mag, rel=1e-5)
    assert np.allclose(f[:, :, 0], 0, atol=1e-7)


def test_mie_force_sign_bound_and_cutoff():
    a = np.array([[0, 10, 1], [40, 10, 1]], np.float32)
    def fy(d, **kw):
        b = np.array([[20, 10 + d, 1], [20, 60, 1]], np.float32)
        f, _ = forces(np.stack([a, b]), mie_params(2, radius=1.5, cutoff=5.0, en=6.0, em=3.0, **kw), domain_size=(100.0, 100.0))
        return f[1, 0, 1]
    assert fy(3.0) == pytest.approx(0.0, abs=1e-6)       # minimum of the potential at r_own + r_ext
    assert fy(2.5) > 0 and fy(3.5) < 0                   # repulsive inside, attractive outside
    assert fy(5.5) == 0.0                                # beyond the cutoff
    assert fy(0.5, bound=1.0) == pytest.approx(0.5)      # clamped to `bound` (x 1/2 weight)
    # analytic value: strength*C/sigma*(n x^(n+1) - m x^(m+1)), C = n/(n-m)*(n/m)^(n/(n-m)), sigma=(m/n)^(1/(n-m)) R
    n, m, R, d = 6.0, 3.0, 3.0, 2.8
    sigma = (m / n) ** (1 / (n - m)) * R
    Cc = n / (n - m) * (n / m) ** (n / (n - m))
    mag = 0.1 * Cc / sigma * (n * (sigma / d) ** (n + 1) - m * (sigma / d) ** (m + 1))
    assert fy(d) == pytest.approx(0.5 * mag, rel=1e-5)


def test_spring_and_bending_increment():
    # one stretched, one bent 3-vertex rod; forces follow docs/source/model.rst eq. force-springs/-stiffness
    cfg = _capi.make_config(dt=1e-3, t_max=1e-3, n_saves=0)
    par = morse_params(1)
    par["spring_tension"], par["rigidity"], par["damping"], par["growth_rate"] = 2.0, 0.0, 0.0, 0.0
    pos = np.array([[[10, 10, 1], [14, 10, 1], [18, 10, 1]]], np.float32)  # segments of 4 vs rest length 3
    with oracle.sim(cfg) as o:
        o.set_agents(pos, np.zeros_like(pos), par)
        o.step(1)
        v = o.get_state()[1][0] / 1e-3  # Euler start-up: dv = F*dt
    assert np.allclose(v[:, 0], [2.0, 0.0, -2.0], atol=1e-5) and np.allclose(v[:, 1:], 0, atol=1e-6)
    par["spring_tension"], par["rigidity"] = 0.0, 3.0
    th = 0.3
    pos = np.array([[[10 - 3 * np.cos(th), 10 - 3 * np.sin(th), 1], [10, 10, 1], [10 + 3 * np.cos(th), 10 - 3 * np.sin(th), 1]]], np.float32)
    with oracle.sim(cfg) as o:
        o.set_agents(pos, np.zeros_like(pos), par)
        o.step(1)
        v = o.get_state()[1][0] / 1e-3
    mag = 3.0 * (2 * th)  # eta * (pi - alpha), alpha = pi - 2 theta; restoring = pushes the apex down
    assert v[1, 1] == pytest.approx(-mag, rel=1e-4) and v[0, 1] == pytest.approx(mag / 2, rel=1e-4)
    assert np.allclose(v.sum(axis=0), 0, atol=1e-5)  # the stiffness force does not move the centre


def test_division_geometry_and_lineage():
    cfg = _capi.make_config(dt=0.1, t_max=0.1, n_saves=0, n_voxels=(2, 2))
    par = morse_params(1, radius=1.0)
    par["spring_length"], par["spring_length_threshold"], par["growth_rate"] = 2.0, 2.0, 0.1
    par["spring_tension"], par["rigidity"], par["damping"] = 0.0, 0.0, 0.0
    pos = straight_rod(60, 70, angle=0.4, sl=2.0)[None]
    with oracle.sim(cfg) as o:
        o.set_agents(pos, np.zeros_like(pos), par)
        o.step(1)
        p, v, pr, ids, parents = o.get_state()
    assert len(p) == 2
    L = 14.0
    # each daughter spans (L - 2 r)/2 along the mother, leaving a gap of 2 r at the septum
    for d in p:
        assert np.linalg.norm(d[-1] - d[0]) == pytest.approx((L - 2.0) / 2, rel=1e-5)
    assert np.allclose(p[0][0], pos[0][0], atol=1e-5) and np.allclose(p[1][-1], pos[0][-1], atol=1e-5)
    assert np.linalg.norm(p[1][0] - p[0][-1]) == pytest.approx(2.0, rel=1e-4)
    sl = 2.0 * (1 + 0.1 * 0.1)
    assert pr["spring_length"][0] == pytest.approx(sl * (0.5 - 1.0 / (8 * sl)), rel=1e-6)
    # both daughters get fresh Division ids in the mother's voxel (60,70 in a 2x2 grid of a 100^2 domain -> 1*2+1)
    assert ids["kind"].tolist() == [0, 0] and ids["voxel"].tolist() == [3, 3] and ids["counter"].tolist() == [1, 2]
    assert parents["kind"].tolist() == [1, 1] and parents["counter"].tolist() == [0, 0]


def test_boundary_reflection_and_error():
    cfg = _capi.make_config(dt=0.1, t_max=1.0, n_saves=0, domain_size=(50.0, 50.0))
    par = morse_params(1, growth_rate=0.0)
    par["spring_tension"], par["rigidity"], par["damping"] = 0.0, 0.0, 0.0
    pos = straight_rod(1.0, 20.0)[None]
    vel = np.zeros_like(pos)
    vel[:, :, 0] = -20.0  # crosses x = 0 by 1 um in one step
    with oracle.sim(cfg) as o:
        o.set_agents(pos, vel, par)
        o.step(1)
        p, v = o.get_state()[:2]
    assert p[0, 0, 0] == pytest.approx(1.0) and v[0, 0, 0] == pytest.approx(20.0)   # mirrored, velocity flipped
    assert p[0, 1, 0] == pytest.approx(2.0) and v[0, 1, 0] == pytest.approx(-20.0)  # stayed inside
    vel[:, :, 0] = -2000.0
    with oracle.sim(cfg) as o:
        o.set_agents(pos, vel, par)
        with pytest.raises(_capi.SimulationError, match="BoundaryError"):
            o.step(1)


def test_neighbor_sensing_and_reduction():
    pos = np.stack([straight_rod(20, 20), straight_rod(20, 21.5), straight_rod(20, 40)])
    par = morse_params(3, radius=1.5, cutoff=6.0, neighbor_reduction=(2, 1.0))
    f, nn = forces(pos, par)
    assert list(nn) == [1, 1, 0]  # |dy| = 1.5 < (1.5+1.5)/sqrt(2)
    cfg = _capi.make_config(dt=0.1, t_max=0.1, n_saves=0)
    with oracle.sim(cfg) as o:
        o.set_agents(pos, np.zeros_like(pos), par)
        o.step(1)
        pr = o.get_state()[2]
    # rate = growth_rate*((max-n)/max)^exp: halved for the two sensing rods
    assert pr["spring_length"][2] == pytest.approx(3.0 * (1 + 0.01 * 0.1), rel=1e-6)
    assert pr["spring_length"][0] == pytest.approx(3.0 * (1 + 0.005 * 0.1), rel=1e-6)


def test_colors_match_reference_formula():
    import ctypes as C
    lib = oracle.library()
    lib.lib.oracle_counter_to_color.argtypes = [C.c_uint32, C.POINTER(C.c_uint8 * 3)]
    lib.lib.oracle_color_to_counter.restype = C.c_uint32
    for counter in range(0, 251**3, 9973):
        rgb = (C.c_uint8 * 3)()
        lib.lib.oracle_counter_to_color(counter, C.byref(rgb))
        c = (counter * 157 * 163 * 173) % 251**3
        assert tuple(rgb) == (c // 251**2, (c % 251**2) // 251, c % 251)
        assert lib.lib.oracle_color_to_counter(C.byref(rgb)) == counter


def test_pair_once_schedule_equals_gather_form(oracle_lib):
    """The timing schedule of bench.py's CPU arm (each unordered pair visited once, four half-weighted results
    scattered to both rods: the reference backend's schedule) gives the gather form's forces up to
    f32 summation order, and the same neighbour counts and trajectory."""
    from cr_mech_coli_b200.testing import lattice_colony, lattice_domain, mie_params
    from oracle import oracle

    n = 400
    L = lattice_domain(n)
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(max(1, int(L // 40)),) * 2, dt=0.01)
    pos = lattice_colony(n, seed=9)
    par = mie_params(n)
    outs = []
    for once in (0, 1):
        with oracle.sim(cfg, openmp=True) as o:
            o.L.check(o.L.fn("sim_set_pair_once")(o.h, once))
            o.set_agents(pos, np.zeros_like(pos), par)
            o.step(5)
            f = np.zeros_like(pos)
            o.L.check(o.L.fn("sim_last_forces")(o.h, f.ctypes.data))
            outs.append((f, o.get_state()))
    (f0, s0), (f1, s1) = outs
    assert np.abs(f0).max() > 0
    assert np.abs(f1 - f0).max() < 2e-6 * np.abs(f0).max()
    assert (s0[2]["n_neighbors"] == s1[2]["n_neighbors"]).all()
    assert np.abs(s0[0] - s1[0]).max() < 1e-5


def test_solver_order_switch(oracle_lib):
    """Euler / AB2 / AB3 on a single stretched rod: all contract the rod, pairwise different at O(dt^2)."""
    from cr_mech_coli_b200.testing import morse_params
    from oracle import oracle

    cfg = _capi.make_config(dt=0.1, t_max=5.0)
    pos = np.zeros((1, 8, 3), np.float32)
    pos[0, :, 0] = 30 + 3.5 * np.arange(8)
    pos[0, :, 1] = 50
    pos[0, :, 2] = 1.25
    par = morse_params(1, growth_rate=0.0)
    ends = []
    for order in (0, 1, 2):
        with oracle.sim(cfg) as o:
            o.L.check(o.L.fn("sim_set_solver_order")(o.h, order))
            o.set_agents(pos, np.zeros_like(pos), par)
            o.step(20)
            ends.append(o.get_state()[0])
    length = [float(p[0, -1, 0] - p[0, 0, 0]) for p in ends]
    assert all(l < 24.5 for l in length)
    assert abs(length[0] - length[1]) > 1e-4 and abs(length[1] - length[2]) > 1e-5
    with oracle.sim(cfg) as o:  # half the spring force per end: slower contraction
        o.L.check(o.L.fn("sim_set_spring_end_weight")(o.h, 0.5))
        o.set_agents(pos, np.zeros_like(pos), par)
        o.step(20)
        half = o.get_state()[0]
    assert float(half[0, -1, 0] - half[0, 0, 0]) > length[2] + 1e-3
