"""The raster oracle (oracle/raster_oracle.c) against the reference's own unit tests of the outline
(src/fitting.rs:755-893) and analytic known answers for the fill and the overlap area."""
import numpy as np
import pytest

from oracle import raster

PI = np.float32(np.pi)


def test_generate_coordinates_sphere_reference_vectors():
    # src/fitting.rs:755-767
    c = raster.arc((0.0, 0.0), 1.0, 0.0, PI, 2 * PI)
    assert len(c) == 1
    assert abs(c[0, 0]) < 1e-6 and abs(c[0, 1] - 1.0) < 1e-6
    # src/fitting.rs:770-791
    c = raster.arc((0.0, 0.0), 1.0, PI / 2, np.float32(1.5) * PI, PI / 5)
    assert len(c) == 3
    s = 1 / np.sqrt(2)
    np.testing.assert_allclose(c, [[-s, s], [-1, 0], [-s, -s]], atol=1e-3)


def test_joint_arcs_reference_vectors():
    # src/fitting.rs:794-819
    s = np.float32(1 / np.sqrt(2))
    right, c = raster.joint((-1, 0), (0, 0), (s, s), 0.5, PI / 4)
    assert not right and len(c) == 1
    np.testing.assert_allclose(c[0], [0.5 * np.sin(np.pi / 8), -0.5 * np.cos(np.pi / 8)], atol=1e-6)
    # src/fitting.rs:822-846
    right, c = raster.joint((0, 1), (0, 0), (1, 0), 0.9, PI / 8)
    assert not right and len(c) == 2
    d = np.pi / 2 / 3
    np.testing.assert_allclose(c, [[-0.9 * np.cos(d), -0.9 * np.sin(d)], [-0.9 * np.cos(2 * d), -0.9 * np.sin(2 * d)]],
                               atol=1e-6)


def test_polygon_hull_reference_vertex_count():
    # src/fitting.rs:849-866: 12 outline points for the three-vertex rod
    h = raster.hull([[-1, 0], [0, 0], [1, -1]], 0.5, PI / 4, 0.01)
    assert len(h) == 12


def test_polygon_hull_reference_edge_lengths():
    # src/fitting.rs:869-893
    p = [[82.003136, 83.25413], [80.756615, 84.66043], [79.50162, 86.086105], [78.22431, 87.53587],
         [76.95105, 88.98563], [75.63393, 90.48065], [74.1932, 92.11886], [72.57576, 93.95425]]
    h = raster.hull(p, 0.5, PI / 15, 0.01)
    d = np.linalg.norm(np.diff(np.vstack([h, h[:1]]), axis=0), axis=1)
    assert d.max() < 3.0
    # the outline encloses the capsule chain: area close to 2 r L + pi r^2
    L = np.linalg.norm(np.diff(np.asarray(p), axis=0), axis=1).sum()
    x, y = h[:, 0].astype(np.float64), h[:, 1].astype(np.float64)
    area = 0.5 * abs(np.dot(x, np.roll(y, -1)) - np.dot(y, np.roll(x, -1)))
    assert abs(area - (2 * 0.5 * L + np.pi * 0.25)) < 0.05


def _straight(x0, y0, angle, nv=8, sl=3.0):
    t = np.arange(nv) * sl
    p = np.zeros((nv, 3), np.float32)
    p[:, 0], p[:, 1], p[:, 2] = x0 + t * np.cos(angle), y0 + t * np.sin(angle), 1.0
    return p


def test_pixel_mapping_is_plotters_floor_with_bias():
    # RangedCoordf32::map: floor(res * v / size + 1e-3)
    assert raster.lib().raster_oracle_to_pixel(0.0, 100.0, 200) == 0
    assert raster.lib().raster_oracle_to_pixel(49.9996, 100.0, 200) == 100  # 99.9992 + 1e-3
    assert raster.lib().raster_oracle_to_pixel(49.99, 100.0, 200) == 99
    assert raster.lib().raster_oracle_to_pixel(100.0, 100.0, 200) == 200


def test_axis_aligned_capsule_pixel_count():
    """A horizontal rod of length 21 and radius 1.5 at 10 px / um: the filled pixels are the capsule outline's
    (two arc points per tip at delta = pi/4) within a pixel row/column of the analytic area."""
    pos = _straight(20.0, 50.0, 0.0)[None]
    mask, over = raster.render(pos, [1.5], [[10, 20, 30]], (100.0, 100.0), (1000, 1000))
    filled = (mask.reshape(-1, 3) != 0).any(axis=1).sum()
    h = raster.hull(pos[0, :, :2], 1.5)
    x, y = h[:, 0].astype(np.float64), h[:, 1].astype(np.float64)
    area = 0.5 * abs(np.dot(x, np.roll(y, -1)) - np.dot(y, np.roll(x, -1)))
    assert abs(filled - area * 100) < 0.02 * area * 100 + 300
    assert (mask[mask.any(axis=2)] == [10, 20, 30]).all()
    assert not over.any()
    # rows are flipped like the reference's invert_axis(0): y = 50 um is row 1000 - 500
    rows = np.nonzero(mask.any(axis=(1, 2)))[0]
    assert rows.min() >= 1000 - 516 and rows.max() <= 1000 - 484


def test_overlap_of_crossing_capsules_known_answer():
    """Two long thin perpendicular rods cross: the intersection is the 2r x 2r square."""
    a = _straight(10.0, 50.0, 0.0, nv=8, sl=10.0)
    b = _straight(45.0, 15.0, np.pi / 2, nv=8, sl=10.0)
    area = raster.overlap(np.stack([a, b]), [1.0, 1.0])
    assert abs(area - 4.0) < 1e-3
    # disjoint rods: zero; identical rods: the whole outline
    c = _straight(10.0, 80.0, 0.0, nv=8, sl=10.0)
    assert raster.overlap(np.stack([a, c]), [1.0, 1.0]) == 0.0
    h = raster.hull(a[:, :2], 1.0)
    x, y = h[:, 0].astype(np.float64), h[:, 1].astype(np.float64)
    own = 0.5 * abs(np.dot(x, np.roll(y, -1)) - np.dot(y, np.roll(x, -1)))
    shifted = a.copy()
    shifted[:, 1] += 1.0  # half overlapping in y
    ov = raster.overlap(np.stack([a, shifted]), [1.0, 1.0])
    assert 0.45 * own < ov < 0.55 * own


def test_render_overlap_pixels_and_draw_order():
    a = _straight(10.0, 50.0, 0.0, nv=8, sl=10.0)
    b = _straight(45.0, 15.0, np.pi / 2, nv=8, sl=10.0)
    mask, over = raster.render(np.stack([a, b]), [1.0, 1.0], [[1, 2, 3], [4, 5, 6]], (100.0, 100.0), (500, 500))
    n_over = over.any(axis=2).sum()
    assert abs(n_over - 4.0 * 25) <= 25  # the 2 x 2 um square at 5 px / um, up to its rim
    assert not mask[over.any(axis=2)].any()  # overlapping pixels are black in the mask
    cols = {tuple(c) for c in mask.reshape(-1, 3)}
    assert cols == {(0, 0, 0), (1, 2, 3), (4, 5, 6)}


def test_parents_diff_mask_rules():
    from cr_mech_coli_b200 import counter_to_color

    c1, c2, c3 = (np.array(counter_to_color(k), np.uint8) for k in (1, 2, 3))
    m1 = np.zeros((2, 3, 3), np.uint8)
    m2 = np.zeros((2, 3, 3), np.uint8)
    m1[0, 0], m2[0, 0] = c1, c1        # same colour -> 0
    m1[0, 1], m2[0, 1] = c1, c2        # cell 1 is the parent of cell 2 -> penalty
    m1[0, 2], m2[0, 2] = c2, c3        # unrelated -> 1
    m1[1, 0], m2[1, 0] = c1, (0, 0, 0)  # background vs cell -> 1
    cell_of_counter = np.array([-1, 0, 1, 2], np.int32)
    parent_of = np.array([-1, 0, -1], np.int32)
    d = raster.parents_diff(m1, m2, cell_of_counter, parent_of, 0.5)
    np.testing.assert_array_equal(d, [[0, 0.5, 1], [1, 0, 0]])
    with pytest.raises(KeyError):
        m2[1, 1] = np.array(counter_to_color(9), np.uint8)
        m1[1, 1] = c1
        raster.parents_diff(m1, m2, cell_of_counter, parent_of, 0.5)
