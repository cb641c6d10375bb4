"""CUDA mask path (csrc/crm_raster.cu, through the C ABI and its Python mirror) against the raster oracle:
outlines within f32 rounding of the trigonometric functions, bitmaps byte-exact, overlap area 1e-4 relative."""
import numpy as np
import pytest

from cr_mech_coli_b200 import fitting
from cr_mech_coli_b200.cell_container import counter_to_color
from cr_mech_coli_b200.testing import lattice_colony, random_colony
from oracle import raster

pytestmark = pytest.mark.gpu


def _readme_colony():
    pos = np.load(__file__.replace("test_gpu_raster.py", "golden/readme_positions_seed0.npy")).astype(np.float32)
    rng = np.random.default_rng(3)
    bent = pos.copy()
    t = np.arange(8) - 3.5
    bent[:, :, 1] += (rng.normal(0, 0.04, (len(pos), 1)) * t[None, :] ** 2).astype(np.float32)  # curved rods
    return np.concatenate([pos, bent + np.float32(1.7)])


def test_outlines_match_the_oracle(product_lib):
    pos = random_colony(300, 200.0, seed=4, bend=0.12)
    radii = np.random.default_rng(1).uniform(0.4, 1.5, len(pos)).astype(np.float32)
    for delta in (np.pi / 4, np.pi / 15):
        hulls = fitting.polygon_hulls(pos, radii, delta)
        for k in range(len(pos)):
            ho = raster.hull(pos[k, :, :2], float(radii[k]), delta)
            assert hulls[k].shape == ho.shape, (k, hulls[k].shape, ho.shape)
            assert np.array_equal(hulls[k], ho), (k, np.abs(hulls[k] - ho).max())  # bit for bit (decision R0)
    # the reference's own vertex-count vector (src/fitting.rs:849-866)
    h = fitting.polygon_hulls(np.array([[[-1, 0], [0, 0], [1, -1]]], np.float32), [0.5], np.pi / 4)
    assert len(h[0]) == 12


@pytest.mark.parametrize("resolution", [(512, 512), (1031, 777)])
def test_readme_colony_bitmaps_byte_exact(product_lib, resolution):
    pos = _readme_colony()
    n = len(pos)
    radii = np.full(n, 3.0, np.float32)
    colors = np.array([counter_to_color(k + 1) for k in range(n)], np.uint8)
    mg, og = fitting.render_mask_arrays(pos, radii, colors, (100.0, 100.0), resolution)
    mo, oo = raster.render(pos, radii, colors, (100.0, 100.0), resolution)
    assert mg.shape == (resolution[1], resolution[0], 3)
    assert mo.any() and oo.any()  # cells and overlaps are there (the shifted copies overlap the originals)
    assert np.array_equal(mg, mo), int((mg != mo).any(axis=2).sum())
    assert np.array_equal(og, oo), int((og != oo).any(axis=2).sum())


def test_dense_colony_bitmap_and_overlap(product_lib):
    pos = random_colony(800, 300.0, seed=9, bend=0.08)
    n = len(pos)
    radii = np.random.default_rng(2).uniform(0.5, 1.2, n).astype(np.float32)
    colors = np.array([counter_to_color(k + 1) for k in range(n)], np.uint8)
    mg, og = fitting.render_mask_arrays(pos, radii, colors, (300.0, 300.0), (1500, 1500))
    mo, oo = raster.render(pos, radii, colors, (300.0, 300.0), (1500, 1500))
    assert np.array_equal(mg, mo), int((mg != mo).any(axis=2).sum())
    assert np.array_equal(og, oo), int((og != oo).any(axis=2).sum())
    ag = fitting.overlap(pos, radii)
    ao = raster.overlap(pos, radii)
    assert ao > 10.0
    assert abs(ag - ao) < 1e-4 * ao, (ag, ao)
    # raster cross-check of the area: overlap pixels x pixel area (boundary pixels: a few per cent)
    px_area = og.any(axis=2).sum() * (300.0 / 1500) ** 2
    assert abs(px_area - ao) < 0.25 * ao + 5.0


def test_overlap_known_answers(product_lib):
    def straight(x0, y0, angle, sl=10.0):
        t = np.arange(8) * sl
        p = np.zeros((8, 3), np.float32)
        p[:, 0], p[:, 1] = x0 + t * np.cos(angle), y0 + t * np.sin(angle)
        return p
    a, b = straight(10.0, 50.0, 0.0), straight(45.0, 15.0, np.pi / 2)
    assert abs(fitting.overlap(np.stack([a, b]), [1.0, 1.0]) - 4.0) < 1e-3  # crossing: the 2r x 2r square
    assert fitting.overlap(np.stack([a, straight(10.0, 80.0, 0.0)]), [1.0, 1.0]) == 0.0


def test_render_mask_2d_and_parents_diff_through_the_mirror(product_lib):
    """The reference-facing signatures: cells dict + cell_to_color in, masks out; parents_diff_mask on two
    masks of a colony before / after a division."""
    import cr_mech_coli_b200 as crm

    config = crm.Configuration(t_max=80.0)
    agents = crm.generate_agents(4, crm.AgentSettings(), config, rng_seed=0)
    container = crm.run_simulation_with_agents(config, agents)
    its = container.get_all_iterations()
    cells0, cells1 = container.get_cells_at_iteration(its[0]), container.get_cells_at_iteration(its[-1])
    assert len(cells1) == 8
    m0, _ = fitting.render_mask_2d(cells0, container.cell_to_color, config.domain_size, (400, 400))
    m1, _ = fitting.render_mask_2d(cells1, container.cell_to_color, config.domain_size, (400, 400))
    assert {tuple(c) for c in m0.reshape(-1, 3)} - {(0, 0, 0)} == {container.cell_to_color[i] for i in cells0}
    d = fitting.parents_diff_mask(m0, m1, container.color_to_cell, container.parent_map, 0.5)
    assert d.shape == (400, 400) and set(np.unique(d)) <= {0.0, 0.5, 1.0}
    assert (d == 0.5).sum() > 100  # daughters lie where their mothers were
    # oracle on the same tables
    cells = sorted(container.parent_map)
    index = {c: k for k, c in enumerate(cells)}
    table = np.full(len(cells) + 1, -1, np.int32)
    for c, col in container.cell_to_color.items():
        table[crm.color_to_counter(col)] = index[c]
    parent_of = np.array([index[container.parent_map[c]] if container.parent_map[c] is not None else -1 for c in cells],
                         np.int32)
    assert np.array_equal(d, raster.parents_diff(m0, m1, table, parent_of, 0.5))
    with pytest.raises(KeyError):  # an unknown colour facing a cell (facing background it is just "different")
        bad = m1.copy()
        y, x = np.argwhere(m0.any(axis=2))[0]
        bad[y, x] = counter_to_color(77)
        fitting.parents_diff_mask(m0, bad, container.color_to_cell, container.parent_map)
    with pytest.raises(ValueError):
        fitting.parents_diff_mask(m0, m1[:10], container.color_to_cell, container.parent_map)


def test_jagged_rod_raises_like_the_reference(product_lib):
    """A zig-zag rod whose offset lines neither cross nor meet within epsilon: the reference returns CalcError
    ("lines should be intersecting", src/fitting.rs:572-577), the oracle and the CUDA path agree."""
    p = np.array([[65.08, 38.62], [68.04, 37.17], [69.97, 38.29], [72.35, 32.63], [75.49, 42.28], [77.72, 43.21],
                  [82.19, 32.83], [86.62, 35.50]], np.float32)
    with pytest.raises(ValueError, match="lines should be intersecting"):
        raster.hull(p, 3.0)
    with pytest.raises(Exception, match="lines should be intersecting"):
        fitting.polygon_hulls(p[None], [3.0])
