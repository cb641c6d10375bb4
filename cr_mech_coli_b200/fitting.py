"""Mask path of the fitting objectives (reference src/fitting.rs, exported through cr_mech_coli/fitting.py):
``render_mask_2d``, ``overlap``, ``parents_diff_mask``, plus the outline they share.  All compute runs in
CUDA kernels behind the C ABI (csrc/crm_raster.cu); there is no CPU fallback and no graphics API."""

from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from .cell_container import color_to_counter

FRAC_PI_4 = float(np.float32(np.pi / 4))


def _rods(cells):
    """(pos, radii) of a ``{CellIdentifier: (RodAgent, parent)}`` mapping in identifier order (the order the
    reference's BTreeMap iterates in) plus the sorted identifiers."""
    idents = sorted(cells)
    n = len(idents)
    nv = cells[idents[0]][0]._pos.shape[0] if n else 8
    pos = np.zeros((n, nv, 3), np.float32)
    radii = np.zeros(n, np.float32)
    for k, i in enumerate(idents):
        agent = cells[i][0]
        pos[k] = agent._pos
        radii[k] = agent.radius
    return idents, pos, radii


def polygon_hulls(positions, radii, delta_angle=FRAC_PI_4, epsilon=0.01, cap=256):
    """``calculate_polygon_hull`` (reference src/fitting.rs:514-620) for every rod: list of (m_i, 2) f32 arrays."""
    lib = _capi.product()
    pos = np.ascontiguousarray(positions, dtype=np.float32)
    if pos.ndim != 3:
        raise ValueError("positions must be (n_rods, n_vertices, 2 or 3)")
    if pos.shape[2] == 2:
        pos = np.ascontiguousarray(np.concatenate([pos, np.zeros(pos.shape[:2] + (1,), np.float32)], axis=2))
    radii = np.ascontiguousarray(radii, dtype=np.float32)
    n = len(pos)
    out = np.zeros((n, cap, 2), np.float32)
    cnt = np.zeros(n, np.uint32)
    lib.check(lib.fn("polygon_hulls")(pos.ctypes.data, radii.ctypes.data, n, pos.shape[1], float(delta_angle),
                                      float(epsilon), out.ctypes.data, cnt.ctypes.data, cap))
    return [out[k, :cnt[k]].copy() for k in range(n)]


def render_mask_arrays(pos, radii, colors, domain_size, resolution, delta_angle=FRAC_PI_4, epsilon=0.01,
                       return_ms: bool = False):
    """Packed-array form of ``render_mask_2d``: rods in draw order with their colours [n, 3] u8."""
    lib = _capi.product()
    pos = np.ascontiguousarray(pos, dtype=np.float32)
    radii = np.ascontiguousarray(radii, dtype=np.float32)
    colors = np.ascontiguousarray(colors, dtype=np.uint8)
    rx, ry = int(resolution[0]), int(resolution[1])
    mask = np.zeros((ry, rx, 3), np.uint8)
    over = np.zeros((ry, rx, 3), np.uint8)
    ds = (C.c_float * 2)(float(domain_size[0]), float(domain_size[1]))
    rs = (C.c_uint32 * 2)(rx, ry)
    ms = C.c_double(0)
    lib.check(lib.fn("render_mask_2d")(pos.ctypes.data, radii.ctypes.data, colors.ctypes.data, len(pos), pos.shape[1],
                                       C.byref(ds), C.byref(rs), float(delta_angle), float(epsilon), mask.ctypes.data,
                                       over.ctypes.data, C.byref(ms)))
    return (mask, over, float(ms.value)) if return_ms else (mask, over)


def render_mask_2d(cells, cell_to_color, domain_size, resolution, delta_angle=FRAC_PI_4, epsilon=0.01):
    """reference src/fitting.rs:111-244: (mask, overlap_mask), both (resolution[1], resolution[0], 3) u8."""
    idents, pos, radii = _rods(cells)
    colors = np.array([cell_to_color[i] for i in idents], np.uint8).reshape(len(idents), 3)
    try:
        return render_mask_arrays(pos, radii, colors, domain_size, resolution, delta_angle, epsilon)
    except _capi.SimulationError as e:
        raise ValueError(str(e)) from e


def overlap(positions, radii, delta_angle=FRAC_PI_4, epsilon=0.01) -> float:
    """reference src/fitting.rs:299-345: approximate overlap of the agents as a 2-D area."""
    lib = _capi.product()
    pos = np.ascontiguousarray(positions, dtype=np.float32)
    if pos.ndim != 3:
        raise ValueError("positions must be (n_rods, n_vertices, 2 or 3)")
    if pos.shape[2] == 2:
        pos = np.ascontiguousarray(np.concatenate([pos, np.zeros(pos.shape[:2] + (1,), np.float32)], axis=2))
    radii = np.ascontiguousarray(radii, dtype=np.float32)
    area = C.c_float(0)
    lib.check(lib.fn("overlap")(pos.ctypes.data, radii.ctypes.data, len(pos), pos.shape[1], float(delta_angle),
                                float(epsilon), C.byref(area)))
    return float(area.value)


def parents_diff_mask(mask1, mask2, color_to_cell, parent_map, parent_penalty=0.5):
    """reference src/fitting.rs:29-109: per-pixel difference of two instance masks; ``parent_penalty`` where one
    cell is the parent of the other.  Raises ValueError on mismatching shapes, KeyError on an unknown colour."""
    m1 = np.ascontiguousarray(mask1, dtype=np.uint8)
    m2 = np.ascontiguousarray(mask2, dtype=np.uint8)
    if m1.shape != m2.shape or m1.ndim != 3 or m1.size == 0:
        raise ValueError(f"Masks need to have matching nonempty shapes. Got shapes: {list(m1.shape)} {list(m2.shape)}")
    cells = sorted(set(color_to_cell.values()) | {p for p in parent_map.values() if p is not None} | set(parent_map))
    index = {c: k for k, c in enumerate(cells)}
    counters = {color_to_counter(col): index[cell] for col, cell in color_to_cell.items()}
    table = np.full(max(counters, default=0) + 1, -1, np.int32)
    for k, v in counters.items():
        table[k] = v
    parent_of = np.full(len(cells), -1, np.int32)
    known = np.zeros(len(cells), bool)
    for c, p in parent_map.items():
        known[index[c]] = True
        parent_of[index[c]] = index[p] if p is not None else -1
    # a coloured cell without an entry in parent_map is a KeyError in the reference (:83-92)
    for k, v in counters.items():
        if not known[v]:
            table[k] = -1
    lib = _capi.product()
    out = np.zeros(m1.shape[:2], np.float32)
    miss = C.c_int64(-1)
    lib.check(lib.fn("parents_diff_mask")(m1.ctypes.data, m2.ctypes.data, m1.shape[0] * m1.shape[1], table.ctypes.data,
                                          len(table), parent_of.ctypes.data, len(parent_of), float(parent_penalty),
                                          out.ctypes.data, C.byref(miss)))
    if miss.value >= 0:
        y, x = divmod(int(miss.value), m1.shape[1])
        raise KeyError(f"could not find color {tuple(m1[y, x])} / {tuple(m2[y, x])}")
    return out
