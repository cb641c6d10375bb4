"""ctypes loader of the raster oracle (TEST INFRASTRUCTURE ONLY — see raster_oracle.c)."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        p = _DIR / "libraster_oracle.so"
        if not p.exists():
            subprocess.run(["make", "-C", str(_DIR)], check=True, capture_output=True)
        L = C.CDLL(str(p))
        vp, f, i = C.c_void_p, C.c_float, C.c_int
        L.raster_oracle_hull.restype = i
        L.raster_oracle_hull.argtypes = [vp, i, f, f, f, vp, i]
        L.raster_oracle_arc.restype = i
        L.raster_oracle_arc.argtypes = [f, f, f, f, f, f, vp]
        L.raster_oracle_joint.restype = i
        L.raster_oracle_joint.argtypes = [vp, f, f, vp, C.POINTER(i)]
        L.raster_oracle_to_pixel.restype = i
        L.raster_oracle_to_pixel.argtypes = [f, f, i]
        L.raster_oracle_render.restype = i
        L.raster_oracle_render.argtypes = [vp, vp, vp, i, i, f, f, i, i, f, f, vp, vp]
        L.raster_oracle_overlap.restype = i
        L.raster_oracle_overlap.argtypes = [vp, vp, i, i, f, f, C.POINTER(C.c_double)]
        L.raster_oracle_parents_diff.restype = C.c_int64
        L.raster_oracle_parents_diff.argtypes = [vp, vp, C.c_int64, vp, C.c_uint32, vp, f, vp]
        _LIB = L
    return _LIB


def hull(p_xy, r, delta=np.pi / 4, eps=0.01, cap=2048):
    p = np.ascontiguousarray(p_xy, dtype=np.float32)
    out = np.zeros((cap, 2), np.float32)
    n = lib().raster_oracle_hull(p.ctypes.data, len(p), r, delta, eps, out.ctypes.data, cap)
    if n < 0:
        raise ValueError(f"CalcError: lines should be intersecting ({n})")
    return out[:n].copy()


def arc(p, r, a0, a1, delta):
    out = np.zeros((512, 2), np.float32)
    n = lib().raster_oracle_arc(p[0], p[1], r, a0, a1, delta, out.ctypes.data)
    return out[:n].copy()


def joint(p1, p2, p3, r, delta):
    p = np.ascontiguousarray([p1, p2, p3], dtype=np.float32)
    out = np.zeros((512, 2), np.float32)
    right = C.c_int(0)
    n = lib().raster_oracle_joint(p.ctypes.data, r, delta, out.ctypes.data, C.byref(right))
    return bool(right.value), out[:n].copy()


def render(pos, radii, colors, domain_size, resolution, delta=np.pi / 4, eps=0.01):
    pos = np.ascontiguousarray(pos, dtype=np.float32)
    radii = np.ascontiguousarray(radii, dtype=np.float32)
    colors = np.ascontiguousarray(colors, dtype=np.uint8)
    rx, ry = int(resolution[0]), int(resolution[1])
    mask = np.zeros((ry, rx, 3), np.uint8)
    over = np.zeros((ry, rx, 3), np.uint8)
    rc = lib().raster_oracle_render(pos.ctypes.data, radii.ctypes.data, colors.ctypes.data, len(pos), pos.shape[1],
                                    domain_size[0], domain_size[1], rx, ry, delta, eps, mask.ctypes.data,
                                    over.ctypes.data)
    if rc < 0:
        raise ValueError(f"CalcError: outline of rod {-1 - rc}")
    return mask, over


def overlap(pos, radii, delta=np.pi / 4, eps=0.01) -> float:
    pos = np.ascontiguousarray(pos, dtype=np.float32)
    radii = np.ascontiguousarray(radii, dtype=np.float32)
    a = C.c_double(0)
    rc = lib().raster_oracle_overlap(pos.ctypes.data, radii.ctypes.data, len(pos), pos.shape[1], delta, eps, C.byref(a))
    if rc < 0:
        raise ValueError(f"CalcError: outline of rod {-1 - rc}")
    return float(a.value)


def parents_diff(m1, m2, cell_of_counter, parent_of, penalty=0.5):
    m1 = np.ascontiguousarray(m1, dtype=np.uint8)
    m2 = np.ascontiguousarray(m2, dtype=np.uint8)
    coc = np.ascontiguousarray(cell_of_counter, dtype=np.int32)
    par = np.ascontiguousarray(parent_of, dtype=np.int32)
    out = np.zeros(m1.shape[:2], np.float32)
    rc = lib().raster_oracle_parents_diff(m1.ctypes.data, m2.ctypes.data, out.size, coc.ctypes.data, len(coc),
                                          par.ctypes.data, penalty, out.ctypes.data)
    if rc < 0:
        raise KeyError(f"could not find color of pixel {-1 - rc}")
    return out
