"""Mask path measured: crm_render_mask_2d on a lattice colony at several resolutions, the device time
of its kernels (outline + scanline raster + resolve, CUDA events inside the entry point) against the HBM roof, the
call through the C ABI with host buffers, and the CPU restatement (oracle/raster_oracle.c) beside it.

Algorithmic bytes per pixel: 8 B cleared + 8 B read by the resolve pass (owner / cover words) + 6 B of output
(two RGB bitmaps) = 22 B, plus 8 B of atomics per COVERED pixel (one atomicMax + one atomicAdd).  One JSON line
per resolution on stdout."""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from cr_mech_coli_b200 import fitting  # noqa: E402
from cr_mech_coli_b200.cell_container import counter_to_color  # noqa: E402
from cr_mech_coli_b200.testing import lattice_colony, lattice_domain  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6500.0))
    L = lattice_domain(n, spacing=(24.0, 4.0))
    pos = lattice_colony(n, seed=5, spacing=(24.0, 4.0))
    radii = np.full(n, 1.5, np.float32)
    colors = np.array([counter_to_color(k + 1) for k in range(n)], np.uint8)
    for res in (1024, 4096, 8192):
        fitting.render_mask_arrays(pos, radii, colors, (L, L), (res, res))  # warm-up
        ks, ws = [], []
        for _ in range(5):
            t0 = time.perf_counter()
            m, o, ms = fitting.render_mask_arrays(pos, radii, colors, (L, L), (res, res), return_ms=True)
            ws.append(time.perf_counter() - t0)
            ks.append(ms)
        covered = int((m.any(axis=2) | o.any(axis=2)).sum())
        npx = res * res
        alg = 22.0 * npx + 8.0 * covered
        k = float(np.median(ks))
        line = {"workload": f"render_mask_2d: {n} rods (lattice 24 x 4 um, r = 1.5) on a {L:.0f} um domain",
                "resolution": [res, res], "kernel_ms": k, "call_ms": float(np.median(ws)) * 1e3,
                "covered_pixels": covered, "rods_per_s_kernels": n / (k * 1e-3), "pixels_per_s_kernels": npx / (k * 1e-3),
                "roofline": {"bound": "hbm", "achieved": alg / (k * 1e-3) * 1e-9, "peak": hbm, "unit": "GB/s",
                             "frac": alg / (k * 1e-3) * 1e-9 / hbm, "algorithmic_bytes": alg,
                             "note": "outline + raster + resolve kernels together (CUDA events inside crm_render_mask_2d); "
                                     "the call time adds the copies of two bitmaps to pageable host memory"}}
        if res <= 4096:
            from oracle import raster

            t0 = time.perf_counter()
            mo, oo = raster.render(pos, radii, colors, (L, L), (res, res))
            line["cpu_oracle_ms"] = (time.perf_counter() - t0) * 1e3
            line["byte_exact_vs_oracle"] = bool(np.array_equal(m, mo) and np.array_equal(o, oo))
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
