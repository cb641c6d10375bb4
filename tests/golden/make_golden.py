"""Golden vectors for the hot path.

1. `generate_positions_seed1.npy`: the reference's OWN golden vector, transcribed from the Rust test
   `backwards_compat_generate_positions` (reference src/simulation.rs:168-219): positions of
   `_generate_positions(4, default, default, seed=1, dx=[0,0], randomize_positions=0.1, 8)`.
2. `readme_positions_seed0.npy`: input of the README / C1 case (seed 0, no jitter) produced by
   oracle/generate_positions.py, which (1) pins.
The reference cannot be imported in this container (Rust extension, no toolchain), so no other
reference-generated outputs exist; see oracle/crm_oracle.c header ("parity unpinned").
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

SEED1 = [
    [[15.782119, 16.658249, 1.4922986], [18.387316, 18.2603, 1.4922986], [21.448421, 19.527, 1.4922986],
     [24.20771, 21.220747, 1.4922986], [26.298336, 22.355158, 1.4922986], [29.717592, 23.55539, 1.4922986],
     [29.781338, 25.25969, 1.4922986], [32.75929, 28.03757, 1.4922986]],
    [[4.2639103, 71.299194, 1.4667264], [7.1084356, 71.53958, 1.4667264], [10.703307, 71.7947, 1.4667264],
     [13.803044, 72.10863, 1.4667264], [15.5561495, 72.22104, 1.4667264], [18.084274, 72.68125, 1.4667264],
     [20.477375, 72.87153, 1.4667264], [23.981491, 73.02278, 1.4667264]],
    [[68.69818, 30.033642, 1.384213], [71.653015, 29.911337, 1.384213], [72.82722, 28.599209, 1.384213],
     [74.60274, 26.107859, 1.384213], [80.05429, 26.160421, 1.384213], [84.007164, 24.251816, 1.384213],
     [83.433105, 23.3395, 1.384213], [90.09657, 18.543726, 1.384213]],
    [[89.117294, 63.976006, 1.4164526], [87.471756, 66.64026, 1.4164526], [84.89454, 67.19988, 1.4164526],
     [84.3147, 70.79428, 1.4164526], [80.57336, 70.544975, 1.4164526], [79.02444, 75.27408, 1.4164526],
     [78.291405, 79.85086, 1.4164526], [74.64508, 82.48188, 1.4164526]],
]

if __name__ == "__main__":
    from oracle.generate_positions import generate_positions

    here = Path(__file__).resolve().parent
    np.save(here / "generate_positions_seed1.npy", np.array(SEED1, dtype=np.float32))
    p = np.stack(generate_positions(4, 3.0, (100.0, 100.0), 2.5, rng_seed=0))
    np.save(here / "readme_positions_seed0.npy", p)
