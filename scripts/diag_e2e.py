"""Where the end-to-end time of bench.py goes (upload / steps / download), 1e6 rods."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench
from cr_mech_coli_b200 import _capi

lib = _capi.product()
cfg, pos, par = bench.workload(1_000_000, seed=12345)
vel = np.zeros_like(pos)
for rep in range(4):
    if rep == 2:  # page-locked caller buffers (crm_host_alloc)
        pos, vel, par = bench.pinned_copy(lib, pos), bench.pinned_copy(lib, vel), bench.pinned_copy(lib, par)
        outb = bench.state_buffers(lib, len(pos), pos.shape[1])
    sim = _capi.Sim(lib, cfg, 0)
    t0 = time.perf_counter(); sim.set_agents(pos, vel, par); t1 = time.perf_counter()
    sim.step(1); t2 = time.perf_counter()
    sim.step(19); t3 = time.perf_counter()
    out = sim.get_state(out=outb if rep >= 2 else None); t4 = time.perf_counter()
    print(f"rep {rep}: set_agents {t1-t0:.4f}  first step {t2-t1:.4f}  19 steps {t3-t2:.4f}  get_state {t4-t3:.4f}")
    sim.close()
