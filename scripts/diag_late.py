"""C3 colony stepped N times, then a few more list steps: run under ncu with -s/-c to capture a late list step."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench
from cr_mech_coli_b200 import _capi

lib = _capi.product()
w = bench.make_workload("c3", 1_000_000, 0, 1)
sim = _capi.Sim(lib, w["cfg"], 0)
sim.set_agents(w["pos"], w["vel"], w["par"])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1100
sim.step(n)
print("stats", sim.list_stats(), sim.list_totals(), flush=True)
import torch
torch.cuda.init()
torch.cuda.cudart().cudaProfilerStart()  # ncu --profile-from-start off: only the steps below are captured
sim.step(4)
torch.cuda.cudart().cudaProfilerStop()
print("launches", sim.launch_count())
sim.close()
