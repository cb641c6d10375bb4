"""torchrun worker of tests/test_gpu_sharded.py::test_two_process_nccl: a growing colony sharded
over WORLD_SIZE GPUs (NCCL), compared on rank 0 with the unsharded CUDA run."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))

import torch
import torch.distributed as dist

from cr_mech_coli_b200 import _capi, sharded
from cr_mech_coli_b200.testing import lattice_colony, morse_params


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl")
    lib = _capi.product()
    n, L = 64, 520.0
    cfg = _capi.make_config(domain_size=(L, L), n_voxels=(3, 3), t_max=25.0, dt=0.1, n_saves=4)
    pos = lattice_colony(n, spacing=(60.0, 40.0), margin=40.0, seed=11, angle_std=0.5)
    par = morse_params(n, growth_rate=0.08)
    par["growth_rate"] = np.random.default_rng(4).uniform(0.05, 0.1, n)
    vel = np.zeros_like(pos)
    vel[:, :, 0] = 0.4 * np.sign(L / 2 - pos[:, :1, 0])  # rods drift towards the middle edge
    ds = sharded.DistributedShard(lib, cfg, sharded.SlabPartition.uniform(L, world), local)
    ds.set_agents(pos, vel, par)
    snaps = ds.run()
    if rank == 0:
        with _capi.Sim(lib, cfg, local) as g:
            g.set_agents(pos, vel, par)
            g.run()
            ref = g.snapshots()
        assert len(snaps) == len(ref)
        for (it, ps, vs, pars, ids, prs), (_, pg, vg, parg, idg, prg) in zip(snaps, ref):
            assert ps.shape == pg.shape, (it, ps.shape, pg.shape)
            assert (ids == idg).all() and (prs == prg).all(), it
            assert (pars["spring_length"] == parg["spring_length"]).all()
            assert np.abs(ps - pg).max() < 1e-3, (it, np.abs(ps - pg).max())
        print(f"dist_worker ok: world {world}, final rods {len(snaps[-1][1])}")
    dist.barrier()
    ds.close()

    # a model error in ONE slab must end the run in ALL slabs (the reference reports one SimulationError,
    # src/simulation.rs:237), not leave the others waiting in a collective: a rod of the last slab flies out
    cfg2 = _capi.make_config(domain_size=(L, L), n_voxels=(3, 3), t_max=5.0, dt=0.1, n_saves=0)
    par2 = morse_params(n, growth_rate=0.0)
    par2["spring_length_threshold"] = 1e30
    vel2 = np.zeros_like(pos)
    far = int(np.argmax(pos[:, :, 0].mean(axis=1)))
    vel2[far, :, 0] = 1e4
    ds2 = sharded.DistributedShard(lib, cfg2, sharded.SlabPartition.uniform(L, world), local)
    ds2.set_agents(pos, vel2, par2)
    try:
        ds2.step(30)
        raised = None
    except _capi.SimulationError as e:
        raised = str(e)
    assert raised is not None and "BoundaryError" in raised, raised
    owner = world - 1
    assert ("Particle is out of domain" in raised) == (rank == owner), (rank, raised)
    print(f"dist_worker error propagation ok on rank {rank}: {raised[:60]}")
    dist.barrier()
    ds2.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
