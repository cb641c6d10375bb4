"""Host-side logic of the voxel-slab sharding: partition, division identifiers across
slabs, snapshot merge, and the torch.distributed transport on CPU with gloo (world_size 2 and 3).
No GPU and no compute call through the C ABI here; the device half is covered by
tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest

from cr_mech_coli_b200 import _capi, sharded


def test_partition_owner_rule():
    p = sharded.SlabPartition.uniform(100.0, 4)
    assert p.world == 4
    assert list(p.owner([0.0, 24.999, 25.0, 74.9, 75.0, 99.0, 150.0, -3.0])) == [0, 0, 1, 2, 3, 3, 3, 0]
    assert p.bounds(0) == (-np.inf, 25.0) and p.bounds(3) == (75.0, np.inf)
    assert p.min_width() == 25.0
    with pytest.raises(ValueError):
        sharded.SlabPartition([10.0, 5.0])


def test_partition_balanced_and_split():
    rng = np.random.default_rng(0)
    pos = rng.uniform(0, 1000, (5000, 8, 3)).astype(np.float32)
    cx = sharded.centroid_x(pos)
    p = sharded.SlabPartition.balanced(cx, 4)
    counts = np.bincount(p.owner(cx), minlength=4)
    assert counts.sum() == 5000 and counts.max() - counts.min() <= 2
    par = np.zeros(5000, _capi.ROD_PARAMS_DTYPE)
    parts = sharded.split_agents(p, pos, np.zeros_like(pos), par)
    gs = np.concatenate([q[3] for q in parts])
    assert sorted(gs.tolist()) == list(range(5000))  # every rod has exactly one owner
    for r, q in enumerate(parts):
        lo, hi = p.bounds(r)
        c = sharded.centroid_x(q[0])
        assert ((c >= lo) & (c < hi)).all()


def _unsharded_ids(gslots, voxels, vox_counter, n_total):
    """The bookkeeping of csrc/crm_sim.cu handle_divisions on one list."""
    order = np.argsort(gslots, kind="stable")
    per_vox = np.bincount(voxels, minlength=len(vox_counter))
    rank_vox = np.zeros(len(vox_counter), int)
    out = {}
    for r, i in enumerate(order):
        v = voxels[i]
        k = rank_vox[v]
        rank_vox[v] += 1
        out[int(gslots[i])] = ((v, int(vox_counter[v]) + 1 + k), (v, int(vox_counter[v]) + per_vox[v] + 1 + k), n_total + r)
    vox_counter += 2 * per_vox.astype(vox_counter.dtype)
    return out


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_division_ids_independent_of_partition(world):
    rng = np.random.default_rng(world)
    vc_a = np.zeros(9, np.uint64)
    vc_b = np.zeros(9, np.uint64)
    n_total = 100
    for step in range(5):
        n = int(rng.integers(0, 30))
        gs = rng.choice(n_total, size=n, replace=False).astype(np.uint64)
        vx = rng.integers(0, 9, n)
        owner = rng.integers(0, world, n)
        per_rank = [(gs[owner == r], vx[owner == r]) for r in range(world)]
        want = _unsharded_ids(gs, vx, vc_a, n_total)
        got = sharded.assign_division_ids(per_rank, vc_b, n_total)
        for (g, v), (idm, idd, gsd) in zip(per_rank, got):
            for i in range(len(g)):
                w = want[int(g[i])]
                assert (int(idm["voxel"][i]), int(idm["counter"][i])) == w[0]
                assert (int(idd["voxel"][i]), int(idd["counter"][i])) == w[1]
                assert int(gsd[i]) == w[2]
                assert idm["kind"][i] == idd["kind"][i] == _capi.ID_DIVISION
        assert (vc_a == vc_b).all()
        n_total += n


def test_merge_snapshots_orders_by_global_slot():
    def snap(it, gs):
        gs = np.asarray(gs, np.uint64)
        n = len(gs)
        pos = np.repeat(gs.astype(np.float32), 24).reshape(n, 8, 3)
        ids = np.zeros(n, _capi.IDENT_DTYPE)
        ids["counter"] = gs
        return (it, pos, pos * 2, np.zeros(n, _capi.ROD_PARAMS_DTYPE), ids, ids.copy(), gs)

    merged = sharded.merge_snapshots([[snap(0, [3, 0]), snap(5, [3, 0, 6])], [snap(0, [1, 2]), snap(5, [])],
                                      [snap(0, []), snap(5, [4, 1, 2, 5])]])
    assert [m[0] for m in merged] == [0, 5]
    assert merged[0][4]["counter"].tolist() == [0, 1, 2, 3]
    assert merged[1][4]["counter"].tolist() == [0, 1, 2, 3, 4, 5, 6]
    assert merged[1][1][:, 0, 0].tolist() == [0, 1, 2, 3, 4, 5, 6]


# ---- torch.distributed transport on CPU (gloo) -------------------------------------------------
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tr = sharded.Transport("cpu")
        assert tr.neighbours() == (rank - 1 if rank else None, rank + 1 if rank + 1 < world else None)
        # fixed-size halo messages: [towards lower, towards upper]
        send = [torch.full((96,), 10 * rank + side, dtype=torch.uint8) for side in (0, 1)]
        recv = [torch.full((96,), 255, dtype=torch.uint8) for _ in (0, 1)]
        tr.sendrecv(send, recv)
        if rank > 0:
            assert (recv[0] == 10 * (rank - 1) + 1).all()  # the lower neighbour's upward message
        else:
            assert (recv[0] == 255).all()
        if rank + 1 < world:
            assert (recv[1] == 10 * (rank + 1)).all()
        else:
            assert (recv[1] == 255).all()
        # variable-length migrant records
        rb = 8
        out = [np.full(rb * (rank + 1), rank, np.uint8) if rank > 0 else np.zeros(0, np.uint8),
               np.full(rb * 2, 100 + rank, np.uint8) if rank + 1 < world else np.zeros(0, np.uint8)]
        got = tr.exchange_records(out, rb)
        assert len(got[0]) == (rb * 2 if rank > 0 else 0) and (got[0] == 100 + rank - 1).all()
        assert len(got[1]) == (rb * (rank + 2) if rank + 1 < world else 0) and (got[1] == rank + 1).all()
        # status all_gather + division coordination: every rank derives the same identifiers
        cfg = _capi.make_config(n_voxels=(2, 2))
        coord = sharded._Coordinator(cfg, 50)
        rng = np.random.default_rng(7)  # same stream everywhere: the "global truth"
        for step in range(3):
            n = int(rng.integers(1, 12))
            gs = rng.choice(coord.n_total, n, replace=False).astype(np.uint64)
            vx = rng.integers(0, 4, n)
            own = rng.integers(0, world, n)
            mine = own == rank
            st = tr.allgather([int(mine.sum()), 0, 1.5 * rank])
            assert st.shape == (world, 3) and st[:, 0].sum() == n and st[:, 2].max() == 1.5 * (world - 1)
            per_rank = tr.gather_divisions(gs[mine], vx[mine], st[:, 0].astype(np.int64))
            for r in range(world):
                assert (per_rank[r][0] == gs[own == r]).all() and (per_rank[r][1] == vx[own == r]).all()
            res = coord.divisions(per_rank)
            ref = sharded.assign_division_ids([(gs, vx)], np.array(q["vc"][step], np.uint64), q["nt"][step])[0]
            sel = np.nonzero(mine)[0]
            assert (res[rank][0] == ref[0][sel]).all() and (res[rank][1] == ref[1][sel]).all()
            assert (res[rank][2] == ref[2][sel]).all()
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_transport_and_division_protocol_gloo(world):
    import torch.multiprocessing as mp

    # the expected (unsharded) counters before each step, computed once here
    cfg = _capi.make_config(n_voxels=(2, 2))
    rng = np.random.default_rng(7)
    vc, nt = np.zeros(4, np.uint64), 50
    q = {"vc": [], "nt": []}
    for step in range(3):
        q["vc"].append(vc.copy().tolist())
        q["nt"].append(nt)
        n = int(rng.integers(1, 12))
        gs = rng.choice(nt, n, replace=False).astype(np.uint64)
        vx = rng.integers(0, 4, n)
        rng.integers(0, world, n)
        sharded.assign_division_ids([(gs, vx)], vc, nt)
        nt += n
    mp.spawn(_worker, args=(world, _free_port(), q), nprocs=world, join=True)
