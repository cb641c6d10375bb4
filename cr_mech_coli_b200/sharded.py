"""Voxel-slab sharding of one colony over several GPUs.

The reference splits its voxel grid into ``Configuration.n_threads`` sub-domains that exchange
boundary-voxel agents and returned forces every step (reference src/simulation.rs:251-256,283-292).
Here every GPU owns the rods whose centroid x lies in its slab ``[lo, hi)``:

    per step   pack halo records -> send/recv to the two slab neighbours (NCCL over NVLink)
               -> ghosts -> binning + force + integrate on the owned rods
               -> all_gather {divisions, leavers} -> division identifiers / migration if any

Forces are gathered by the owner, so no force travels back (one exchange instead of the
reference's two).  Identifiers, division order and snapshots are keyed by the *global slot* (the
slot a rod has in the unsharded run), which makes lineage independent of the partition.

Host-side logic (``SlabPartition``, ``assign_division_ids``, ``merge_snapshots``) is plain numpy and
is exercised on CPU with ``gloo`` (tests/test_sharded_host.py); device work goes through the C ABI
(``crm_sim_halo_pack`` ... ``crm_sim_apply_divisions``), the transport through ``torch.distributed``.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _capi

HEADER_BYTES = 64  # message header: int32 n_records, float32 max_extent
# steps between two rebuilds run stream-ordered (no host wait between pack, transport, ghosts, step and
# the flag all_gather); CRM_B200_SYNC_HALO=1 selects the synchronous protocol for comparison
STREAM_ORDERED = os.environ.get("CRM_B200_SYNC_HALO", "0") != "1"
# the step loop between two rebuilds runs in the C++ host with NCCL called directly (crm_sim_steps_sharded: whole
# batches of steps incl. their exchanges, one status read-back per batch); CRM_B200_PY_COORD=1 keeps the
# per-step Python coordinator (torch.distributed transport) for comparison
CPP_COORDINATOR = os.environ.get("CRM_B200_PY_COORD", "0") != "1"


# --------------------------------------------------------------------------------------------
# host logic (numpy only)

class SlabPartition:
    """Ownership intervals along x: rank r owns centroid x in [edges[r], edges[r+1]); the outer
    edges are -inf/+inf so that every rod has an owner."""

    def __init__(self, inner_edges):
        inner = np.asarray(inner_edges, dtype=np.float32)
        if np.any(np.diff(inner) <= 0):
            raise ValueError("slab edges must increase")
        self.edges = np.concatenate([[-np.inf], inner, [np.inf]]).astype(np.float32)

    @property
    def world(self) -> int:
        return len(self.edges) - 1

    @classmethod
    def uniform(cls, domain_x: float, world: int) -> "SlabPartition":
        return cls(np.linspace(0.0, float(domain_x), world + 1)[1:-1])

    @classmethod
    def balanced(cls, centroid_x, world: int) -> "SlabPartition":
        """Equal rod counts per slab (weak scaling): edges at the quantiles of the centroids."""
        cx = np.sort(np.asarray(centroid_x, dtype=np.float64))
        if world == 1 or len(cx) == 0:
            return cls([])
        cuts = [(cx[min(len(cx) - 1, (len(cx) * r) // world)] + cx[max(0, (len(cx) * r) // world - 1)]) / 2
                for r in range(1, world)]
        return cls(np.maximum.accumulate(np.asarray(cuts)) + np.arange(world - 1) * 1e-3)

    def owner(self, centroid_x) -> np.ndarray:
        """Same rule as the device (``x >= lo && x < hi`` on the f32 centroid)."""
        cx = np.asarray(centroid_x, dtype=np.float32)
        return np.searchsorted(self.edges[1:-1], cx, side="right").astype(np.int64)

    def bounds(self, rank: int):
        return float(self.edges[rank]), float(self.edges[rank + 1])

    def min_width(self) -> float:
        inner = self.edges[1:-1]
        return float(np.min(np.diff(inner))) if len(inner) > 1 else float("inf")


def centroid_x(pos) -> np.ndarray:
    """f32 centroid x of every rod as the device computes it (sum of the vertices * 1/Nv)."""
    pos = np.asarray(pos, dtype=np.float32)
    return (pos[:, :, 0].sum(axis=1, dtype=np.float32) * np.float32(1.0 / pos.shape[1])).astype(np.float32)


def assign_division_ids(per_rank, vox_counter: np.ndarray, n_total: int):
    """Identifiers of one division step from the division lists of all slabs.

    ``per_rank[r] = (gslot_mother[], user_voxel[])``.  Mirrors the unsharded bookkeeping
    (csrc/crm_sim.cu handle_divisions; cellular_raza update_cell_cycle): mothers ordered by global
    slot; inside a voxel the mothers are renamed first (counter+1+k), then the new cells
    (counter+count+1+k); daughter k of the step gets global slot n_total+k.  Updates
    ``vox_counter`` in place and returns ``per_rank`` lists of (id_mother, id_daughter,
    gslot_daughter) in each rank's own order."""
    gs = np.concatenate([np.asarray(g, dtype=np.uint64) for g, _ in per_rank]) if per_rank else np.zeros(0, np.uint64)
    vx = np.concatenate([np.asarray(v, dtype=np.int64) for _, v in per_rank]) if per_rank else np.zeros(0, np.int64)
    order = np.argsort(gs, kind="stable")
    n = len(gs)
    rank_in_step = np.empty(n, dtype=np.int64)
    rank_in_step[order] = np.arange(n)
    per_vox = np.bincount(vx, minlength=len(vox_counter)).astype(np.uint64)
    k_in_vox = np.empty(n, dtype=np.uint64)
    seen = {}
    for i in order:
        v = int(vx[i])
        k_in_vox[i] = seen.get(v, 0)
        seen[v] = seen.get(v, 0) + 1
    idm = np.zeros(n, dtype=_capi.IDENT_DTYPE)
    idd = np.zeros(n, dtype=_capi.IDENT_DTYPE)
    idm["kind"] = idd["kind"] = _capi.ID_DIVISION
    idm["voxel"] = idd["voxel"] = vx.astype(np.uint64)
    base = vox_counter[vx].astype(np.uint64)
    idm["counter"] = base + np.uint64(1) + k_in_vox
    idd["counter"] = base + per_vox[vx] + np.uint64(1) + k_in_vox
    gsd = (np.uint64(n_total) + rank_in_step.astype(np.uint64)).astype(np.uint64)
    vox_counter += (2 * per_vox).astype(vox_counter.dtype)
    out, o = [], 0
    for g, _ in per_rank:
        m = len(g)
        out.append((idm[o:o + m].copy(), idd[o:o + m].copy(), gsd[o:o + m].copy()))
        o += m
    return out


def merge_snapshots(per_rank):
    """``per_rank[r]`` = list of (iteration, pos, vel, par, ids, parents, gslot) -> one list of
    (iteration, pos, vel, par, ids, parents) with the rods in global-slot order, i.e. exactly the
    slot order of the unsharded run."""
    n_snap = len(per_rank[0])
    out = []
    for k in range(n_snap):
        it = per_rank[0][k][0]
        assert all(p[k][0] == it for p in per_rank), "ranks disagree on the save iterations"
        parts = [p[k] for p in per_rank if len(p[k][6])]
        if not parts:
            parts = [per_rank[0][k]]
        gs = np.concatenate([p[6] for p in parts])
        order = np.argsort(gs, kind="stable")
        out.append((it,) + tuple(np.concatenate([p[j] for p in parts])[order] for j in range(1, 6)))
    return out


# --------------------------------------------------------------------------------------------
# one slab on one device

class Shard:
    """Device side of one slab: a ``crm_sim`` in sharded mode plus its exchange buffers."""

    def __init__(self, lib, cfg, device: int, rank: int, partition: SlabPartition, halo_cap: int = 0):
        import torch

        self.torch = torch
        self.lib, self.cfg, self.rank, self.partition = lib, cfg, rank, partition
        self.device = torch.device("cuda", device)
        self.sim = _capi.Sim(lib, cfg, device)
        lo, hi = partition.bounds(rank)
        lib.check(lib.fn("sim_set_slab")(self.sim.h, lo, hi))
        self.halo_cap = halo_cap
        self.msg = None
        self.emax_nb = 0.0
        # the sim runs on one of torch's pool streams (they are never destroyed, so tensors recorded on it
        # cannot outlive it): transport enqueued under it is ordered with the kernels
        self.stream = torch.cuda.Stream(device=self.device)
        lib.check(lib.fn("sim_set_stream")(self.sim.h, C.c_void_p(self.stream.cuda_stream)))
        self.flags = torch.zeros(4, dtype=torch.int32, device=self.device)  # [divisions, rebuild wish, -error, 0]
        self.n_send = [0, 0]
        self.n_recv = [0, 0]

    def close(self):
        self.torch.cuda.synchronize(self.device)
        self.sim.close()

    # -- setup ---------------------------------------------------------------
    def set_agents(self, pos, vel, par, gslots):
        L, h = self.lib, self.sim.h
        self.sim.set_agents(pos, vel, par)
        gs = np.ascontiguousarray(gslots, dtype=np.uint64)
        L.check(L.fn("sim_set_global_slots")(h, gs.ctypes.data_as(C.c_void_p), len(gs)))
        self.rb = int(L.fn("sim_halo_record_bytes")(h))
        self.mb = int(L.fn("sim_migrant_record_bytes")(h))
        if not self.halo_cap:
            self.halo_cap = max(4096, len(gs) // 8)
        t = self.torch
        self.send = [t.zeros(HEADER_BYTES + self.halo_cap * self.rb, dtype=t.uint8, device=self.device) for _ in range(2)]
        self.recv = [t.zeros_like(self.send[0]) for _ in range(2)]
        self.selfbuf = t.zeros(self.halo_cap * self.rb, dtype=t.uint8, device=self.device)
        t.cuda.synchronize(self.device)  # (filled on torch's current stream, used on the sim's)
        self.n_self = 0
        self.n_leave = 0
        g = np.asarray(par["growth_rate"]) if len(gs) else np.zeros(0)
        thr = np.asarray(par["spring_length_threshold"]) if len(gs) else np.zeros(0)
        self.can_divide = bool(np.any((g != 0) & np.isfinite(thr) & (thr < 1e29)))

    @property
    def n_rods(self):
        return self.sim.n_rods

    def max_extent(self) -> float:
        return float(self.lib.fn("sim_max_extent")(self.sim.h))

    def halo_width(self, emax_global: float) -> float:
        """A rod interacts with rods whose centroid is at most ext_a + ext_b + range away."""
        r = float(self.lib.fn("sim_search_radius")(self.sim.h))  # includes the skin of the persistent lists
        skin = float(self.lib.fn("sim_skin")(self.sim.h))  # frozen halos: rods may drift skin/2 each before a rebuild
        return (self.max_extent() + emax_global) * 1.02 + r * 1.02 + skin + 1e-3

    # -- halo ----------------------------------------------------------------
    def begin_step(self, emax_nb: float, rebuild: bool = True):
        """Pack both halo messages.  On a rebuild step the rods that left the slab are handed over
        first (their halo records stay here as ghosts for this step) and the halo bands are
        selected afresh; otherwise the frozen selection of the last rebuild is re-sent and the
        persistent neighbour structure is kept.  Returns the migrant records (host bytes) for the
        lower / upper neighbour."""
        L, h = self.lib, self.sim.h
        out = (np.zeros(0, np.uint8), np.zeros(0, np.uint8))
        self.n_self = 0
        if not rebuild:
            L.check(L.fn("sim_halo_pack_msgs")(h, 0.0, C.c_void_p(self.send[0].data_ptr()) if self.rank > 0 else None,
                                               C.c_void_p(self.send[1].data_ptr()) if self.rank + 1 < self.partition.world else None,
                                               self.halo_cap, 0, 0, None, None))
            L.fn("sim_keep_lists")(h)
            return out
        L.check(L.fn("sim_refresh_bounds")(h))  # centroids / extents of the current positions
        if self.n_leave:
            n = C.c_uint32(0)
            L.check(L.fn("sim_pack_leavers")(h, C.c_void_p(self.selfbuf.data_ptr()), self.halo_cap, C.byref(n)))
            self.n_self = int(n.value)
            out = self.migrate_out()
            self.n_leave = 0
        flags = (1 if len(out[0]) else 0) | (2 if len(out[1]) else 0)
        width = self.halo_width(max(emax_nb, self.max_extent()))
        lo = C.c_void_p(self.send[0].data_ptr()) if self.rank > 0 else None
        hi = C.c_void_p(self.send[1].data_ptr()) if self.rank + 1 < self.partition.world else None
        n_lo, n_hi = C.c_uint32(0), C.c_uint32(0)
        L.check(L.fn("sim_halo_pack_msgs")(h, width, lo, hi, self.halo_cap, flags, 1, C.byref(n_lo), C.byref(n_hi)))
        self.n_send = [int(n_lo.value) if lo else 0, int(n_hi.value) if hi else 0]
        return out

    def _msg_ptrs(self, bufs):
        return (C.c_void_p(bufs[0].data_ptr()) if self.rank > 0 else None,
                C.c_void_p(bufs[1].data_ptr()) if self.rank + 1 < self.partition.world else None)

    def frozen_pack(self):
        """Stream-ordered half of a step between rebuilds: the frozen selections packed again (no
        host wait).  Returns the message slices to send / receive (the counts are those of the last
        rebuild step on both ends)."""
        L, h = self.lib, self.sim.h
        L.check(L.fn("sim_halo_repack")(h, *self._msg_ptrs(self.send)))
        self.n_self = 0
        cut = lambda bufs, n: [b[:HEADER_BYTES + k * self.rb] if (side == 0 and self.rank > 0) or
                               (side == 1 and self.rank + 1 < self.partition.world) else b[:0]
                               for side, (b, k) in enumerate(zip(bufs, n))]
        return cut(self.send, self.n_send), cut(self.recv, self.n_recv)

    def frozen_ghosts(self):
        self.lib.check(self.lib.fn("sim_ghosts_refresh")(self.sim.h, *self._msg_ptrs(self.recv)))

    def keep_lists(self):
        """All slabs agreed after the last step that nothing moved far enough: reuse the structure."""
        self.lib.fn("sim_keep_lists")(self.sim.h)

    def step_begin(self):
        """Enqueue the step; ``self.flags`` (device int32[2]) receives [divisions, reason to rebuild]."""
        self.lib.check(self.lib.fn("sim_step_begin")(self.sim.h, C.c_void_p(self.flags.data_ptr())))

    def step_end(self):
        self.lib.check(self.lib.fn("sim_step_end")(self.sim.h))
        return self._step_result()

    def set_comm(self, comm):
        """Steps with the (job-wide) NCCL communicator of all slabs."""
        self.lib.check(self.lib.fn("sim_set_comm")(self.sim.h, comm))

    def steps_sharded(self, n_max: int, world: int):
        """Up to ``n_max`` list steps incl. their halo exchanges, enqueued by the C++ host (collective).  Returns
        (steps done, flags[world, 4], SimulationError of this slab or None)."""
        done = C.c_uint32(0)
        flags = np.zeros(4 * world, dtype=np.uint32)
        send, recv = self._msg_ptrs(self.send), self._msg_ptrs(self.recv)
        rc = self.lib.fn("sim_steps_sharded")(self.sim.h, int(n_max), send[0], send[1], recv[0], recv[1], C.byref(done),
                                              flags.ctypes.data_as(C.c_void_p))
        err = None
        if rc < 0:
            kind = _capi.STATUS_NAMES.get(rc, f"error {rc}")
            err = _capi.SimulationError(f"{kind}: {self.lib.last_error()}")
            if rc not in (_capi.ERR_BOUNDARY, _capi.ERR_NUMERIC, _capi.ERR_DIVISION):
                raise err  # not a model error: the transport or the device failed
        self.n_self = 0
        return int(done.value), flags.reshape(world, 4).astype(np.float64), err

    def install_ghosts(self):
        """Ghosts of this step from the received messages + the own leavers.  Returns (counts,
        incoming-migrant flags per side, largest extent reported by the neighbours)."""
        L, h = self.lib, self.sim.h
        lo = C.c_void_p(self.recv[0].data_ptr()) if self.rank > 0 else None
        hi = C.c_void_p(self.recv[1].data_ptr()) if self.rank + 1 < self.partition.world else None
        counts = (C.c_uint32 * 2)()
        flags = (C.c_uint32 * 2)()
        emax = C.c_float(0)
        L.check(L.fn("sim_set_ghosts_msgs")(h, lo, hi, C.c_void_p(self.selfbuf.data_ptr()), self.n_self, counts, flags,
                                            C.byref(emax)))
        # the lower neighbour announces migrants for us in bit 1 (its upper side), the upper one in bit 0
        incoming = (bool(flags[0] & 2), bool(flags[1] & 1))
        self.n_recv = [int(counts[0]), int(counts[1])]
        return (int(counts[0]), int(counts[1])), incoming, float(emax.value)

    # -- step ----------------------------------------------------------------
    def step_local(self):
        """One local step; returns (pending divisions, 1 if this slab has a reason to rebuild)."""
        self.sim.step(1)
        return self._step_result()

    def _step_result(self):
        n_div = int(self.lib.fn("sim_pending_divisions")(self.sim.h, None, None, 0))
        self.n_leave = int(self.lib.fn("sim_n_leaving")(self.sim.h))
        return n_div, int(self.lib.fn("sim_wants_rebuild")(self.sim.h))

    def pending_divisions(self):
        L, h = self.lib, self.sim.h
        n = int(L.fn("sim_pending_divisions")(h, None, None, 0))
        gs = np.zeros(n, dtype=np.uint64)
        vx = np.zeros(n, dtype=np.uint32)
        L.fn("sim_pending_divisions")(h, gs.ctypes.data_as(C.c_void_p), vx.ctypes.data_as(C.c_void_p), n)
        return gs, vx

    def apply_divisions(self, idm, idd, gsd):
        L = self.lib
        idm, idd = np.ascontiguousarray(idm), np.ascontiguousarray(idd)
        gsd = np.ascontiguousarray(gsd, dtype=np.uint64)
        L.check(L.fn("sim_apply_divisions")(self.sim.h, idm.ctypes.data_as(C.c_void_p), idd.ctypes.data_as(C.c_void_p),
                                            gsd.ctypes.data_as(C.c_void_p), len(gsd)))

    # -- migration -------------------------------------------------------------
    def migrate_out(self, cap: int = 4096):
        """Returns two uint8 host arrays (records for the lower / upper neighbour)."""
        L, h = self.lib, self.sim.h
        bufs = [np.zeros(cap * self.mb, dtype=np.uint8) for _ in range(2)]
        n_lo, n_hi = C.c_uint32(0), C.c_uint32(0)
        L.check(L.fn("sim_migrate_out")(h, bufs[0].ctypes.data_as(C.c_void_p), cap, C.byref(n_lo),
                                        bufs[1].ctypes.data_as(C.c_void_p), cap, C.byref(n_hi)))
        return bufs[0][:n_lo.value * self.mb], bufs[1][:n_hi.value * self.mb]

    def migrate_in(self, records: np.ndarray):
        n = len(records) // self.mb
        if n:
            rec = np.ascontiguousarray(records)
            self.lib.check(self.lib.fn("sim_migrate_in")(self.sim.h, rec.ctypes.data_as(C.c_void_p), n))

    # -- results ---------------------------------------------------------------
    def take_snapshot(self):
        self.lib.check(self.lib.fn("sim_take_snapshot")(self.sim.h))

    def snapshots(self):
        out = []
        for k, s in enumerate(self.sim.snapshots()):
            gs = np.zeros(len(s[1]), dtype=np.uint64)
            self.lib.check(self.lib.fn("sim_snapshot_global_slots")(self.sim.h, k, gs.ctypes.data_as(C.c_void_p)))
            out.append(s + (gs,))
        return out

    def global_slots(self):
        gs = np.zeros(self.n_rods, dtype=np.uint64)
        self.lib.fn("sim_get_global_slots")(self.sim.h, gs.ctypes.data_as(C.c_void_p))
        return gs


def split_agents(partition: SlabPartition, pos, vel, par):
    """Per-rank (pos, vel, par, gslots) of an agent list given in global-slot order."""
    owner = partition.owner(centroid_x(pos)) if len(pos) else np.zeros(0, np.int64)
    out = []
    for r in range(partition.world):
        idx = np.nonzero(owner == r)[0]
        out.append((pos[idx], vel[idx], par[idx], idx.astype(np.uint64)))
    return out


# --------------------------------------------------------------------------------------------
# drivers

class _Coordinator:
    """State every rank replicates: per-voxel identifier counters and the global rod count."""

    def __init__(self, cfg, n_total: int):
        self.vox_counter = np.zeros(int(cfg.n_voxels[0]) * int(cfg.n_voxels[1]), dtype=np.uint64)
        self.n_total = int(n_total)

    def divisions(self, per_rank):
        res = assign_division_ids(per_rank, self.vox_counter, self.n_total)
        self.n_total += sum(len(g) for g, _ in per_rank)
        return res


class LocalCluster:
    """All slabs in ONE process (several sims on one or more devices); the transport is a tensor
    copy.  Used by the tests to exercise halo exchange, migration and division bookkeeping on a
    single GPU, and as the reference for the distributed driver's protocol."""

    def __init__(self, lib, cfg, partition: SlabPartition, devices=None, halo_cap: int = 0):
        w = partition.world
        devices = devices or [0] * w
        self.cfg, self.partition = cfg, partition
        self.shards = [Shard(lib, cfg, devices[r], r, partition, halo_cap) for r in range(w)]
        self.coord = None
        self.iteration = 0

    def close(self):
        for s in self.shards:
            s.close()

    def set_agents(self, pos, vel, par):
        for s, part in zip(self.shards, split_agents(self.partition, pos, vel, par)):
            s.set_agents(*part)
        self.coord = _Coordinator(self.cfg, len(pos))
        self.emax_nb = [max(s.max_extent() for s in self.shards)] * len(self.shards)
        self.rebuild_next = True

    def _exchange_halos(self):
        w = len(self.shards)
        rebuild = self.rebuild_next
        outs = [s.begin_step(self.emax_nb[r], rebuild) for r, s in enumerate(self.shards)]
        # ownership changed hands: the hand-over ghosts exist for this step only, rebuild again next step
        self.moved = any(len(o[0]) or len(o[1]) for o in outs)
        for r, s in enumerate(self.shards):
            if r > 0:
                s.recv[0].copy_(self.shards[r - 1].send[1])
            if r + 1 < w:
                s.recv[1].copy_(self.shards[r + 1].send[0])
        self.shards[0].torch.cuda.synchronize()
        for r, s in enumerate(self.shards):
            counts, incoming, emax = s.install_ghosts()
            self.emax_nb[r] = emax
            got = False
            if r > 0 and incoming[0]:
                assert len(outs[r - 1][1])
                s.migrate_in(outs[r - 1][1])
                got = True
            if r + 1 < w and incoming[1]:
                assert len(outs[r + 1][0])
                s.migrate_in(outs[r + 1][0])
                got = True
            if got:  # the new owned rods took the first ghost slots
                s.install_ghosts()

    def _frozen_step(self):
        """A step between two rebuilds through the stream-ordered entry points (the protocol the
        distributed driver runs without host waits; here the copies need the streams joined)."""
        t, w = self.shards[0].torch, len(self.shards)
        cuts = [s.frozen_pack() for s in self.shards]
        for s in self.shards:
            s.stream.synchronize()
        for r in range(w):
            if r > 0:
                cuts[r][1][0].copy_(cuts[r - 1][0][1])
            if r + 1 < w:
                cuts[r][1][1].copy_(cuts[r + 1][0][0])
        t.cuda.synchronize()
        for s in self.shards:
            s.frozen_ghosts()
            s.keep_lists()
            s.step_begin()
        status = [s.step_end() for s in self.shards]
        for s, st in zip(self.shards, status):  # the device-written flags are what the ranks all_gather
            assert s.flags.tolist()[:3] == [st[0], int(bool(st[1])), 0], (s.flags.tolist(), st)
        self.moved = False
        return status

    def step(self, n_steps: int = 1):
        for _ in range(n_steps):
            if STREAM_ORDERED and not self.rebuild_next:
                status = self._frozen_step()
            else:
                self._exchange_halos()
                status = [s.step_local() for s in self.shards]
            self.iteration += 1
            self.rebuild_next = any(w for _, w in status) or self.moved  # rebuilds are collective
            if any(nd for nd, _ in status):
                res = self.coord.divisions([s.pending_divisions() for s in self.shards])
                for s, r in zip(self.shards, res):
                    s.apply_divisions(*r)

    def compute_forces(self):
        """Forces / neighbour counts of the current state in global-slot order."""
        self.rebuild_next = True
        self._exchange_halos()
        parts = [s.sim.compute_forces() + (s.global_slots(),) for s in self.shards]
        gs = np.concatenate([p[2] for p in parts])
        order = np.argsort(gs, kind="stable")
        return np.concatenate([p[0] for p in parts])[order], np.concatenate([p[1] for p in parts])[order]

    def run(self):
        its, n_it = self.shards[0].lib.save_schedule(self.cfg)
        saves = set(int(i) for i in its)
        if 0 in saves:
            for s in self.shards:
                s.take_snapshot()
        while self.iteration < n_it:
            self.step(1)
            if self.iteration in saves:
                for s in self.shards:
                    s.take_snapshot()
        return merge_snapshots([s.snapshots() for s in self.shards])

    def get_state(self):
        parts = [s.sim.get_state() + (s.global_slots(),) for s in self.shards]
        order = np.argsort(np.concatenate([p[5] for p in parts]), kind="stable")
        return tuple(np.concatenate([p[j] for p in parts])[order] for j in range(5))


class Transport:
    """Neighbour send/recv and small collectives over ``torch.distributed`` (NCCL for cuda tensors,
    gloo for the CPU tests of the host protocol)."""

    def __init__(self, device, group=None):
        import torch
        import torch.distributed as dist

        self.torch, self.dist, self.group, self.device = torch, dist, group, torch.device(device)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def neighbours(self):
        return (self.rank - 1 if self.rank > 0 else None, self.rank + 1 if self.rank + 1 < self.world else None)

    def sendrecv(self, send, recv, wait: bool = True):
        """send[0]/recv[0] <-> lower neighbour, send[1]/recv[1] <-> upper neighbour; empty tensors
        are skipped (both sides know the sizes).  ``wait=False`` (NCCL): the transfers are only
        ordered on the current stream, the host does not wait."""
        dist, ops = self.dist, []
        for side, nb in enumerate(self.neighbours()):
            if nb is None:
                continue
            if send[side].numel():
                ops.append(dist.P2POp(dist.isend, send[side], nb, group=self.group))
            if recv[side].numel():
                ops.append(dist.P2POp(dist.irecv, recv[side], nb, group=self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            if self.device.type == "cuda" and wait:
                self.torch.cuda.current_stream(self.device).synchronize()

    def allgather(self, values) -> np.ndarray:
        t = self.torch.tensor(values, dtype=self.torch.float64, device=self.device)
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t, group=self.group)
        return self.torch.stack(out).cpu().numpy()

    def gather_divisions(self, gs, vx, counts):
        """Division lists (global slot of the mother, user voxel) of all ranks."""
        t = self.torch
        m = int(max(counts))
        buf = np.zeros((m, 2), dtype=np.int64)
        buf[:len(gs), 0] = np.asarray(gs).astype(np.int64)
        buf[:len(gs), 1] = vx
        mine = t.from_numpy(buf).to(self.device)
        out = [t.zeros_like(mine) for _ in range(self.world)]
        self.dist.all_gather(out, mine, group=self.group)
        out = [o.cpu().numpy() for o in out]
        return [(out[r][:int(counts[r]), 0].astype(np.uint64), out[r][:int(counts[r]), 1]) for r in range(self.world)]

    def exchange_records(self, out, record_bytes: int, sides=(True, True)):
        """Variable-length byte records with the neighbours on the given sides (both ends of a
        link must pass the same flag): counts first, then the payload."""
        t = self.torch
        e = t.zeros(0, dtype=t.int64, device=self.device)
        cs = [t.tensor([len(o) // record_bytes], dtype=t.int64, device=self.device) if sd else e for o, sd in zip(out, sides)]
        cr = [t.zeros(1, dtype=t.int64, device=self.device) if sd else e for sd in sides]
        self.sendrecv(cs, cr)
        e8 = t.zeros(0, dtype=t.uint8, device=self.device)
        send = [t.from_numpy(np.ascontiguousarray(o)).to(self.device) if sd else e8 for o, sd in zip(out, sides)]
        recv = [t.zeros(int(c.item()) * record_bytes, dtype=t.uint8, device=self.device) if sd else e8 for c, sd in zip(cr, sides)]
        self.sendrecv(send, recv)
        return [r.cpu().numpy() for r in recv]


_COMMS: dict = {}


def _job_comm(lib, tr: "Transport", device: int, group):
    """The NCCL communicator of the C++ step loop for this process group, created once per job (NCCL connects
    lazily on the first transfers: hundreds of milliseconds).  The unique id travels over the process group."""
    key = (id(group), int(device), tr.world)
    if key not in _COMMS:
        uid = [None]
        if tr.rank == 0:
            b = (C.c_char * 128)()
            lib.check(lib.fn("comm_unique_id")(b))
            uid[0] = bytes(b.raw)
        tr.dist.broadcast_object_list(uid, src=tr.dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        comm = C.c_void_p()
        buf = (C.c_char * 128).from_buffer_copy(uid[0])
        lib.check(lib.fn("comm_create")(buf, tr.rank, tr.world, int(device), C.byref(comm)))
        _COMMS[key] = comm
    return _COMMS[key]


class DistributedShard:
    """One slab per process; transport = ``torch.distributed`` (NCCL on GPUs).  Every rank calls
    the same methods in the same order."""

    def __init__(self, lib, cfg, partition: SlabPartition, device: int, halo_cap: int = 0, group=None):
        import torch

        self.tr = Transport(torch.device("cuda", device), group)
        self.rank, self.world = self.tr.rank, self.tr.world
        if partition.world != self.world:
            raise ValueError("partition and process group disagree on the number of slabs")
        self.cfg, self.partition = cfg, partition
        self.shard = Shard(lib, cfg, device, self.rank, partition, halo_cap)
        self.iteration = 0
        self.coord = None
        self.emax_global = 0.0
        self._flags_all = self._flags_host = None
        self._halo_ahead = False
        self.cpp = False
        if CPP_COORDINATOR and self.tr.device.type == "cuda" and self.world > 1:
            self.shard.set_comm(_job_comm(lib, self.tr, device, group))
            self.cpp = True

    def close(self):
        self.shard.close()

    def set_agents_local(self, pos, vel, par, gslots, n_total: int):
        self.shard.set_agents(pos, vel, par, gslots)
        self.coord = _Coordinator(self.cfg, n_total)
        st = self.tr.allgather([self.shard.max_extent(), float(self.shard.can_divide)])
        self.emax_global = float(st[:, 0].max())
        self.can_divide = bool(st[:, 1].max() > 0)
        self.use_lists = float(self.shard.lib.fn("sim_skin")(self.shard.sim.h)) > 0.0
        self.rebuild_next = True
        need = 2.0 * self.shard.halo_width(self.emax_global)
        if self.partition.min_width() < need:
            raise ValueError(f"slabs narrower than two halo widths ({need:.1f} um): use fewer GPUs")

    def set_agents(self, pos, vel, par):
        part = split_agents(self.partition, pos, vel, par)[self.rank]
        self.set_agents_local(*part, n_total=len(pos))

    def _exchange_halos(self):
        s = self.shard
        out = s.begin_step(self.emax_global, self.rebuild_next)
        self.tr.sendrecv(s.send, s.recv)
        counts, incoming, emax = s.install_ghosts()
        self.last_counts = counts
        self.emax_global = max(emax, s.max_extent())  # (refreshed on rebuild steps)
        # migrants travel only between neighbours that announced them in the message headers
        sides = [bool(len(out[k])) or incoming[k] for k in (0, 1)]
        self.moved = any(sides)  # the hand-over ghosts exist for this step only: rebuild again next step
        if any(sides):
            got = self.tr.exchange_records(out, s.mb, sides)
            for r in got:
                s.migrate_in(r)
            if any(len(r) for r in got):
                s.install_ghosts()

    def _frozen_step(self):
        """A step between two rebuilds, enqueued on the sim's stream end to end: pack -> NCCL
        send/recv -> ghost scatter -> force + integrate -> all_gather of the device-written
        [divisions, rebuild] flags -> one host wait."""
        s, t = self.shard, self.tr.torch
        if self._flags_all is None:
            self._flags_all = t.zeros(4 * self.world, dtype=t.int32, device=s.device)
            self._flags_host = t.zeros(4 * self.world, dtype=t.int32).pin_memory()
            self._flags_event = t.cuda.Event()
        with t.cuda.stream(s.stream):
            if not self._halo_ahead:
                self._frozen_halo()
            s.keep_lists()
            s.step_begin()
            self.tr.dist.all_gather_into_tensor(self._flags_all, s.flags, group=self.tr.group)
            self._flags_host.copy_(self._flags_all, non_blocking=True)
            self._flags_event.record(s.stream)
            # the halo of the NEXT step, enqueued while this one runs: it is what the next step needs if
            # no slab asks for a rebuild (the common case), and harmless otherwise (it only touches the
            # exchange buffers and ghost slots, which a rebuild step fills again)
            self._frozen_halo()
            self._halo_ahead = True
        # the host waits for the flags only: the exchange above is still running when the next step's
        # kernels are enqueued behind it, so the device never waits for the host
        self._flags_event.synchronize()
        st = self._flags_host.numpy().reshape(self.world, 4).astype(np.float64)
        if st[:, 2].max() > 0:  # a BoundaryError / CalcError in some slab ends the run in ALL slabs
            s.stream.synchronize()
            self._halo_ahead = False
            self._raise_collective(st[:, 2])
        s.step_end()
        self.moved = False
        if st[:, 1].max() > 0 or st[:, 0].sum() > 0:
            s.stream.synchronize()  # the synchronous rebuild / division protocol follows: drain the stream first
        return st

    def _raise_collective(self, codes, own: Exception | None = None):
        """Every rank raises the SimulationError of the run (the reference reports one error for the whole
        simulation, src/simulation.rs:237): the owner of the failing rod with the position dump, the
        others naming the rank."""
        if own is None and codes[self.rank] > 0:
            try:
                self.shard.step_end()
            except _capi.SimulationError as e:
                own = e
        if own is not None:
            raise own
        r = int(np.argmax(codes > 0))
        kind = _capi.STATUS_NAMES.get(-int(codes[r]), f"error {-int(codes[r])}")
        raise _capi.SimulationError(f"{kind}: raised in slab {r} at iteration {self.iteration + 1}")

    def _frozen_halo(self):
        s = self.shard
        send, recv = s.frozen_pack()
        self.tr.sendrecv(send, recv, wait=False)
        s.frozen_ghosts()

    def step(self, n_steps: int = 1):
        """``n_steps`` collective steps.  The cyclic garbage collector is paused meanwhile: the ranks run in
        lock step, so a collection on any of them is a stall of all of them."""
        import gc
        was_on = gc.isenabled()
        gc.disable()
        try:
            self._steps(n_steps)
        finally:
            if was_on:
                gc.enable()

    def _steps(self, n_steps: int):
        remaining = int(n_steps)
        while remaining > 0:
            done = 1
            if self.cpp and STREAM_ORDERED and self.use_lists and not self.rebuild_next:
                # between two rebuilds: whole batches of steps from the C++ host, one host wait per batch
                with self.tr.torch.cuda.stream(self.shard.stream):
                    done, st, own = self.shard.steps_sharded(remaining, self.world)
                self._halo_ahead = False
                self.moved = False
                if st[:, 2].max() > 0:
                    self.iteration += done - 1
                    self._raise_collective(st[:, 2], own)
                self.shard._step_result()
                self.iteration += done
            elif STREAM_ORDERED and self.use_lists and not self.rebuild_next:
                st = self._frozen_step()
                self.iteration += 1
            else:
                self._halo_ahead = False
                self._exchange_halos()
                own, code = None, 0
                try:
                    n_div, want = self.shard.step_local()
                except _capi.SimulationError as e:  # the other slabs must learn of it before anyone raises
                    own, n_div, want = e, 0, 1
                    code = next((-k for k, v in _capi.STATUS_NAMES.items() if str(e).startswith(v)), 1)
                # one small all_gather per step: divisions need the lists of all slabs for the
                # identifiers, the persistent neighbour structure is rebuilt in all slabs at once, and
                # an error in one slab ends the run in all of them
                st = self.tr.allgather([n_div, int(want or self.moved), code])
                if st[:, 2].max() > 0:
                    self._raise_collective(st[:, 2], own)
                self.iteration += 1
            remaining -= done
            self.rebuild_next = bool(st[:, 1].max() > 0) or not self.use_lists
            if st[:, 0].sum() > 0:
                gs, vx = self.shard.pending_divisions()
                res = self.coord.divisions(self.tr.gather_divisions(gs, vx, st[:, 0].astype(np.int64)))
                self.shard.apply_divisions(*res[self.rank])

    def run(self):
        """Steps to t_max; returns the merged snapshots on every rank."""
        its, n_it = self.shard.lib.save_schedule(self.cfg)
        saves = set(int(i) for i in its)
        if 0 in saves:
            self.shard.take_snapshot()
        while self.iteration < n_it:
            self.step(1)
            if self.iteration in saves:
                self.shard.take_snapshot()
        mine = self.shard.snapshots()
        gathered = [None] * self.world
        self.tr.dist.all_gather_object(gathered, mine, group=self.tr.group)
        return merge_snapshots(gathered)
