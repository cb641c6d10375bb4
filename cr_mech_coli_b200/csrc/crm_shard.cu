// crm_shard.cu — voxel-slab sharding of a colony over several GPUs: the device and
// host halves of the per-step halo exchange and of the migration of rods that change owner.
//
// The reference decomposes the voxel grid into `n_threads` sub-domains that exchange the positions
// of boundary-voxel agents and the returned forces every step (cellular_raza chili backend,
// reference src/simulation.rs:251-256,283-292).  Here a slab of the domain, x in [lo, hi), is owned
// by one GPU; a rod belongs to the slab that holds its centroid.  Because forces are gathered by
// the owner, a neighbour only needs the *halo records* of the rods near the common edge (vertex
// records, bounds, potential: (LP + 5) * 16 B per rod) and sends no force back.  The transport
// between GPUs (NCCL send/recv) lives above the C ABI (cr_mech_coli_b200/sharded.py); this file
// packs and unpacks the device buffers it moves.
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include "crm_sim.h"

using namespace crm;

namespace {

struct InBand {  // owned rods whose centroid x lies in [lo, hi)
    const float4* bs;
    float lo, hi;
    __device__ bool operator()(uint32_t a) const {
        const float x = bs[a].x;
        return x >= lo && x < hi;
    }
};

// record i = [vrec[LP] | bsphere | chord0 | chord1 | pot(2 x float4)] of slot sel[i]
__global__ void k_halo_gather(Dev d, const uint32_t* __restrict__ sel, uint32_t n_sel, int lp, float4* __restrict__ out) {
    const int rq = lp + 5;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i = t / rq;
    const int q = t % rq;
    if (i >= n_sel) return;
    const uint32_t a = sel[i];
    float4 v;
    if (q < lp) v = d.vrec[(size_t)a * lp + q];
    else if (q == lp) v = d.bsphere[a];
    else if (q == lp + 1) v = d.chord[2 * a];
    else if (q == lp + 2) v = d.chord[2 * a + 1];
    else v = reinterpret_cast<const float4*>(d.pot + a)[q - lp - 3];
    out[(size_t)i * rq + q] = v;
}

// message header (first 64 B of an exchange buffer): {n_records, max_extent bits, flags}
__global__ void k_write_header(uint32_t* hdr, const uint32_t* count, float emax, uint32_t flags) {
    hdr[0] = *count;
    hdr[1] = __float_as_uint(emax);
    hdr[2] = flags;
}

// inverse: record i of `in` becomes slot first + i
__global__ void k_ghost_scatter(Dev d, const float4* __restrict__ in, uint32_t n_rec, uint32_t first, int lp) {
    const int rq = lp + 5;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i = t / rq;
    const int q = t % rq;
    if (i >= n_rec) return;
    const uint32_t a = first + i;
    const float4 v = in[(size_t)i * rq + q];
    if (q < lp) d.vrec[(size_t)a * lp + q] = v;
    else if (q == lp) d.bsphere[a] = v;
    else if (q == lp + 1) d.chord[2 * a] = v;
    else if (q == lp + 2) d.chord[2 * a + 1] = v;
    else reinterpret_cast<float4*>(d.pot + a)[q - lp - 3] = v;
}

// ---- one launch per direction of a sharded step between rebuilds (crm_sim_steps_sharded) ---------------------
// pack: both halo messages (headers + records of the frozen selections, [count | slots] lists) and this slab's
// flags {divisions, rebuild wish, -error, 0} of the step that has just been integrated
__global__ void k_halo_pack_all(Dev d, const uint32_t* __restrict__ sel_lo, uint32_t n_lo, const uint32_t* __restrict__ sel_hi,
                                uint32_t n_hi, int lp, uint32_t* msg_lo, uint32_t* msg_hi, float emax, float skin,
                                uint32_t* flags) {
    const int rq = lp + 5;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) {
        if (msg_lo) { msg_lo[0] = sel_lo[0]; msg_lo[1] = __float_as_uint(emax); msg_lo[2] = 0u; }
        if (msg_hi) { msg_hi[0] = sel_hi[0]; msg_hi[1] = __float_as_uint(emax); msg_hi[2] = 0u; }
        const Status* st = d.status;  // (the same conditions step_conclude evaluates on the host)
        const float disp = sqrtf(__uint_as_float(st->disp2_bits));
        uint32_t want = (st->n_div > 0) | (st->n_leave > 0) | (st->err_code != 0);
        if (skin > 0.0f) want |= (st->plist_overflow != 0) | !(disp < 0.4999f * skin);
        flags[0] = st->n_div;
        flags[1] = want ? 1u : 0u;
        flags[2] = (uint32_t)(-st->err_code);
        flags[3] = 0u;
    }
    const uint32_t i = t / rq;
    const int q = t % rq;
    if (i >= n_lo + n_hi) return;
    const bool lo = i < n_lo;
    const uint32_t k = lo ? i : i - n_lo;
    const uint32_t a = (lo ? sel_lo : sel_hi)[1 + k];
    float4 v;
    if (q < lp) v = d.vrec[(size_t)a * lp + q];
    else if (q == lp) v = d.bsphere[a];
    else if (q == lp + 1) v = d.chord[2 * a];
    else if (q == lp + 2) v = d.chord[2 * a + 1];
    else v = reinterpret_cast<const float4*>(d.pot + a)[q - lp - 3];
    reinterpret_cast<float4*>((lo ? msg_lo : msg_hi) + 16)[(size_t)k * rq + q] = v;
}

// unpack: both received messages scattered to the ghost slots of the last rebuild step, and the gathered flags of
// the PREVIOUS step turned into the batch's latch (first step in which any slab wants the host)
__global__ void k_ghost_unpack_all(Dev d, const uint32_t* __restrict__ msg_lo, uint32_t n_lo, const uint32_t* __restrict__ msg_hi,
                                   uint32_t n_hi, uint32_t first, int lp, const uint32_t* __restrict__ flags_all, int world,
                                   uint32_t* latch, uint32_t iteration_prev) {
    const int rq = lp + 5;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) {
        uint32_t any = 0u;
        for (int r = 0; r < world; ++r) any |= flags_all[4 * r] | flags_all[4 * r + 1] | flags_all[4 * r + 2];
        if (any && *latch == 0u) *latch = iteration_prev;
    }
    const uint32_t i = t / rq;
    const int q = t % rq;
    if (i >= n_lo + n_hi) return;
    const bool lo = i < n_lo;
    const uint32_t k = lo ? i : i - n_lo;
    const uint32_t a = first + i;  // lower neighbour's ghosts first, then the upper one's
    const float4 v = reinterpret_cast<const float4*>((lo ? msg_lo : msg_hi) + 16)[(size_t)k * rq + q];
    if (q < lp) d.vrec[(size_t)a * lp + q] = v;
    else if (q == lp) d.bsphere[a] = v;
    else if (q == lp + 1) d.chord[2 * a] = v;
    else if (q == lp + 2) d.chord[2 * a + 1] = v;
    else reinterpret_cast<float4*>(d.pot + a)[q - lp - 3] = v;
}

// full device state of a rod as 32-bit words: six [3][LP] matrices, then pot, mech, grow, cyc, nn, nhist
__host__ __device__ inline int state_words(int lp) { return 18 * lp + 8 + 4 + 4 + 12 + 2; }

__global__ void k_state_gather(Dev d, const uint32_t* __restrict__ sel, uint32_t n_sel, int lp, uint32_t* __restrict__ out) {
    const int W = state_words(lp), M = 3 * lp;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i = t / W;
    const int q = t % W;
    if (i >= n_sel) return;
    const uint32_t a = sel[i];
    uint32_t v;
    if (q < 6 * M) {
        const float* src[6] = {d.pos, d.vel, d.dx1, d.dv1, d.dx2, d.dv2};
        v = __float_as_uint(src[q / M][(size_t)a * M + q % M]);
    } else {
        const int r = q - 6 * M;
        if (r < 8) v = reinterpret_cast<const uint32_t*>(d.pot + a)[r];
        else if (r < 12) v = reinterpret_cast<const uint32_t*>(d.mech + a)[r - 8];
        else if (r < 16) v = reinterpret_cast<const uint32_t*>(d.grow + a)[r - 12];
        else if (r < 28) v = reinterpret_cast<const uint32_t*>(d.cyc + a)[r - 16];
        else if (r == 28) v = d.nn[a];
        else v = d.nhist[a];
    }
    out[(size_t)i * W + q] = v;
}

__global__ void k_state_scatter(Dev d, const uint32_t* __restrict__ in, uint32_t n_rec, uint32_t first, int lp) {
    const int W = state_words(lp), M = 3 * lp;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i = t / W;
    const int q = t % W;
    if (i >= n_rec) return;
    const uint32_t a = first + i;
    const uint32_t v = in[(size_t)i * W + q];
    if (q < 6 * M) {
        float* dst[6] = {d.pos, d.vel, d.dx1, d.dv1, d.dx2, d.dv2};
        dst[q / M][(size_t)a * M + q % M] = __uint_as_float(v);
    } else {
        const int r = q - 6 * M;
        if (r < 8) reinterpret_cast<uint32_t*>(d.pot + a)[r] = v;
        else if (r < 12) reinterpret_cast<uint32_t*>(d.mech + a)[r - 8] = v;
        else if (r < 16) reinterpret_cast<uint32_t*>(d.grow + a)[r - 12] = v;
        else if (r < 28) reinterpret_cast<uint32_t*>(d.cyc + a)[r - 16] = v;
        else if (r == 28) d.nn[a] = v;
        else d.nhist[a] = (uint8_t)v;
    }
}

// slot moves of the compaction after rods left: state of slot src[i] -> slot dst[i] (disjoint sets)
__global__ void k_state_move(Dev d, const uint32_t* __restrict__ src, const uint32_t* __restrict__ dst, uint32_t n_mv, int lp) {
    const int W = state_words(lp), M = 3 * lp;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i = t / W;
    const int q = t % W;
    if (i >= n_mv) return;
    const uint32_t a = src[i], b = dst[i];
    if (q < 6 * M) {
        float* arr[6] = {d.pos, d.vel, d.dx1, d.dv1, d.dx2, d.dv2};
        arr[q / M][(size_t)b * M + q % M] = arr[q / M][(size_t)a * M + q % M];
    } else {
        const int r = q - 6 * M;
        if (r < 8) reinterpret_cast<uint32_t*>(d.pot + b)[r] = reinterpret_cast<const uint32_t*>(d.pot + a)[r];
        else if (r < 12) reinterpret_cast<uint32_t*>(d.mech + b)[r - 8] = reinterpret_cast<const uint32_t*>(d.mech + a)[r - 8];
        else if (r < 16) reinterpret_cast<uint32_t*>(d.grow + b)[r - 12] = reinterpret_cast<const uint32_t*>(d.grow + a)[r - 12];
        else if (r < 28) reinterpret_cast<uint32_t*>(d.cyc + b)[r - 16] = reinterpret_cast<const uint32_t*>(d.cyc + a)[r - 16];
        else if (r == 28) d.nn[b] = d.nn[a];
        else d.nhist[b] = d.nhist[a];
    }
}

struct HostTail {  // host half of a migrant record
    crm_rod_params base;
    uint32_t pad;
    crm_ident id, parent;
    uint64_t gslot;
};
static_assert(sizeof(HostTail) % 8 == 0, "migrant records stay 8-byte aligned");

// ordered (by slot) list of the owned rods with centroid x in [lo, hi) -> s->select_out; count -> *n
int select_band(crm_sim* s, float lo, float hi, uint32_t* n) {
    *n = 0;
    if (s->n == 0) return 0;
    if (!s->select_count) CK(cudaMalloc(&s->select_count, sizeof(uint32_t)));
    if (!s->select_out || s->select_cap < s->cap) {  // (re)plan for the current slot capacity
        if (s->select_out) CK(cudaFree(s->select_out));
        if (s->select_tmp) CK(cudaFree(s->select_tmp));
        s->select_out = nullptr;
        s->select_tmp = nullptr;
        CK(cudaMalloc(&s->select_out, sizeof(uint32_t) * s->cap));
        size_t bytes = 0;
        thrust::counting_iterator<uint32_t> it0(0);
        CK(cub::DeviceSelect::If(nullptr, bytes, it0, s->select_out, s->select_count, (int)s->cap,
                                 InBand{s->d.bsphere, lo, hi}, s->stream));
        CK(cudaMalloc(&s->select_tmp, bytes));
        s->select_tmp_bytes = bytes;
        s->select_cap = s->cap;
    }
    thrust::counting_iterator<uint32_t> it(0);
    size_t bytes = s->select_tmp_bytes;
    CK(cub::DeviceSelect::If(s->select_tmp, bytes, it, s->select_out, s->select_count, (int)s->n,
                             InBand{s->d.bsphere, lo, hi}, s->stream));
    CK(cudaMemcpyAsync(n, s->select_count, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    return 0;
}

}  // namespace

// launchers of the two fused kernels for crm_sim.cu
int crm_shard_pack_all(crm_sim* s, void* msg_lo, void* msg_hi, uint32_t* flags) {
    if (!s->halo_sel_cap) return crm_fail(CRM_ERR_INVALID, "no halo selection yet (crm_sim_halo_pack_msgs with reselect)");
    const uint32_t n_lo = msg_lo ? s->halo_sel_n[0] : 0, n_hi = msg_hi ? s->halo_sel_n[1] : 0;
    const uint64_t threads = std::max<uint64_t>(1, (uint64_t)(n_lo + n_hi) * (s->lp + 5));
    k_halo_pack_all<<<blocks_for(threads, 256), 256, 0, s->stream>>>(s->d, s->halo_sel[0], n_lo, s->halo_sel[1], n_hi, s->lp,
                                                                     (uint32_t*)msg_lo, (uint32_t*)msg_hi, s->emax, s->skin, flags);
    s->launches += 1;
    return 0;
}
int crm_shard_unpack_all(crm_sim* s, const void* msg_lo, const void* msg_hi, const uint32_t* flags_all, int world,
                         uint32_t* latch, uint32_t iteration_prev) {
    if (s->n_ghost != s->ghost_n[0] + s->ghost_n[1])
        return crm_fail(CRM_ERR_INVALID, "the ghosts of the last rebuild step include hand-over records: rebuild first");
    const uint32_t n_lo = msg_lo ? s->ghost_n[0] : 0, n_hi = msg_hi ? s->ghost_n[1] : 0;
    if ((s->ghost_n[0] && !msg_lo) || (s->ghost_n[1] && !msg_hi)) return crm_fail(CRM_ERR_INVALID, "missing halo message");
    const uint64_t threads = std::max<uint64_t>(1, (uint64_t)(n_lo + n_hi) * (s->lp + 5));
    k_ghost_unpack_all<<<blocks_for(threads, 256), 256, 0, s->stream>>>(s->d, (const uint32_t*)msg_lo, n_lo, (const uint32_t*)msg_hi,
                                                                        n_hi, s->n, s->lp, flags_all, world, latch, iteration_prev);
    s->launches += 1;
    return 0;
}

extern "C" {

int crm_sim_set_slab(crm_sim* s, float x_lo, float x_hi) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    if (!(x_lo < x_hi)) return crm_fail(CRM_ERR_INVALID, "empty slab [%g, %g)", x_lo, x_hi);
    if (s->n) return crm_fail(CRM_ERR_INVALID, "declare the slab before setting the agents");
    s->slab_lo = s->d.slab_lo = x_lo;
    s->slab_hi = s->d.slab_hi = x_hi;
    s->sharded = true;
    return 0;
}

int crm_sim_set_global_slots(crm_sim* s, const uint64_t* gslots, uint32_t n) {
    if (!s || n != s->n) return crm_fail(CRM_ERR_INVALID, "global slot list must cover the %u owned rods", s ? s->n : 0);
    for (uint32_t a = 0; a < n; ++a) {
        s->gslot[a] = gslots[a];
        if (s->ids[a].kind == CRM_ID_INITIAL) s->ids[a].counter = gslots[a];  // Initial(index in the input list)
    }
    CK(cudaSetDevice(s->device));
    s->rkey_on = true;
    return crm_upload_rkeys(s, 0, n);
}

uint32_t crm_sim_halo_record_bytes(const crm_sim* s) { return (uint32_t)(s->lp + 5) * 16u; }
uint32_t crm_sim_migrant_record_bytes(const crm_sim* s) {
    return (uint32_t)state_words(s->lp) * 4u + (uint32_t)sizeof(HostTail);
}
float crm_sim_max_extent(const crm_sim* s) { return s->emax; }
int crm_sim_refresh_bounds(crm_sim* s) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    CK(cudaSetDevice(s->device));
    return crm_refresh_bounds(s);
}
float crm_sim_search_radius(const crm_sim* s) { return s->rsearch + (s->skin > 0.0f ? s->skin : 0.0f); }
float crm_sim_skin(const crm_sim* s) { return s->skin; }
int crm_sim_wants_rebuild(const crm_sim* s) { return s->want_rebuild ? 1 : 0; }
// All slabs agreed that nothing moved far enough: the next step reuses the persistent structure.
int crm_sim_keep_lists(crm_sim* s) {
    if (s->have_lists && !s->want_rebuild) s->need_rebuild = false;
    return s->need_rebuild ? 0 : 1;
}

int crm_sim_halo_pack(crm_sim* s, int side, float width, void* dev_buf, uint32_t cap, uint32_t* n_out) {
    if (!s || !n_out) return crm_fail(CRM_ERR_INVALID, "null argument");
    CK(cudaSetDevice(s->device));
    {
        int rc = crm_refresh_bounds(s);
        if (rc) return rc;
    }
    uint32_t n = 0;
    const float lo = side == 0 ? -INFINITY : s->slab_hi - width;
    const float hi = side == 0 ? s->slab_lo + width : INFINITY;
    int rc = select_band(s, lo, hi, &n);
    if (rc) return rc;
    if (n > cap) return crm_fail(CRM_ERR_NOMEM, "halo of %u rods exceeds the exchange buffer (%u records)", n, cap);
    if (n) {
        const uint64_t threads = (uint64_t)n * (s->lp + 5);
        k_halo_gather<<<blocks_for(threads, 256), 256, 0, s->stream>>>(s->d, s->select_out, n, s->lp, (float4*)dev_buf);
        s->launches += 2;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(s->stream));  // the transport above the ABI runs on other streams
    }
    *n_out = n;
    return 0;
}

int crm_sim_set_ghosts(crm_sim* s, const void* dev_lo, uint32_t n_lo, const void* dev_hi, uint32_t n_hi,
                       float max_extent) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    if (s->pending_div) return crm_fail(CRM_ERR_INVALID, "divisions pending");
    CK(cudaSetDevice(s->device));
    int rc = crm_reserve(s, s->n + n_lo + n_hi);
    if (rc) return rc;
    if (n_lo) {
        const uint64_t threads = (uint64_t)n_lo * (s->lp + 5);
        k_ghost_scatter<<<blocks_for(threads, 256), 256, 0, s->stream>>>(s->d, (const float4*)dev_lo, n_lo, s->n, s->lp);
    }
    if (n_hi) {
        const uint64_t threads = (uint64_t)n_hi * (s->lp + 5);
        k_ghost_scatter<<<blocks_for(threads, 256), 256, 0, s->stream>>>(s->d, (const float4*)dev_hi, n_hi, s->n + n_lo, s->lp);
    }
    s->launches += (n_lo ? 1 : 0) + (n_hi ? 1 : 0);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s->stream));  // the receive buffers may be reused by the caller
    s->n_ghost = n_lo + n_hi;
    s->emax_ghost = max_extent;
    return 0;
}


// Both halo messages of a step in one call: [64-byte header {n, max_extent, flags} | records].
// One stream synchronisation; the counts come back for the caller's bookkeeping.
int crm_sim_halo_pack_msgs(crm_sim* s, float width, void* dev_msg_lo, void* dev_msg_hi, uint32_t cap,
                           uint32_t flags, int reselect, uint32_t* n_lo, uint32_t* n_hi) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null argument");
    CK(cudaSetDevice(s->device));
    void* msg[2] = {dev_msg_lo, dev_msg_hi};
    if (reselect || s->halo_sel_cap < cap) {
        int rc = crm_refresh_bounds(s);  // the bands are selected on the centroids; records carry the bounds
        if (rc) return rc;
    }
    if (s->halo_sel_cap < cap) {
        for (int side = 0; side < 2; ++side) {
            if (s->halo_sel[side]) CK(cudaFree(s->halo_sel[side]));
            CK(cudaMalloc(&s->halo_sel[side], sizeof(uint32_t) * (cap + 1)));  // [count | slots]
            s->halo_sel_n[side] = 0;
        }
        s->halo_sel_cap = cap;
        reselect = 1;
    }
    for (int side = 0; side < 2; ++side) {
        if (!msg[side]) continue;
        if (reselect) {  // otherwise the selection of the last rebuild stays frozen (ghost slots stay put)
            uint32_t n = 0;
            const float lo = side == 0 ? -INFINITY : s->slab_hi - width;
            const float hi = side == 0 ? s->slab_lo + width : INFINITY;
            int rc = select_band(s, lo, hi, &n);
            if (rc) return rc;
            if (n > cap) return crm_fail(CRM_ERR_NOMEM, "halo of %u rods exceeds the exchange buffer (%u records)", n, cap);
            s->halo_sel_n[side] = n;
            CK(cudaMemcpyAsync(s->halo_sel[side], &s->halo_sel_n[side], sizeof(uint32_t), cudaMemcpyHostToDevice, s->stream));
            if (n) CK(cudaMemcpyAsync(s->halo_sel[side] + 1, s->select_out, sizeof(uint32_t) * n, cudaMemcpyDeviceToDevice, s->stream));
        }
        const uint32_t n = s->halo_sel_n[side];
        k_write_header<<<1, 1, 0, s->stream>>>((uint32_t*)msg[side], s->halo_sel[side], s->emax, flags);
        if (n) {
            const uint64_t threads = (uint64_t)n * (s->lp + 5);
            k_halo_gather<<<blocks_for(threads, 256), 256, 0, s->stream>>>(
                s->d, s->halo_sel[side] + 1, n, s->lp, reinterpret_cast<float4*>((char*)msg[side] + 64));
        }
        s->launches += 2 + (reselect ? 1 : 0);
    }
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s->stream));  // the transport above the ABI runs on other streams
    if (n_lo) *n_lo = s->halo_sel_n[0];
    if (n_hi) *n_hi = s->halo_sel_n[1];
    return 0;
}

// The frozen selections of the last rebuild packed again, stream-ordered: nothing waits for the
// device, the caller's transport has to run on crm_sim_stream (or wait for it).
int crm_sim_halo_repack(crm_sim* s, void* dev_msg_lo, void* dev_msg_hi) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null argument");
    if (!s->halo_sel_cap) return crm_fail(CRM_ERR_INVALID, "no halo selection yet (crm_sim_halo_pack_msgs with reselect)");
    CK(cudaSetDevice(s->device));
    void* msg[2] = {dev_msg_lo, dev_msg_hi};
    for (int side = 0; side < 2; ++side) {
        if (!msg[side]) continue;
        const uint32_t n = s->halo_sel_n[side];
        k_write_header<<<1, 1, 0, s->stream>>>((uint32_t*)msg[side], s->halo_sel[side], s->emax, 0u);
        if (n) {
            const uint64_t threads = (uint64_t)n * (s->lp + 5);
            k_halo_gather<<<blocks_for(threads, 256), 256, 0, s->stream>>>(
                s->d, s->halo_sel[side] + 1, n, s->lp, reinterpret_cast<float4*>((char*)msg[side] + 64));
        }
        s->launches += n ? 2 : 1;
    }
    CK(cudaGetLastError());
    return 0;
}

// Halo records of the owned rods whose centroid is OUTSIDE the slab (about to migrate): the old
// owner keeps them as ghosts for the step in which the new owner has not announced them yet.
int crm_sim_pack_leavers(crm_sim* s, void* dev_buf, uint32_t cap, uint32_t* n_out) {
    if (!s || !n_out) return crm_fail(CRM_ERR_INVALID, "null argument");
    CK(cudaSetDevice(s->device));
    {
        int rc = crm_refresh_bounds(s);
        if (rc) return rc;
    }
    uint32_t total = 0;
    for (int side = 0; side < 2; ++side) {
        if ((side == 0 && !(s->slab_lo > -INFINITY)) || (side == 1 && !(s->slab_hi < INFINITY))) continue;
        uint32_t n = 0;
        int rc = select_band(s, side == 0 ? -INFINITY : s->slab_hi, side == 0 ? s->slab_lo : INFINITY, &n);
        if (rc) return rc;
        if (total + n > cap) return crm_fail(CRM_ERR_NOMEM, "%u leaving rods exceed the buffer", total + n);
        if (n) {
            const uint64_t threads = (uint64_t)n * (s->lp + 5);
            k_halo_gather<<<blocks_for(threads, 256), 256, 0, s->stream>>>(
                s->d, s->select_out, n, s->lp, reinterpret_cast<float4*>((char*)dev_buf + (size_t)total * (s->lp + 5) * 16));
            s->launches += 2;
        }
        total += n;
    }
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s->stream));
    *n_out = total;
    return 0;
}

// Install the ghosts of a step from the two received messages (headers read here) plus the
// caller's own leavers; returns the neighbours' counts, flags and the largest reported extent.
int crm_sim_set_ghosts_msgs(crm_sim* s, const void* dev_msg_lo, const void* dev_msg_hi, const void* dev_self,
                            uint32_t n_self, uint32_t* counts, uint32_t* flags, float* max_extent) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    if (s->pending_div) return crm_fail(CRM_ERR_INVALID, "divisions pending");
    CK(cudaSetDevice(s->device));
    uint32_t* h = reinterpret_cast<uint32_t*>(s->h_status + 1);  // pinned scratch behind the status block
    const void* msg[2] = {dev_msg_lo, dev_msg_hi};
    for (int side = 0; side < 2; ++side) {
        h[4 * side] = h[4 * side + 1] = h[4 * side + 2] = 0;
        if (msg[side]) CK(cudaMemcpyAsync(h + 4 * side, msg[side], 12, cudaMemcpyDeviceToHost, s->stream));
    }
    CK(cudaStreamSynchronize(s->stream));
    const uint32_t n_lo = h[0], n_hi = h[4];
    float emax = std::max(__builtin_bit_cast(float, h[1]), __builtin_bit_cast(float, h[5]));
    int rc = crm_reserve(s, s->n + n_lo + n_hi + n_self);
    if (rc) return rc;
    const uint32_t cnt[3] = {n_lo, n_hi, n_self};
    const void* src[3] = {msg[0] ? (const char*)msg[0] + 64 : nullptr, msg[1] ? (const char*)msg[1] + 64 : nullptr, dev_self};
    uint32_t first = s->n;
    for (int k = 0; k < 3; ++k) {
        if (!cnt[k]) continue;
        const uint64_t threads = (uint64_t)cnt[k] * (s->lp + 5);
        k_ghost_scatter<<<blocks_for(threads, 256), 256, 0, s->stream>>>(s->d, (const float4*)src[k], cnt[k], first, s->lp);
        first += cnt[k];
        s->launches += 1;
    }
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s->stream));  // the receive buffers may be reused by the caller
    s->n_ghost = n_lo + n_hi + n_self;
    s->ghost_n[0] = n_lo;
    s->ghost_n[1] = n_hi;
    if (n_self) emax = std::max(emax, s->emax);
    s->emax_ghost = emax;
    if (counts) { counts[0] = n_lo; counts[1] = n_hi; }
    if (flags) { flags[0] = h[2]; flags[1] = h[6]; }
    if (max_extent) *max_extent = emax;
    return 0;
}

// The ghosts of a step between two rebuilds: both neighbours re-sent their frozen selections, so
// the counts and ghost slots are those of the last crm_sim_set_ghosts_msgs.  Stream-ordered, the
// headers are not read (a neighbour that reselected would have announced a rebuild to all slabs).
int crm_sim_ghosts_refresh(crm_sim* s, const void* dev_msg_lo, const void* dev_msg_hi) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    if (s->pending_div) return crm_fail(CRM_ERR_INVALID, "divisions pending");
    if (s->n_ghost != s->ghost_n[0] + s->ghost_n[1])
        return crm_fail(CRM_ERR_INVALID, "the ghosts of the last rebuild step include hand-over records: rebuild first");
    if ((s->ghost_n[0] && !dev_msg_lo) || (s->ghost_n[1] && !dev_msg_hi)) return crm_fail(CRM_ERR_INVALID, "missing halo message");
    CK(cudaSetDevice(s->device));
    const void* msg[2] = {dev_msg_lo, dev_msg_hi};
    uint32_t first = s->n;
    for (int side = 0; side < 2; ++side) {
        const uint32_t n = s->ghost_n[side];
        if (!n) continue;
        const uint64_t threads = (uint64_t)n * (s->lp + 5);
        k_ghost_scatter<<<blocks_for(threads, 256), 256, 0, s->stream>>>(
            s->d, reinterpret_cast<const float4*>((const char*)msg[side] + 64), n, first, s->lp);
        first += n;
        s->launches += 1;
    }
    CK(cudaGetLastError());
    return 0;
}

// Rods whose centroid left the slab: their records go to the two host buffers (lower / upper
// neighbour), the remaining owned rods are compacted (the last rods fill the holes).
int crm_sim_migrate_out(crm_sim* s, void* host_lo, uint32_t cap_lo, uint32_t* n_lo_out, void* host_hi, uint32_t cap_hi,
                        uint32_t* n_hi_out) {
    if (!s || !n_lo_out || !n_hi_out) return crm_fail(CRM_ERR_INVALID, "null argument");
    if (s->pending_div) return crm_fail(CRM_ERR_INVALID, "divisions pending");
    CK(cudaSetDevice(s->device));
    *n_lo_out = *n_hi_out = 0;
    if (s->n == 0) return 0;
    {
        int rc = crm_refresh_bounds(s);
        if (rc) return rc;
    }
    const int W = state_words(s->lp);
    const size_t rec = (size_t)W * 4 + sizeof(HostTail);
    std::vector<uint32_t> leaving;  // slots, lower-bound leavers first
    uint32_t counts[2] = {0, 0};
    for (int side = 0; side < 2; ++side) {
        uint32_t n = 0;
        const float lo = side == 0 ? -INFINITY : s->slab_hi;
        const float hi = side == 0 ? s->slab_lo : INFINITY;
        if ((side == 0 && !(s->slab_lo > -INFINITY)) || (side == 1 && !(s->slab_hi < INFINITY))) continue;
        int rc = select_band(s, lo, hi, &n);
        if (rc) return rc;
        if (n > (side == 0 ? cap_lo : cap_hi))
            return crm_fail(CRM_ERR_NOMEM, "%u migrating rods exceed the exchange buffer", n);
        if (!n) continue;
        std::vector<uint32_t> sel(n);
        CK(cudaMemcpyAsync(sel.data(), s->select_out, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, s->stream));
        // device state -> staging -> host record
        uint32_t* stage = reinterpret_cast<uint32_t*>(s->stage);
        if ((size_t)n * W > (size_t)s->cap * s->nv * 3)
            return crm_fail(CRM_ERR_NOMEM, "too many migrating rods in one step (%u)", n);
        k_state_gather<<<blocks_for((uint64_t)n * W, 256), 256, 0, s->stream>>>(s->d, s->select_out, n, s->lp, stage);
        s->launches += 2;
        CK(cudaGetLastError());
        std::vector<uint32_t> words((size_t)n * W);
        CK(cudaMemcpyAsync(words.data(), stage, sizeof(uint32_t) * words.size(), cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        unsigned char* out = static_cast<unsigned char*>(side == 0 ? host_lo : host_hi);
        for (uint32_t i = 0; i < n; ++i) {
            memcpy(out + i * rec, words.data() + (size_t)i * W, (size_t)W * 4);
            HostTail t{};
            t.base = s->base[sel[i]];
            t.id = s->ids[sel[i]];
            t.parent = s->parents[sel[i]];
            t.gslot = s->gslot[sel[i]];
            memcpy(out + i * rec + (size_t)W * 4, &t, sizeof(t));
        }
        counts[side] = n;
        leaving.insert(leaving.end(), sel.begin(), sel.end());
    }
    *n_lo_out = counts[0];
    *n_hi_out = counts[1];
    const uint32_t n_rm = (uint32_t)leaving.size();
    if (!n_rm) return 0;
    // compaction: survivors of the tail [n - n_rm, n) fill the holes below n - n_rm
    const uint32_t n_new = s->n - n_rm;
    std::vector<uint8_t> gone(s->n, 0);
    for (uint32_t a : leaving) gone[a] = 1;
    std::vector<uint32_t> src, dst;
    uint32_t tail = n_new;
    for (uint32_t a = 0; a < n_new; ++a) {
        if (!gone[a]) continue;
        while (gone[tail]) ++tail;
        src.push_back(tail++);
        dst.push_back(a);
    }
    const uint32_t n_mv = (uint32_t)src.size();
    if (n_mv) {
        uint32_t* dsrc = s->select_out;  // reuse: capacity >= n
        uint32_t* ddst = s->d.cell_rank;  // scratch of the binning, rewritten every step
        CK(cudaMemcpyAsync(dsrc, src.data(), sizeof(uint32_t) * n_mv, cudaMemcpyHostToDevice, s->stream));
        CK(cudaMemcpyAsync(ddst, dst.data(), sizeof(uint32_t) * n_mv, cudaMemcpyHostToDevice, s->stream));
        k_state_move<<<blocks_for((uint64_t)n_mv * W, 256), 256, 0, s->stream>>>(s->d, dsrc, ddst, n_mv, s->lp);
        s->launches += 1;
        CK(cudaGetLastError());
        for (uint32_t i = 0; i < n_mv; ++i) {
            s->base[dst[i]] = s->base[src[i]];
            s->ids[dst[i]] = s->ids[src[i]];
            s->parents[dst[i]] = s->parents[src[i]];
            s->gslot[dst[i]] = s->gslot[src[i]];
        }
    }
    s->base.resize(n_new);
    s->ids.resize(n_new);
    s->parents.resize(n_new);
    s->gslot.resize(n_new);
    s->n = n_new;
    s->n_ghost = 0;
    s->need_rebuild = true;
    s->have_lists = false;
    s->d.n = s->d.n_own = n_new;
    {
        int rc = crm_upload_rkeys(s, 0, n_new);
        if (rc) return rc;
    }
    // bounds (vertex records, spheres, chords) of the moved slots
    CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
    launch_bounds_kernel(s->nv, s->d, s->stream);
    s->launches += 1;
    CK(cudaGetLastError());
    int rc = crm_read_status(s);
    if (rc) return rc;
    s->emax = __builtin_bit_cast(float, s->h_status->emax_bits);
    s->bounds_fresh = true;
    return 0;
}

int crm_sim_migrate_in(crm_sim* s, const void* host_buf, uint32_t n_in) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    if (s->pending_div) return crm_fail(CRM_ERR_INVALID, "divisions pending");
    if (!n_in) return 0;
    if (s->nv == 0) return crm_fail(CRM_ERR_INVALID, "set agents (possibly none) before migrating rods in");
    CK(cudaSetDevice(s->device));
    const int W = state_words(s->lp);
    const size_t rec = (size_t)W * 4 + sizeof(HostTail);
    int rc = crm_reserve(s, s->n + n_in);
    if (rc) return rc;
    if ((size_t)n_in * W > (size_t)s->cap * s->nv * 3)
        return crm_fail(CRM_ERR_NOMEM, "too many migrating rods in one step (%u)", n_in);
    std::vector<uint32_t> words((size_t)n_in * W);
    const unsigned char* in = static_cast<const unsigned char*>(host_buf);
    const uint32_t n0 = s->n;
    s->base.resize(n0 + n_in);
    s->ids.resize(n0 + n_in);
    s->parents.resize(n0 + n_in);
    s->gslot.resize(n0 + n_in);
    for (uint32_t i = 0; i < n_in; ++i) {
        memcpy(words.data() + (size_t)i * W, in + i * rec, (size_t)W * 4);
        HostTail t;
        memcpy(&t, in + i * rec + (size_t)W * 4, sizeof(t));
        s->base[n0 + i] = t.base;
        s->ids[n0 + i] = t.id;
        s->parents[n0 + i] = t.parent;
        s->gslot[n0 + i] = t.gslot;
    }
    uint32_t* stage = reinterpret_cast<uint32_t*>(s->stage);
    CK(cudaMemcpyAsync(stage, words.data(), sizeof(uint32_t) * words.size(), cudaMemcpyHostToDevice, s->stream));
    k_state_scatter<<<blocks_for((uint64_t)n_in * W, 256), 256, 0, s->stream>>>(s->d, stage, n_in, n0, s->lp);
    s->launches += 1;
    CK(cudaGetLastError());
    s->n = n0 + n_in;
    if ((rc = crm_upload_rkeys(s, n0, n_in))) return rc;
    s->n_ghost = 0;
    s->need_rebuild = true;
    s->have_lists = false;
    s->d.n = s->d.n_own = s->n;
    CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
    launch_bounds_kernel(s->nv, s->d, s->stream);
    s->launches += 1;
    CK(cudaGetLastError());
    if ((rc = crm_read_status(s))) return rc;
    s->emax = __builtin_bit_cast(float, s->h_status->emax_bits);
    s->bounds_fresh = true;
    return 0;
}

uint32_t crm_sim_n_leaving(const crm_sim* s) { return s->h_status ? s->h_status->n_leave : 0; }

uint32_t crm_sim_pending_divisions(const crm_sim* s, uint64_t* gslot_mother, uint32_t* user_voxel, uint32_t cap) {
    const uint32_t n = s->pending_div;
    for (uint32_t r = 0; r < n && r < cap; ++r) {
        if (gslot_mother) gslot_mother[r] = s->gslot[s->h_div[r].slot];
        if (user_voxel) user_voxel[r] = s->h_div[r].uvox;
    }
    return n;
}

int crm_sim_apply_divisions(crm_sim* s, const crm_ident* id_mother, const crm_ident* id_daughter,
                            const uint64_t* gslot_daughter, uint32_t n) {
    if (!s || n != s->pending_div) return crm_fail(CRM_ERR_INVALID, "expected the %u pending divisions", s ? s->pending_div : 0);
    if (!n) return 0;
    CK(cudaSetDevice(s->device));
    return crm_apply_divisions(s, id_mother, id_daughter, gslot_daughter);
}

int crm_sim_take_snapshot(crm_sim* s) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    CK(cudaSetDevice(s->device));
    return crm_take_snapshot(s);
}

int crm_sim_get_global_slots(const crm_sim* s, uint64_t* out) {
    memcpy(out, s->gslot.data(), sizeof(uint64_t) * s->n);
    return 0;
}

int crm_sim_snapshot_global_slots(const crm_sim* s, uint32_t k, uint64_t* out) {
    if (k >= s->snaps.size()) return crm_fail(CRM_ERR_INVALID, "snapshot %u out of range", k);
    memcpy(out, s->snaps[k].gslot.data(), sizeof(uint64_t) * s->snaps[k].n);
    return 0;
}

}  // extern "C"
