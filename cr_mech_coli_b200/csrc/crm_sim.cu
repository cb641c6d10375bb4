// crm_sim.cu — host orchestration of the colony-mechanics path + the C ABI of
// include/crm_b200.h.  Replaces the run_main! driver of the reference
// (src/simulation.rs:234-319): save schedule, per-step kernel sequence, division
// bookkeeping (CellIdentifier / parent assignment), snapshots.
#include <chrono>
#include <mutex>

#include "crm_sim.h"

using namespace crm;

static thread_local std::string g_err;

// CRM_B200_TRACE=1: wall-clock phases of the upload / download entry points on stderr (diagnostics only)
static bool trace_on() {
    static const bool on = [] { const char* e = getenv("CRM_B200_TRACE"); return e && *e && *e != '0'; }();
    return on;
}
struct Trace {
    const char* what;
    std::chrono::steady_clock::time_point t0, t;
    std::string line;
    explicit Trace(const char* w) : what(w) { if (trace_on()) t0 = t = std::chrono::steady_clock::now(); }
    void mark(const char* phase) {
        if (!trace_on()) return;
        const auto now = std::chrono::steady_clock::now();
        char buf[96];
        snprintf(buf, sizeof buf, " %s %.2f", phase, std::chrono::duration<double, std::milli>(now - t).count());
        line += buf;
        t = now;
    }
    ~Trace() {
        if (!trace_on()) return;
        fprintf(stderr, "[crm trace] %s:%s | total %.2f ms\n", what, line.c_str(),
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
};

int crm_fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define fail crm_fail

// Work split over host threads (upload / download of 1e6-rod parameter blocks is host-bound otherwise).
template <class F>
static void parallel_for(size_t n, F f) {
    // (several ranks of one job share the host's cores: LOCAL_WORLD_SIZE, set by torchrun, divides them up;
    // CRM_B200_HOST_THREADS overrides)
    static const unsigned share = [] {
        const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        if (const char* e = getenv("CRM_B200_HOST_THREADS")) return (unsigned)std::max(1, atoi(e));
        unsigned ranks = 1;
        if (const char* e = getenv("LOCAL_WORLD_SIZE")) ranks = (unsigned)std::max(1, atoi(e));
        return std::max(1u, std::min(16u, hw / ranks));
    }();
    const size_t nt = std::min<size_t>(share, n / 16384);
    if (nt <= 1) {
        f((size_t)0, n);
        return;
    }
    std::vector<std::thread> th;
    const size_t chunk = (n + nt - 1) / nt;
    for (size_t t = 1; t < nt; ++t) th.emplace_back([=] { f(std::min(n, t * chunk), std::min(n, (t + 1) * chunk)); });
    f((size_t)0, std::min(n, chunk));
    for (auto& t : th) t.join();
}

// Process-wide page-locked staging for the parameter blocks of upload / download: converted records cross the bus
// from (into) it at link speed and asynchronously, which pageable std::vectors never do (the driver bounces them
// through its own small pinned buffer, ~3x slower, and blocks the host meanwhile).  One buffer, grown on demand
// up to POOL_LIMIT and kept until the process ends (cudaHostAlloc costs ~0.3 ms/MB: paid once, not per sim); a
// call that finds it in use by another thread, or needs more than the limit, falls back to pageable memory.
namespace {
constexpr size_t POOL_LIMIT = (size_t)1 << 30;
struct PinnedPool {
    std::mutex m;
    char* p = nullptr;
    size_t cap = 0;
    bool busy = false;
} g_pool;
struct PinnedLease {
    char* p = nullptr;
    bool pooled = false;
    crm_hvec<char> pageable;
    explicit PinnedLease(size_t bytes) {
        if (bytes <= POOL_LIMIT) {
            std::lock_guard<std::mutex> lock(g_pool.m);
            if (!g_pool.busy) {
                if (g_pool.cap < bytes) {
                    if (g_pool.p) cudaFreeHost(g_pool.p);
                    g_pool.p = nullptr;
                    g_pool.cap = 0;
                    const size_t want = (bytes + (bytes >> 3) + ((size_t)1 << 20) - 1) >> 20 << 20;
                    void* q = nullptr;
                    if (cudaHostAlloc(&q, want, cudaHostAllocPortable) == cudaSuccess) {
                        g_pool.p = static_cast<char*>(q);
                        g_pool.cap = want;
                    } else {
                        cudaGetLastError();
                    }
                }
                if (g_pool.cap >= bytes) {
                    g_pool.busy = true;
                    pooled = true;
                    p = g_pool.p;
                }
            }
        }
        if (!pooled) {
            pageable.resize(bytes + 256);
            p = pageable.data() + ((256 - (reinterpret_cast<uintptr_t>(pageable.data()) & 255)) & 255);
        }
    }
    ~PinnedLease() {
        if (pooled) {
            std::lock_guard<std::mutex> lock(g_pool.m);
            g_pool.busy = false;
        }
    }
    PinnedLease(const PinnedLease&) = delete;
    PinnedLease& operator=(const PinnedLease&) = delete;
    // carve `count` items of T (256-B aligned parts)
    template <class T> T* take(size_t& off, size_t count) {
        T* r = reinterpret_cast<T*>(p + off);
        off += (sizeof(T) * count + 255) / 256 * 256;
        return r;
    }
    template <class T> static size_t part(size_t count) { return (sizeof(T) * count + 255) / 256 * 256; }
};
}  // namespace

// All per-rod arrays live in ONE device allocation (an arena): a colony of 1e6 rods needs ~45 arrays, and 45
// cudaMalloc calls cost more than moving the colony across the bus (they dominated the upload of the end-to-end
// path and varied 4x between hosts).  Growth allocates a new arena, copies what must survive, frees the old one.
struct ArenaItem {
    void** pptr;
    size_t elem, old_count, new_count;
    int fill;  // byte the fresh part is set to (0 or 0xff)
};

// grow every per-rod array to `need` slots, preserving the first s->n
int crm_reserve(crm_sim* s, uint32_t need) {
    if (need <= s->cap) return 0;
    // first allocation: what the caller asks for (set_agents adds head room for growth); later: doubling
    uint32_t cap = s->cap ? s->cap : (need + 4095u) / 4096u * 4096u;
    while (cap < need) cap *= 2;
    const size_t st_ = (size_t)3 * s->lp;
    const size_t on = s->n, oc = (size_t)s->n * st_;
    std::vector<ArenaItem> items;
#define G(p, cnt_old, cnt_new) items.push_back(ArenaItem{reinterpret_cast<void**>(&(p)), sizeof(*(p)), (size_t)(cnt_old), (size_t)(cnt_new), 0})
    G(s->d.pos, oc, cap * st_);
    G(s->d.vel, oc, cap * st_);
    G(s->d.force, oc, cap * st_);
    G(s->d.dx1, oc, cap * st_);
    G(s->d.dv1, oc, cap * st_);
    G(s->d.dx2, oc, cap * st_);
    G(s->d.dv2, oc, cap * st_);
    G(s->d.nhist, on, (size_t)cap);
    G(s->d.pot, on, (size_t)cap);
    G(s->d.mech, on, (size_t)cap);
    G(s->d.grow, on, (size_t)cap);
    G(s->d.cyc, on, (size_t)cap);
    G(s->d.nn, on, (size_t)cap);
    G(s->tile_buf, on, (size_t)cap);
    G(s->rkey_buf, on, (size_t)cap);
    G(s->d.bsphere, on, (size_t)cap);
    G(s->d.chord, 2 * on, (size_t)2 * cap);
    G(s->d.vrec, on * s->lp, (size_t)cap * s->lp);
    G(s->vrec_alt, 0, (size_t)cap * s->lp);
    G(s->d.plist, 0, (size_t)cap * PL_MAXP_MAX);
    G(s->d.pqo, 0, (size_t)cap * PL_QP * s->lp);
    G(s->d.pqe, 0, (size_t)cap * PL_QP * s->lp);
    G(s->d.pqs, 0, (size_t)cap * PL_QS * s->lp);
    G(s->d.pnq, 0, (size_t)cap * s->lp);
    G(s->d.pslow, 0, (size_t)cap);
    G(s->d.vbuild, 0, (size_t)cap * s->lp);
    G(s->d.gof, 0, (size_t)cap);
    G(s->d.dcnt, 0, (size_t)cap);
    G(s->d.dextra, 0, (size_t)cap * PATCH_EXTRA);
    G(s->d.dmark, 0, (size_t)cap);
    G(s->d.dsib, 0, (size_t)cap);
    items.back().fill = 0xff;
    G(s->d.dirty_list, 0, (size_t)cap);
    G(s->long_rows, 0, (size_t)cap);
    G(s->d.dirty_n, 0, (size_t)1);
    G(s->d.cell_of, 0, (size_t)cap);
    G(s->d.cell_rank, 0, (size_t)cap);
    G(s->d.sorted, 0, (size_t)cap);
    G(s->d.bs_sorted, 0, (size_t)cap);
    G(s->d.chord_sorted, 0, (size_t)2 * cap);
    G(s->d.divlist, 0, (size_t)cap);
    G(s->stage, 0, (size_t)cap * s->nv * 3);
    G(s->stage2, 0, (size_t)cap * s->nv * 3);
#undef G
    size_t total = 0;
    std::vector<size_t> off(items.size());
    for (size_t i = 0; i < items.size(); ++i) {
        off[i] = total;
        total += (items[i].elem * items[i].new_count + 255) / 256 * 256;
    }
    char* arena = nullptr;
    if (cudaMalloc(&arena, total) != cudaSuccess) {
        cudaGetLastError();
        return fail(CRM_ERR_NOMEM, "out of device memory: %zu bytes for %u rod slots", total, cap);
    }
    // fresh parts are zeroed (only what is not overwritten by the copy), surviving parts copied
    for (size_t i = 0; i < items.size(); ++i) {
        const ArenaItem& it = items[i];
        char* dst = arena + off[i];
        const size_t keep = (*it.pptr && it.old_count) ? it.elem * it.old_count : 0;
        if (keep) CK(cudaMemcpyAsync(dst, *it.pptr, keep, cudaMemcpyDeviceToDevice, s->stream));
        CK(cudaMemsetAsync(dst + keep, it.fill, it.elem * it.new_count - keep, s->stream));
        *it.pptr = dst;
    }
    if (s->arena) {  // the copies out of the old arena must have finished
        CK(cudaStreamSynchronize(s->stream));
        CK(cudaFree(s->arena));
    }
    s->arena = arena;
    s->n_dirty = 0;
    s->n_long = 0;
    s->need_rebuild = true;  // the structure does not survive a reallocation
    s->have_lists = false;
    s->d.divcap = cap;
    s->d.tile = s->sim_of.empty() ? nullptr : s->tile_buf;
    s->d.rkey = s->rkey_on ? s->rkey_buf : nullptr;
    s->cap = cap;
    return 0;
}

// device copy of the global slots [first, first + count): the key of the rods' random streams
int crm_upload_rkeys(crm_sim* s, uint32_t first, uint32_t count) {
    if (!s->rkey_on || !count) return 0;
    std::vector<uint32_t> k(count);
    for (uint32_t i = 0; i < count; ++i) k[i] = (uint32_t)s->gslot[first + i];
    CK(cudaMemcpyAsync(s->rkey_buf + first, k.data(), sizeof(uint32_t) * count, cudaMemcpyHostToDevice, s->stream));
    CK(cudaStreamSynchronize(s->stream));  // (k is a temporary)
    s->d.rkey = s->rkey_buf;
    return 0;
}

static int reserve_cells(crm_sim* s, uint32_t ncells) {
    if (ncells + 1 <= s->ncells_cap) return 0;
    uint32_t c = s->ncells_cap ? s->ncells_cap : 4096;
    while (c < ncells + 1) c *= 2;
    if (s->d.cell_start) CK(cudaFree(s->d.cell_start));
    if (s->d.cell_count) CK(cudaFree(s->d.cell_count));
    CK(cudaMalloc(&s->d.cell_start, sizeof(uint32_t) * c));
    CK(cudaMalloc(&s->d.cell_count, sizeof(uint32_t) * c));
    s->ncells_cap = c;
    return 0;
}

// Centroid cell list geometry: the 3x3 stencil is complete when the cell edge is at
// least 2*max_extent + search radius.
// interaction range used by the search: inflated by the skin while the persistent structure is in use
static bool persistent(const crm_sim* s) { return s->skin > 0.0f; }
static float search_radius(const crm_sim* s) { return s->rsearch + (persistent(s) ? s->skin : 0.0f); }

static int configure_grid(crm_sim* s, float emax_own = -1.0f) {
    if (emax_own < 0.0f) emax_own = s->emax;
    const float h = (2.0f * std::max(emax_own, s->n_ghost ? s->emax_ghost : 0.0f) + search_radius(s)) * 1.0005f + 1e-4f;
    double ncx = std::floor((double)s->cfg.domain_size[0] / h), ncy = std::floor((double)s->cfg.domain_size[1] / h);
    if (!(ncx >= 1.0)) ncx = 1.0;
    if (!(ncy >= 1.0)) ncy = 1.0;
    const double MAXC = 4.0 * 1024 * 1024 / (double)s->n_sims;  // batched sims: one tile of cells per sim
    while (ncx * ncy > MAXC && ncx * ncy > 1.0) {
        ncx = std::max(1.0, std::floor(ncx / 1.5));
        ncy = std::max(1.0, std::floor(ncy / 1.5));
    }
    s->d.ncx = (int)ncx;
    s->d.ncy = (int)ncy;
    s->d.inv_wx = (float)(ncx / (double)s->cfg.domain_size[0]);
    s->d.inv_wy = (float)(ncy / (double)s->cfg.domain_size[1]);
    s->d.cpt = (uint32_t)(ncx * ncy);
    return reserve_cells(s, (uint32_t)(ncx * ncy) * s->n_sims);
}

// ------------------------------------------------------------------------------------
// Largest f32 x with sqrtf(x) <= cutoff: lets the kernel test the reference's `r > cutoff`
// on r^2 without changing a single outcome.
static float cutoff_threshold(float c) {
    if (!(c >= 0.0f) || !std::isfinite(c)) return c >= 0.0f ? INFINITY : -1.0f;
    float x = c * c;
    while (x > 0.0f && std::sqrt(x) > c) x = std::nextafter(x, 0.0f);
    while (std::sqrt(std::nextafter(x, INFINITY)) <= c) x = std::nextafter(x, INFINITY);
    return x;
}

// The constants that cost something (two powf, the cutoff search) repeat from rod to rod in every real
// colony: remembered for the last distinct (cutoff), (en, em) seen.  Same bits as computing them afresh.
struct ParamCache {
    bool has_c = false, has_m = false;
    float cutoff = 0, cutoff2 = 0, en = 0, em = 0, C = 0, d = 0;
};

static void fill_device_params_cached(const crm_rod_params& p, ParamCache& k, PotParams& pot, MechParams& mech,
                                      GrowState& grow, CycleParams& cyc) {
    pot.type = p.potential;
    pot.radius = p.radius;
    if (!k.has_c || memcmp(&k.cutoff, &p.cutoff, sizeof(float)) != 0) {
        k.cutoff = p.cutoff;
        k.cutoff2 = cutoff_threshold(p.cutoff);
        k.has_c = true;
    }
    pot.cutoff2 = k.cutoff2;
    if (p.potential == CRM_POT_MORSE) {
        pot.strength = p.strength;
        pot.a = p.potential_stiffness;
        pot.b = pot.c = pot.d = 0.0f;
    } else {
        // same f32 operation order as the reference's `strength * mie_constant * part`
        const float en = p.en, em = p.em;
        if (!k.has_m || memcmp(&k.en, &en, sizeof(float)) != 0 || memcmp(&k.em, &em, sizeof(float)) != 0) {
            k.en = en;
            k.em = em;
            k.C = en / (en - em) * powf(en / em, en / (en - em));
            k.d = powf(em / en, 1.0f / (en - em));
            k.has_m = true;
        }
        pot.strength = p.strength * k.C;
        pot.a = p.bound;
        pot.b = en;
        pot.c = em;
        pot.d = k.d;
    }
    mech = make_float4(p.spring_tension, p.rigidity, p.damping, p.diffusion_constant);
    grow = make_float4(p.growth_rate, p.spring_length_threshold, p.spring_length, p.has_neighbor_reduction ? 1.0f : 0.0f);
    memset(&cyc, 0, sizeof(cyc));
    cyc.growth_kind = p.growth_setter_kind;
    cyc.growth_a = p.growth_setter_a;
    cyc.growth_b = p.growth_setter_b;
    cyc.thr_kind = p.threshold_setter_kind;
    cyc.thr_a = p.threshold_setter_a;
    cyc.thr_b = p.threshold_setter_b;
    cyc.has_nr = p.has_neighbor_reduction;
    cyc.nr_max = p.neighbor_reduction_max;
    cyc.nr_exp = p.neighbor_reduction_exp;
}

void crm_fill_device_params(const crm_rod_params& p, PotParams& pot, MechParams& mech, GrowState& grow,
                            CycleParams& cyc) {
    ParamCache k;
    fill_device_params_cached(p, k, pot, mech, grow, cyc);
}

// what the (re)builds of the persistent structure reported about its rods; sticky until the next full build
static void crm_note_list_flags(crm_sim* s) {
    s->lists_maxl = std::max(s->lists_maxl, s->h_status->max_list);
    if (s->h_status->any_slow) s->lists_slow = true;
    if (s->h_status->seen_morse) s->lists_morse = true;
    if (s->h_status->seen_mie) s->lists_mie = true;
}

int crm_read_status(crm_sim* s) {
    CK(cudaMemcpyAsync(s->h_status, s->d.status, sizeof(Status), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    crm_note_list_flags(s);
    return 0;
}

int crm_report_device_error(crm_sim* s) {
    const Status& st = *s->h_status;
    if (st.err_code == 0) return 0;
    std::vector<float> rod((size_t)3 * s->lp);
    cudaMemcpy(rod.data(), s->d.pos + (size_t)st.err_slot * 3 * s->lp, sizeof(float) * rod.size(),
               cudaMemcpyDeviceToHost);
    std::string dump;
    char tmp[64];
    for (int c = 0; c < 3; ++c)
        for (int i = 0; i < s->nv; ++i) {
            snprintf(tmp, sizeof(tmp), "%s%g", dump.empty() ? "" : ", ", rod[c * s->lp + i]);
            dump += tmp;
        }
    if (st.err_code == CRM_ERR_BOUNDARY)
        return fail(CRM_ERR_BOUNDARY, "BoundaryError: Particle is out of domain at pos [%s] (rod slot %u, iteration %llu)",
                    dump.c_str(), st.err_slot, (unsigned long long)s->iteration);
    return fail(CRM_ERR_NUMERIC, "CalcError: non-finite position [%s] (rod slot %u, iteration %llu)", dump.c_str(),
                st.err_slot, (unsigned long long)s->iteration);
}

// Divisions of this step.  fetch_divisions orders the mothers by global slot (= slot when the run
// is not sharded); apply_divisions runs the geometry kernel and installs the identifiers of the
// backend's update_cell_cycle_4 (mothers are renamed first, then the new cells, with the per-voxel
// pre-incremented counter).  Unsharded runs derive the identifiers here; sharded runs get them from
// the coordinator, which sees the divisions of all slabs (crm_sim_apply_divisions).
// one block: the division list ordered by (global slot, slot) on the device (bitonic sort in shared memory).  The
// atomic appends of k_integrate arrive in any order; the order fixes identifiers and daughter slots.
constexpr uint32_t DIV_SORT_MAX = 2048;
__global__ void __launch_bounds__(1024) k_div_sort(DivRecord* list, uint32_t n, const uint32_t* __restrict__ rkey) {
    __shared__ unsigned long long key[DIV_SORT_MAX];
    __shared__ uint32_t vox[DIV_SORT_MAX];
    uint32_t m = 1;
    while (m < n) m <<= 1;
    for (uint32_t i = threadIdx.x; i < m; i += blockDim.x) {
        if (i < n) {
            const DivRecord r = list[i];
            key[i] = ((unsigned long long)(rkey ? rkey[r.slot] : r.slot) << 32) | r.slot;
            vox[i] = r.uvox;
        } else {
            key[i] = ~0ull;
            vox[i] = 0;
        }
    }
    __syncthreads();
    for (uint32_t k = 2; k <= m; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t i = threadIdx.x; i < m; i += blockDim.x) {
                const uint32_t l = i ^ j;
                if (l > i) {
                    const bool up = (i & k) == 0;
                    if ((key[i] > key[l]) == up) {
                        const unsigned long long t = key[i]; key[i] = key[l]; key[l] = t;
                        const uint32_t v = vox[i]; vox[i] = vox[l]; vox[l] = v;
                    }
                }
            }
            __syncthreads();
        }
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        list[i].slot = (uint32_t)(key[i] & 0xffffffffull);
        list[i].uvox = vox[i];
    }
}

// Divisions of this step.  The mothers are ordered by global slot (= slot when the run is not sharded) on the
// device; the geometry kernel, the patch kernels and the read-back of the ordered list are enqueued together and
// the host waits ONCE, then installs the identifiers of the backend's update_cell_cycle_4 (mothers are renamed
// first, then the new cells, with the per-voxel pre-incremented counter).  Unsharded runs derive the
// identifiers here; sharded runs get them from the coordinator, which sees the divisions of all slabs
// (fetch_divisions -> crm_sim_pending_divisions -> crm_sim_apply_divisions).
static int sort_divisions_on_device(crm_sim* s, uint32_t n_div) {
    if (n_div > 1 && n_div <= DIV_SORT_MAX) {
        k_div_sort<<<1, 1024, 0, s->stream>>>(s->d.divlist, n_div, s->d.rkey);
        s->launches += 1;
        return 1;
    }
    return n_div <= 1 ? 1 : 0;
}

static int fetch_divisions(crm_sim* s, uint32_t n_div) {
    const int sorted = sort_divisions_on_device(s, n_div);
    s->h_div.resize(n_div);
    CK(cudaMemcpyAsync(s->h_div.data(), s->d.divlist, sizeof(DivRecord) * n_div, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    if (!sorted) {  // (more divisions in one step than the one-block sort takes)
        std::sort(s->h_div.begin(), s->h_div.end(), [s](const DivRecord& a, const DivRecord& b) {
            return s->gslot[a.slot] != s->gslot[b.slot] ? s->gslot[a.slot] < s->gslot[b.slot] : a.slot < b.slot;
        });
        CK(cudaMemcpyAsync(s->d.divlist, s->h_div.data(), sizeof(DivRecord) * n_div, cudaMemcpyHostToDevice, s->stream));
    }
    for (uint32_t r = 0; r < n_div; ++r) {
        const crm_rod_params& p = s->base[s->h_div[r].slot];
        if (!(p.radius > 0.0f))
            return fail(CRM_ERR_DIVISION, "DivisionError: cannot divide rod slot %u (radius %g)", s->h_div[r].slot, p.radius);
    }
    return 0;
}

// geometry + (optionally) the patch of the live neighbour structure for the ordered list on the device; status
// zeroed in front, nothing waits
static int divisions_enqueue(crm_sim* s, uint32_t n_div, uint32_t n0, bool patch, uint32_t* h_dirty) {
    CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
    launch_divide(s->nv, s->d, n_div, n0, s->stream);
    s->launches += 1;
    if (patch) {
        CK(cudaMemsetAsync(s->d.dirty_n, 0, sizeof(uint32_t), s->stream));
        s->launches += launch_patch(s->d, n_div, n0, s->lp, s->lp > 8 ? 16 : 32, s->stream);
        CK(cudaMemcpyAsync(h_dirty, s->d.dirty_n, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    }
    CK(cudaGetLastError());
    return 0;
}

// host bookkeeping of the divisions in s->h_div (ordered), after the kernels' status has arrived
static int divisions_bookkeep(crm_sim* s, uint32_t n0, bool patch, const uint32_t* h_dirty, const crm_ident* id_mother,
                              const crm_ident* id_daughter, const uint64_t* gslot_daughter) {
    const uint32_t n_div = (uint32_t)s->h_div.size();
    int rc;
    crm_note_list_flags(s);
    if ((rc = crm_report_device_error(s))) return rc;
    bool patched = false;
    if (patch) {
        patched = s->h_status->patch_overflow == 0;
        if (patched) {
            s->n_dirty = *h_dirty;
            s->n_patches += 1;
        } else {  // a row ran out of room for the announced rods: search again, scratch arrays back to empty
            CK(cudaMemsetAsync(s->d.dcnt, 0, sizeof(uint32_t) * s->cap, s->stream));
            CK(cudaMemsetAsync(s->d.dmark, 0, sizeof(uint32_t) * s->cap, s->stream));
            CK(cudaMemsetAsync(s->d.dsib, 0xff, sizeof(uint32_t) * s->cap, s->stream));
        }
    }
    s->base.resize(n0 + n_div);
    s->ids.resize(n0 + n_div);
    s->parents.resize(n0 + n_div);
    s->gslot.resize(n0 + n_div);
    for (uint32_t r = 0; r < n_div; ++r) {
        const uint32_t a = s->h_div[r].slot, b = n0 + r;
        const crm_ident mother = s->ids[a];
        s->ids[a] = id_mother[r];
        s->parents[a] = mother;
        s->ids[b] = id_daughter[r];
        s->parents[b] = mother;
        s->base[b] = s->base[a];
        s->gslot[b] = gslot_daughter[r];
        if (!s->sim_of.empty()) s->sim_of.push_back(s->sim_of[a]);  // b == sim_of.size(): slots are appended in order
    }
    s->n = n0 + n_div;
    if ((rc = crm_upload_rkeys(s, n0, n_div))) return rc;
    s->n_ghost = 0;  // the new cells took the first ghost slots; ghosts are re-sent every step
    s->d.n = s->d.n_own = s->n;
    s->bounds_fresh = false;
    s->h_div.clear();
    s->pending_div = 0;
    if (!patched) {
        s->need_rebuild = true;
        s->have_lists = false;
        s->n_dirty = 0;
    }
    return 0;
}

// sharded runs: the ordered list is already on both sides (fetch_divisions), the identifiers come from the coordinator
int crm_apply_divisions(crm_sim* s, const crm_ident* id_mother, const crm_ident* id_daughter,
                           const uint64_t* gslot_daughter) {
    const uint32_t n0 = s->n, n_div = (uint32_t)s->h_div.size();
    int rc = crm_reserve(s, n0 + n_div);
    if (rc) return rc;
    // (the list on the device does not survive a reallocation: the host's ordered copy goes back)
    CK(cudaMemcpyAsync(s->d.divlist, s->h_div.data(), sizeof(DivRecord) * n_div, cudaMemcpyHostToDevice, s->stream));
    s->patch_divisions = false;  // (sharded runs search again after a division: it is a collective decision)
    if ((rc = divisions_enqueue(s, n_div, n0, false, nullptr))) return rc;
    if ((rc = crm_read_status(s))) return rc;
    return divisions_bookkeep(s, n0, false, nullptr, id_mother, id_daughter, gslot_daughter);
}

// identifiers of an unsharded division step from the ordered list in s->h_div, then the bookkeeping
static int divisions_ids_and_bookkeep(crm_sim* s, uint32_t n0, bool patch, const uint32_t* h_dirty) {
    const uint32_t n_div = (uint32_t)s->h_div.size();
    const uint32_t nvox = s->cfg.n_voxels[0] * s->cfg.n_voxels[1];
    const bool batched = !s->sim_of.empty();
    // counters are per (sim, user voxel); mothers come ordered by slot, which inside a sim is the
    // slot order of that sim run on its own
    auto key = [&](uint32_t r) { return (size_t)(batched ? s->sim_of[s->h_div[r].slot] : 0u) * nvox + s->h_div[r].uvox; };
    std::vector<uint32_t> per_vox((size_t)nvox * s->n_sims, 0), rank_vox((size_t)nvox * s->n_sims, 0);
    for (uint32_t r = 0; r < n_div; ++r) per_vox[key(r)]++;
    std::vector<crm_ident> idm(n_div), idd(n_div);
    std::vector<uint64_t> gs(n_div);
    for (uint32_t r = 0; r < n_div; ++r) {
        const uint32_t v = s->h_div[r].uvox;
        const size_t kv = key(r);
        const uint32_t k = rank_vox[kv]++;
        idm[r] = crm_ident{CRM_ID_DIVISION, v, s->vox_counter[kv] + 1 + k};
        idd[r] = crm_ident{CRM_ID_DIVISION, v, s->vox_counter[kv] + per_vox[kv] + 1 + k};
        gs[r] = batched ? s->sim_rods[s->sim_of[s->h_div[r].slot]]++ : (uint64_t)n0 + r;
    }
    for (size_t kv = 0; kv < per_vox.size(); ++kv) s->vox_counter[kv] += 2ull * per_vox[kv];
    return divisions_bookkeep(s, n0, patch, h_dirty, idm.data(), idd.data(), gs.data());
}

static int handle_divisions(crm_sim* s, uint32_t n_div) {
    const uint32_t n0 = s->n;
    int rc;
    // Divisions were the only reason to search again and the structure is live: patch it (the rows of the
    // daughters and of the mothers' neighbours are rebuilt from candidate lists at the start of the next step)
    const bool patch = s->patch_divisions && s->have_lists && !s->need_rebuild && !s->sharded && s->skin > 0.0f;
    s->patch_divisions = false;
    uint32_t* const h_dirty = reinterpret_cast<uint32_t*>(s->h_status + 1) + 16;  // pinned scratch behind the status
    const bool grow = n0 + n_div > s->cap;
    if (grow || n_div > DIV_SORT_MAX || n_div > s->h_div_pinned_cap) {
        // the arrays move (the live structure does not survive that) or the list is too long for the one-block
        // sort: order it through the host first
        if ((rc = fetch_divisions(s, n_div))) return rc;
        if ((rc = crm_reserve(s, n0 + n_div))) return rc;
        // (the list on the device does not survive a reallocation: the host's ordered copy goes back)
        CK(cudaMemcpyAsync(s->d.divlist, s->h_div.data(), sizeof(DivRecord) * n_div, cudaMemcpyHostToDevice, s->stream));
        const bool p2 = patch && !grow;
        if ((rc = divisions_enqueue(s, n_div, n0, p2, h_dirty))) return rc;
        if ((rc = crm_read_status(s))) return rc;
        return divisions_ids_and_bookkeep(s, n0, p2, h_dirty);
    }
    // everything on the device, one wait: order the list, read it back (page-locked), divide, patch, status
    sort_divisions_on_device(s, n_div);
    CK(cudaMemcpyAsync(s->h_div_pinned, s->d.divlist, sizeof(DivRecord) * n_div, cudaMemcpyDeviceToHost, s->stream));
    if ((rc = divisions_enqueue(s, n_div, n0, patch, h_dirty))) return rc;
    if ((rc = crm_read_status(s))) return rc;
    s->h_div.assign(s->h_div_pinned, s->h_div_pinned + n_div);
    for (uint32_t r = 0; r < n_div; ++r) {
        const crm_rod_params& p = s->base[s->h_div[r].slot];
        if (!(p.radius > 0.0f))
            return fail(CRM_ERR_DIVISION, "DivisionError: cannot divide rod slot %u (radius %g)", s->h_div[r].slot, p.radius);
    }
    return divisions_ids_and_bookkeep(s, n0, patch, h_dirty);
}

// Bounds of the search (bounding sphere, chord, cull thresholds, max extent) for the current
// positions.  The integrator only maintains the vertex records, so this runs before every search.
int crm_refresh_bounds(crm_sim* s) {
    if (s->bounds_fresh || s->n == 0) return 0;
    s->d.n_own = s->n;
    s->d.rsearch = search_radius(s);
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

__global__ void k_grid_check(const Status* st, uint32_t limit_bits, uint32_t* skip) { *skip = st->emax_bits > limit_bits ? 1u : 0u; }

// `async_bounds` (search steps of crm_sim_step & co.): the bounds kernel is enqueued in front of the search without
// waiting for its result; the grid is sized with head-room over the last known extent and the kernels of the step
// return at once should a rod exceed it (Dev::emax_limit_bits; step_finish then repeats the step the slow way).
static int launch_neighbour_and_force(crm_sim* s, bool count, bool build = false, bool async_bounds = false) {
    int rc;
    s->d.emax_limit_bits = 0u;
    s->d.skip = nullptr;
    s->async_search = false;
    if (async_bounds && !s->bounds_fresh && !s->sharded && s->n > 0 && s->emax > 0.0f && std::isfinite(s->emax)) {
        const float emax_cfg = s->emax * 1.05f + 0.05f;
        if ((rc = configure_grid(s, emax_cfg))) return rc;
        s->d.n = s->n + s->n_ghost;
        s->d.n_own = s->n;
        s->d.rsearch = search_radius(s);
        s->d.skin = s->skin;
        CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
        launch_bounds_kernel(s->nv, s->d, s->stream);  // status->emax_bits: read by the kernels below, and by the host after the step
        s->d.emax_limit_bits = __builtin_bit_cast(uint32_t, emax_cfg);
        k_grid_check<<<1, 1, 0, s->stream>>>(s->d.status, s->d.emax_limit_bits, s->latch_buf + 2);
        s->launches += 2;
        s->d.skip = s->latch_buf + 2;
        s->async_search = true;
    } else {
        if ((rc = crm_refresh_bounds(s))) return rc;
        if ((rc = configure_grid(s))) return rc;
        s->d.n = s->n + s->n_ghost;
        s->d.n_own = s->n;
        s->d.rsearch = search_radius(s);
        s->d.skin = s->skin;
        CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
    }
    s->launches += launch_binning(s->d, s->n_sims, s->stream);
    if (s->timers) CK(cudaEventRecord(s->ev0, s->stream));
    if (build) {  // until this step's status says otherwise (read before any list step)
        s->lists_slow = s->lists_morse = s->lists_mie = false;
        s->lists_maxl = 0;
    }
    launch_force(s->nv, s->d, s->fast_math, count, build, s->stream);
    if (s->timers) CK(cudaEventRecord(s->ev1, s->stream));
    s->launches += 1;
    return 0;
}

// list entries the list kernel stages per rod: exactly the longest list of the live structure (rows that a division
// is about to patch grow by at most PATCH_EXTRA candidates before the host hears of their new length).  Not rounded
// up: at 8 vertices 8 entries leave room for 8 resident blocks per SM, 9 for 7, 12 for 6 (a slab of a sharded
// lattice lists 9 where the whole colony lists 8: rounding to 12 cost 13 % of the list step).
static void set_staging(crm_sim* s, bool rows_pending) {
    const int cap = s->lp > 8 ? 16 : 32;
    const int m = std::max(1, (int)(s->lists_maxl + (rows_pending ? (uint32_t)PATCH_EXTRA : 0u)));
    s->d.maxl = std::min(m, cap);
    // (Staging the records in passes of 8 rods, so that the footprint does not follow the longest list, was built and
    // measured: the per-entry pass check and the cursors held across passes cost 10 % of a single-pass step, and
    // the relaxed lattice with one list of 13 rods ran no faster: 0.99 vs 0.93 ms.  Not kept.)
}

// Rows in two classes.  The staging area of the list kernel is sized for the longest list it serves, and one rod that
// lists 13 others leaves 6 resident blocks per SM where 8 entries leave 8.  So lists of up to short_maxl rods
// (the longest staging that still leaves 7 blocks) are served by the main launch over all rows, which skips
// longer rows, and those (k_classify_rows collects them after every (re)build of rows) by a second launch
// over just them.  Results do not depend on the split: a row's arithmetic is the same in either launch.
static int short_maxl(const crm_sim* s) {
    const int cap = s->lp > 8 ? 16 : 32;
    int t = 4;
    while (t < cap && (232448 / (list_kernel_smem(s->lp, t + 1, true) + 1024)) >= 7) ++t;
    return t;
}

__global__ void k_classify_rows(const uint32_t* __restrict__ pnq, const uint32_t* __restrict__ sorted, uint32_t n, uint32_t n_own,
                                int lp, uint32_t thr, uint32_t* __restrict__ rows, Status* st) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n || sorted[g] >= n_own) return;  // (rows of ghost rods are never written: whatever they hold is stale)
    const uint32_t c = pnq[(size_t)g * lp] >> 24;
    if (c != 0xffu && c > thr) rows[atomicAdd(&st->n_long, 1u)] = g;
}

// after a (re)build of rows, in the same step: the long rows of the structure as it is now
static int classify_rows(crm_sim* s) {
    const uint32_t n = s->n + s->n_ghost;
    if (n == 0) return 0;
    k_classify_rows<<<blocks_for(n, 256), 256, 0, s->stream>>>(s->d.pnq, s->d.sorted, n, s->n, s->lp, (uint32_t)short_maxl(s),
                                                               s->long_rows, s->d.status);
    s->launches += 1;
    s->classified = true;
    CK(cudaGetLastError());
    return 0;
}

// the list kernel(s) of one step; s->d.maxl = longest list (set_staging).  `all_rows`: rows were rebuilt in front of
// this launch, the classes are stale: one launch staged for the longest list.
static int launch_list_step(crm_sim* s, bool uniform, int pt, bool fuse, bool all_rows) {
    s->d.rows = nullptr;
    s->d.n_rows = 0;
    if (all_rows || s->n_long == 0 || s->d.maxl <= short_maxl(s)) {
        launch_force_eval(s->nv, s->d, s->fast_math, uniform, pt, fuse, s->stream);
        s->launches += 1;
        return 0;
    }
    Dev da = s->d;
    da.maxl = short_maxl(s);
    launch_force_eval(s->nv, da, s->fast_math, uniform, pt, fuse, s->stream);
    Dev db = s->d;
    db.rows = s->long_rows;
    db.n_rows = s->n_long;
    launch_force_eval(s->nv, db, s->fast_math, uniform, pt, fuse, s->stream);
    s->launches += 2;
    s->n_two_class_steps += 1;
    return 0;
}

// the list step can run force + mechanics in ONE kernel (k_force_eval<.., FUSE>): fast math, no mixed pairs
static bool fusable(const crm_sim* s) { return s->fused && s->fast_math && !s->lists_slow; }
// after a fused step the buffer it wrote holds the current vertex records
static void swap_vrec(crm_sim* s) { std::swap(s->d.vrec, s->vrec_alt); }

// force of a step between two rebuilds: no cell list, no search (k_force_eval)
static int launch_force_from_lists(crm_sim* s) {
    s->d.emax_limit_bits = 0u;
    s->d.skip = nullptr;
    s->async_search = false;
    s->d.n = s->n + s->n_ghost;  // the build order covers the ghosts as well
    s->d.n_own = s->n;
    CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
    set_staging(s, s->n_dirty != 0);
    const bool rows_rebuilt = s->n_dirty != 0;
    if (s->n_dirty) {  // rows touched by the divisions of the last step: rebuilt with twice the skin as
        // margin, because their partners' displacement budgets keep counting from the last full build
        Dev dl = s->d;
        dl.n_dirty = s->n_dirty;
        dl.rsearch = s->rsearch + 2.0f * s->skin;
        dl.skin = 2.0f * s->skin;
        launch_force_lists(s->nv, dl, s->fast_math, s->stream);
        s->launches += 1;
        s->n_dirty = 0;
    }
    if (s->timers) CK(cudaEventRecord(s->ev0, s->stream));
    const int pt = s->lists_morse == s->lists_mie ? -1 : (s->lists_mie ? CRM_POT_MIE : CRM_POT_MORSE);
    // force + mechanics in one kernel, also when rows were rebuilt in front of it (a row that did not fit keeps
    // the force the rebuild computed: the fused kernel integrates such a rod with that force)
    s->step_fused = fusable(s);
    if (s->step_fused) s->d.vrec_out = s->vrec_alt;
    int rc = launch_list_step(s, !s->lists_slow, pt, s->step_fused, rows_rebuilt);
    if (s->timers) CK(cudaEventRecord(s->ev1, s->stream));
    if (rc) return rc;
    return rows_rebuilt ? classify_rows(s) : 0;
}

// [n_div, reason to rebuild, error code, 0] of the step just enqueued, for the sharded coordinator's all_gather
// (the same conditions step_finish evaluates on the host)
__global__ void k_status_flags(const Status* st, float skin, uint32_t host_want, uint32_t* flags) {
    const float disp = sqrtf(__uint_as_float(st->disp2_bits));
    uint32_t want = host_want | (st->n_div > 0) | (st->n_leave > 0) | (st->err_code != 0);
    if (skin > 0.0f) want |= (st->plist_overflow != 0) | !(disp < 0.4999f * skin);
    flags[0] = st->n_div;
    flags[1] = want ? 1u : 0u;
    flags[2] = (uint32_t)(-st->err_code);  // 0 | -CRM_ERR_BOUNDARY | -CRM_ERR_NUMERIC: every slab learns of the error
    flags[3] = 0u;
}

// A step is enqueued (step_enqueue: nothing waits for the device unless a search needs fresh bounds)
// and finished (step_finish: one stream synchronisation, then the host's bookkeeping).
static int step_enqueue_open(crm_sim* s, uint32_t* dev_flags);
static int step_enqueue(crm_sim* s, uint32_t* dev_flags) {
    if (s->pending_div) return fail(CRM_ERR_INVALID, "divisions of the last step are pending (crm_sim_apply_divisions)");
    if (s->step_open) return fail(CRM_ERR_INVALID, "the last step was not finished (crm_sim_step_end)");
    s->step_open = true;
    const int rc = step_enqueue_open(s, dev_flags);
    if (rc) s->step_open = false;  // nothing to finish
    return rc;
}
static int step_enqueue_open(crm_sim* s, uint32_t* dev_flags) {
    if (s->n == 0) {
        s->iteration += 1;
        if (dev_flags) CK(cudaMemsetAsync(dev_flags, 0, 4 * sizeof(uint32_t), s->stream));
        return 0;
    }
    s->iteration += 1;
    s->rod_steps += s->n;
    s->d.iteration = s->iteration;
    int rc;
    const bool use_lists = persistent(s) && !s->need_rebuild && s->have_lists;
    if (use_lists) {
        rc = launch_force_from_lists(s);
        s->n_eval_steps += 1;
        s->evals_since_build += 1;
    } else {
        const bool build = persistent(s) && s->holdoff == 0;
        rc = launch_neighbour_and_force(s, false, build, s->async_bounds);
        if (build) {
            // lists that repeatedly are not used (a colony with divisions in every step): stop writing them
            // for a while, twice as long every time it happens again (8 .. 512 steps)
            if (s->evals_since_build >= 2 || s->n_builds == 0) {
                s->short_builds = 0;
                if (s->evals_since_build >= 2) s->backoff = 8;
            } else {
                s->short_builds += 1;
            }
            if (s->short_builds >= 4) {
                s->holdoff = s->backoff;
                s->backoff = std::min<uint32_t>(s->backoff * 2, 512);
                s->short_builds = 0;
            }
            if (!rc) rc = classify_rows(s);
            s->n_builds += 1;
            s->steps_since_build = 0;
            s->evals_since_build = 0;
            s->need_rebuild = false;
            s->have_lists = true;
        } else {
            s->have_lists = false;  // a plain search: whatever structure is on the device is stale
            if (s->holdoff) s->holdoff -= 1;
        }
    }
    if (rc) return rc;
    s->steps_since_build += 1;
    if (use_lists && s->step_fused) {
        swap_vrec(s);
        s->n_fused_steps += 1;
    } else {
        launch_integrate(s->nv, s->d, s->stream);
        s->launches += 1;
    }
    s->step_fused = false;
    s->bounds_fresh = false;
    if (dev_flags) {
        k_status_flags<<<1, 1, 0, s->stream>>>(s->d.status, persistent(s) ? s->skin : 0.0f, s->have_lists ? 0u : 1u, dev_flags);
        s->launches += 1;
    }
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(s->h_status, s->d.status, sizeof(Status), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaEventRecord(s->ev_status, s->stream));
    return 0;
}

// host half of a step (or of a batch of list steps) once its status block has arrived
static int step_conclude(crm_sim* s) {
    int rc;
    crm_note_list_flags(s);
    if ((rc = crm_report_device_error(s))) return rc;
    const uint32_t n_div = s->h_status->n_div;
    bool want = s->h_status->n_leave > 0;
    if (s->classified) {  // rows were (re)built in this step: the long rows of the structure as it is now
        s->n_long = s->h_status->n_long;
        s->classified = false;
    }
    s->why_rebuild[7] += 1;
    if (want) s->why_rebuild[1] += 1;
    if (persistent(s)) {
        if (!s->have_lists) s->why_rebuild[5] += 1;
        if (s->h_status->plist_overflow) {  // a rod did not fit: search every step for a while, then retry
            s->have_lists = false;
            s->holdoff = 64;
            s->overflow_reasons |= s->h_status->plist_overflow;
            s->n_overflows += 1;
            s->why_rebuild[2] += 1;
        }
        if (s->h_status->patch_overflow) {  // a patched row is full: a fresh search empties it
            want = true;
            s->why_rebuild[3] += 1;
        }
        const float disp = std::sqrt(__builtin_bit_cast(float, s->h_status->disp2_bits));
        if (!(disp < 0.4999f * s->skin)) {
            want = true;
            if (s->have_lists) s->why_rebuild[0] += 1;
        }
    }
    // divisions alone do not force a new search while the structure is live: its rows are patched
    s->patch_divisions = n_div > 0 && !want && persistent(s) && s->have_lists && !s->need_rebuild && !s->sharded &&
                         s->patching;
    if (n_div > 0 && !s->patch_divisions) {
        if (!want && persistent(s) && s->have_lists) s->why_rebuild[4] += 1;
        want = true;
    }
    s->want_rebuild = want || !s->have_lists;
    // sharded runs rebuild in all slabs at once: the coordinator clears the flag (crm_sim_keep_lists)
    if (s->sharded || s->want_rebuild) s->need_rebuild = true;
    if (n_div) {
        if (s->sharded) {  // identifiers need the divisions of all slabs: crm_sim_apply_divisions
            if ((rc = fetch_divisions(s, n_div))) return rc;
            s->pending_div = n_div;
        } else if ((rc = handle_divisions(s, n_div))) {
            return rc;
        }
    }
    return 0;
}

static int step_finish(crm_sim* s) {
    if (!s->step_open) return 0;
    s->step_open = false;
    if (s->n == 0) return 0;
    // the step's own work only: what a caller enqueued behind it on the stream (the sharded coordinator's
    // next halo exchange) keeps running while the host decides what the next step is
    CK(cudaEventSynchronize(s->ev_status));
    if (s->async_search) {
        s->async_search = false;
        const uint32_t bits = s->h_status->emax_bits;
        s->emax = __builtin_bit_cast(float, bits);  // the extent k_bounds measured in front of this step's search
        if (!std::isfinite(s->emax)) return fail(CRM_ERR_NUMERIC, "CalcError: non-finite positions (iteration %llu)",
                                                 (unsigned long long)s->iteration);
        if (bits > s->d.emax_limit_bits) {
            // a rod outgrew the head-room the grid was sized with: the force and integrate kernels of the step
            // returned without touching anything.  Repeat it with the bounds read back first.
            s->iteration -= 1;
            s->rod_steps -= s->n;
            s->n_grid_redos += 1;
            s->why_rebuild[6] += 1;
            s->need_rebuild = true;
            s->have_lists = false;
            s->bounds_fresh = true;  // (k_bounds ran on the current positions and s->emax is its result)
            const bool keep = s->async_bounds;
            s->async_bounds = false;
            const int rc = step_enqueue(s, nullptr);
            s->async_bounds = keep;
            return rc ? rc : step_finish(s);
        }
    }
    if (s->timers) {
        float ms = 0.0f;
        CK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
        s->force_ms += ms;
        s->force_launches += 1;
    }
    return step_conclude(s);
}

// ---- the host out of the step loop -----------------------------------------------------------------
// Between two searches nothing the host decides depends on a step's outcome unless that step divides a rod,
// fails, or moves a vertex beyond skin/2.  Up to `n_max` list steps are therefore enqueued back to back
// (k_force_eval + k_integrate each, no host wait); the first step after which the host has to act latches
// its iteration on the device (Status::stop_iter), the kernels of the steps behind it return at once, and ONE
// status block comes back per batch.  The accumulators of the block (max displacement, division list, error)
// are only ever written by the latching step, so they read exactly as after a single step.  Results are
// identical to stepping one by one: the same kernels see the same state.
static bool batchable(const crm_sim* s) {
    return s->batching && persistent(s) && !s->need_rebuild && s->have_lists && s->n_dirty == 0 && !s->sharded && s->n > 0 &&
           s->pending_div == 0 && !s->step_open;
}

static int step_batch(crm_sim* s, uint32_t n_max, uint32_t* done) {
    const uint64_t it0 = s->iteration;
    const uint32_t B = std::max<uint32_t>(1u, std::min<uint32_t>(n_max, s->batch_len));
    s->d.n = s->n + s->n_ghost;
    s->d.n_own = s->n;
    s->d.emax_limit_bits = 0u;
    s->d.skip = nullptr;
    s->async_search = false;
    CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
    CK(cudaMemsetAsync(s->latch_buf, 0, sizeof(uint32_t), s->stream));
    s->d.latch = s->latch_buf;
    uint32_t* const h_latch = reinterpret_cast<uint32_t*>(s->h_status + 1) + 24;  // pinned scratch behind the status
    const bool uniform = !s->lists_slow;
    const int pt = s->lists_morse == s->lists_mie ? -1 : (s->lists_mie ? CRM_POT_MIE : CRM_POT_MORSE);
    set_staging(s, false);
    const bool fuse = fusable(s);
    for (uint32_t k = 0; k < B; ++k) {
        s->d.iteration = it0 + k + 1;
        if (s->timers && k == 0) CK(cudaEventRecord(s->ev0, s->stream));
        if (fuse) s->d.vrec_out = s->vrec_alt;
        launch_list_step(s, uniform, pt, fuse, false);
        if (s->timers && k == 0) CK(cudaEventRecord(s->ev1, s->stream));
        if (fuse) {
            swap_vrec(s);
        } else {
            if (s->timers && k == 0) CK(cudaEventRecord(s->ev2, s->stream));
            launch_integrate(s->nv, s->d, s->stream);
            if (s->timers && k == 0) CK(cudaEventRecord(s->ev3, s->stream));
        }
    }
    s->d.latch = nullptr;
    s->launches += (fuse ? 0ull : 1ull) * B;
    s->bounds_fresh = false;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(s->h_status, s->d.status, sizeof(Status), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaMemcpyAsync(h_latch, s->latch_buf, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaEventRecord(s->ev_status, s->stream));
    CK(cudaEventSynchronize(s->ev_status));
    const uint32_t stop = *h_latch;
    const uint32_t n_done = stop ? (uint32_t)(stop - (uint32_t)it0) : B;
    if (n_done == 0 || n_done > B) return fail(CRM_ERR_CUDA, "batched steps: latch out of range (%u of %u)", n_done, B);
    if (fuse) {  // the steps behind the latch wrote nothing: the records of the last executed step are current
        if ((B - n_done) & 1u) swap_vrec(s);
        s->n_fused_steps += n_done;
    }
    s->iteration = it0 + n_done;
    s->d.iteration = s->iteration;
    s->rod_steps += (uint64_t)s->n * n_done;
    s->n_eval_steps += n_done;
    s->evals_since_build += n_done;
    s->steps_since_build += n_done;
    s->n_batches += 1;
    s->batch_len = stop ? std::max<uint32_t>(1u, n_done) : std::min<uint32_t>(2u * B, 64u);
    if (s->timers) {
        float ms = 0.0f;
        CK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
        s->force_ms += ms;
        s->force_launches += 1;
        if (!fuse) {
            CK(cudaEventElapsedTime(&ms, s->ev2, s->ev3));
            s->integrate_ms += ms;
            s->integrate_launches += 1;
        }
    }
    *done = n_done;
    return step_conclude(s);
}

static int step_once(crm_sim* s) {
    const int rc = step_enqueue(s, nullptr);
    return rc ? rc : step_finish(s);
}

// ---- the same for a slab of a sharded colony: batches of list steps INCLUDING their halo exchange -----------
// Between two collective rebuilds a sharded step is: repack the frozen halo selections -> send/recv with the two
// slab neighbours -> scatter the received records to the ghost slots -> k_force_eval -> k_integrate -> this
// slab's flags {divisions, rebuild wish, error} -> all-gather of the flags.  All of it is enqueued on the sim's
// stream from here (NCCL through crm_comm.cu), up to `n_max` steps back to back.  A one-thread kernel turns the
// gathered flags into the latch of step_batch: the first step in which ANY slab has a reason to stop latches in
// ALL slabs in that same step, the kernels of later steps return at once (their transfers still run, every
// rank keeps issuing the same collectives), and the host reads one status block per batch.  No step is ever
// computed with stale neighbour lists, so results equal the step-by-step protocol.
__global__ void k_latch_from_flags(const uint32_t* __restrict__ flags_all, int world, uint32_t* latch, uint32_t iteration) {
    if (threadIdx.x != 0) return;
    uint32_t any = 0u;
    for (int r = 0; r < world; ++r) any |= flags_all[4 * r] | flags_all[4 * r + 1] | flags_all[4 * r + 2];
    if (any && *latch == 0u) *latch = iteration;
}

// n_steps steps: batches of list steps where nothing needs the host, single steps otherwise
static int advance(crm_sim* s, uint64_t n_steps) {
    while (n_steps) {
        uint32_t done = 1;
        int rc;
        if (batchable(s)) rc = step_batch(s, (uint32_t)std::min<uint64_t>(n_steps, 1u << 20), &done);
        else rc = step_once(s);
        if (rc) return rc;
        n_steps -= done;
    }
    return 0;
}

static int download_state(crm_sim* s, float* pos, float* vel, crm_rod_params* par, crm_ident* ids = nullptr,
                          crm_ident* parents = nullptr) {
    const size_t m = (size_t)s->n * s->nv * 3;
    if (s->n == 0) return 0;
    Trace tr("download_state");
    // Everything is enqueued first, the small parameter blocks in front: the host threads merge them into the
    // caller's records while the positions and velocities (two staging buffers) are still crossing the bus; into
    // pinned caller buffers those copies run at link speed.
    const size_t np_ = par ? s->n : 0;
    PinnedLease lease(PinnedLease::part<GrowState>(np_) + PinnedLease::part<uint32_t>(np_));
    size_t off = 0;
    GrowState* const g = lease.take<GrowState>(off, np_);
    uint32_t* const nn = lease.take<uint32_t>(off, np_);
    if (par) {
        CK(cudaMemcpyAsync(g, s->d.grow, sizeof(GrowState) * s->n, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaMemcpyAsync(nn, s->d.nn, sizeof(uint32_t) * s->n, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaEventRecord(s->ev_status, s->stream));
    }
    if (pos) {
        launch_pack(s->nv, s->d.pos, s->stage, s->n, s->stream);
        CK(cudaMemcpyAsync(pos, s->stage, sizeof(float) * m, cudaMemcpyDeviceToHost, s->stream));
    }
    if (vel) {
        launch_pack(s->nv, s->d.vel, s->stage2, s->n, s->stream);
        CK(cudaMemcpyAsync(vel, s->stage2, sizeof(float) * m, cudaMemcpyDeviceToHost, s->stream));
    }
    tr.mark("enqueue");
    if (ids || parents) {  // host-only bookkeeping: copied while the device data is on its way
        parallel_for(s->n, [&](size_t a0, size_t a1) {
            if (ids) memcpy(ids + a0, s->ids.data() + a0, sizeof(crm_ident) * (a1 - a0));
            if (parents) memcpy(parents + a0, s->parents.data() + a0, sizeof(crm_ident) * (a1 - a0));
        });
        tr.mark("ids");
    }
    if (par) {
        CK(cudaEventSynchronize(s->ev_status));
        tr.mark("wait_params");
        parallel_for(s->n, [&](size_t a0, size_t a1) {
            for (size_t a = a0; a < a1; ++a) {
                par[a] = s->base[a];
                par[a].growth_rate = g[a].x;
                par[a].spring_length_threshold = g[a].y;
                par[a].spring_length = g[a].z;
                par[a].n_neighbors = nn[a];
            }
        });
        tr.mark("merge");
    }
    CK(cudaStreamSynchronize(s->stream));
    tr.mark("wait");
    return 0;
}

int crm_take_snapshot(crm_sim* s) {
    s->snaps.emplace_back();
    HostSnap& sn = s->snaps.back();
    sn.iteration = s->iteration;
    sn.n = s->n;
    const size_t m = (size_t)s->n * s->nv * 3;
    sn.pos.resize(m);
    sn.vel.resize(m);
    sn.par.resize(s->n);
    sn.ids.assign(s->ids.begin(), s->ids.begin() + s->n);
    sn.parents.assign(s->parents.begin(), s->parents.begin() + s->n);
    sn.gslot.assign(s->gslot.begin(), s->gslot.begin() + s->n);
    sn.sim_of = s->sim_of;
    return download_state(s, sn.pos.data(), sn.vel.data(), sn.par.data());
}

// ====================================================================================
// C ABI
extern "C" {

int crm_abi_version(void) { return CRM_ABI_VERSION; }
const char* crm_last_error(void) { return g_err.c_str(); }

int crm_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        fail(CRM_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return -1;
    }
    return n;
}

int crm_device_mem_info(int device, size_t* free_bytes, size_t* total_bytes) {
    CK(cudaSetDevice(device));
    size_t f = 0, t = 0;
    CK(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return 0;
}

int crm_save_schedule(const crm_config* config, uint64_t* out, uint32_t cap, uint64_t* n_iterations) {
    std::vector<uint64_t> it;
    uint64_t mx = 0;
    std::string err;
    int rc = build_save_schedule(config, it, mx, err);
    if (rc) return fail(rc, "%s", err.c_str());
    for (size_t i = 0; i < it.size() && i < cap; ++i) out[i] = it[i];
    if (n_iterations) *n_iterations = mx;
    return (int)it.size();
}

int crm_sim_create(const crm_config* c, int device, crm_sim** out) {
    if (!c || !out) return fail(CRM_ERR_INVALID, "null argument");
    if (c->n_voxels[0] == 0 || c->n_voxels[1] == 0 || !(c->domain_size[0] > 0) || !(c->domain_size[1] > 0) ||
        !(c->domain_height > 0))
        return fail(CRM_ERR_INVALID, "invalid domain configuration");
    std::vector<uint64_t> it;
    uint64_t mx = 0;
    std::string err;
    int rc = build_save_schedule(c, it, mx, err);
    if (rc) return fail(rc, "%s", err.c_str());
    int ndev = crm_device_count();
    if (ndev <= 0) return fail(CRM_ERR_CUDA, "no CUDA device available (cr_mech_coli_b200 has no CPU fallback): %s", g_err.c_str());
    if (device < 0 || device >= ndev) return fail(CRM_ERR_INVALID, "device %d out of range (%d visible)", device, ndev);
    CK(cudaSetDevice(device));
    crm_sim* s = new crm_sim();
    s->cfg = *c;
    s->device = device;
    s->save_iter = it;
    s->max_iter = mx;
    s->vox_counter.assign((size_t)c->n_voxels[0] * c->n_voxels[1], 0);
    const char* ex = getenv("CRM_B200_EXACT_MATH");
    s->fast_math = !(ex && ex[0] == '1');
    if (const char* sk = getenv("CRM_B200_SKIN")) s->skin = std::max(0.0f, (float)atof(sk));
    if (const char* pt = getenv("CRM_B200_PATCH")) s->patching = pt[0] != '0';
    CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CK(cudaMalloc(&s->d.status, sizeof(Status)));
    CK(cudaMemset(s->d.status, 0, sizeof(Status)));
    s->h_div_pinned_cap = DIV_SORT_MAX;
    CK(cudaMallocHost(&s->h_div_pinned, sizeof(DivRecord) * s->h_div_pinned_cap));
    CK(cudaMalloc(&s->latch_buf, 256));
    CK(cudaMemset(s->latch_buf, 0, 256));
    CK(cudaMallocHost(&s->h_status, sizeof(Status) + 128));  // + pinned scratch (message headers)
    CK(cudaEventCreate(&s->ev0));
    CK(cudaEventCreate(&s->ev1));
    CK(cudaEventCreate(&s->ev2));
    CK(cudaEventCreate(&s->ev3));
    CK(cudaEventCreateWithFlags(&s->ev_status, cudaEventDisableTiming));
    if (const char* bt = getenv("CRM_B200_BATCH")) s->batching = bt[0] != '0';
    if (const char* ab = getenv("CRM_B200_ASYNC_BOUNDS")) s->async_bounds = ab[0] != '0';
    if (const char* fu = getenv("CRM_B200_FUSED")) s->fused = fu[0] != '0';
    s->d.lx = c->domain_size[0];
    s->d.ly = c->domain_size[1];
    s->d.lz = c->domain_height;
    s->d.unx = (int)c->n_voxels[0];
    s->d.uny = (int)c->n_voxels[1];
    s->d.uvx_w = c->domain_size[0] / (float)c->n_voxels[0];
    s->d.uvy_w = c->domain_size[1] / (float)c->n_voxels[1];
    s->d.gel_pressure = c->gel_pressure;
    s->d.surface_friction = c->surface_friction;
    s->d.surface_friction_distance = c->surface_friction_distance;
    s->d.dt = c->dt;
    s->d.slab_lo = -INFINITY;
    s->d.slab_hi = INFINITY;
    s->d.seed = c->rng_seed;
    s->d.ab_hist = 2;
    s->d.spring_w = 1.0f;
    if (const char* so = getenv("CRM_B200_SOLVER_ORDER")) s->d.ab_hist = std::min(2, std::max(0, atoi(so)));
    if (const char* sw = getenv("CRM_B200_SPRING_END_WEIGHT")) s->d.spring_w = (float)atof(sw);
    *out = s;
    return 0;
}

void crm_sim_destroy(crm_sim* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    // (the per-rod arrays are parts of the arena)
    void* ptrs[] = {s->arena, s->d.cell_start, s->d.cell_count, s->select_tmp, s->select_out, s->select_count, s->halo_sel[0],
                    s->halo_sel[1], s->sim_fail, s->d.status, s->latch_buf};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (s->h_status) cudaFreeHost(s->h_status);
    if (s->flags_host) cudaFreeHost(s->flags_host);
    if (s->h_div_pinned) cudaFreeHost(s->h_div_pinned);
    if (s->flags_dev) cudaFree(s->flags_dev);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->ev2) cudaEventDestroy(s->ev2);
    if (s->ev3) cudaEventDestroy(s->ev3);
    if (s->ev_status) cudaEventDestroy(s->ev_status);
    if (s->ev_t0) cudaEventDestroy(s->ev_t0);
    if (s->ev_t1) cudaEventDestroy(s->ev_t1);
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->stream && s->own_stream) cudaStreamDestroy(s->stream);
    delete s;
}

int crm_sim_set_skin(crm_sim* s, float skin) {
    if (!s || !(skin >= 0.0f)) return fail(CRM_ERR_INVALID, "skin must be >= 0");
    if (s->n) return fail(CRM_ERR_INVALID, "set the skin before the agents");
    s->skin = skin;
    s->need_rebuild = true;
    return 0;
}

// Why builds of the neighbour structure were abandoned: reason bits (1 more than 32 / 16 neighbour rods, 2 / 4 more
// than 128 queued own / ext evaluations per vertex, 8 more than 8 sensing candidates per vertex), how often, and the
// longest list of the live structure.
int crm_sim_list_overflows(const crm_sim* s, uint32_t* reasons, uint64_t* count, uint32_t* longest_list) {
    if (reasons) *reasons = s->overflow_reasons;
    if (count) *count = s->n_overflows;
    if (longest_list) *longest_list = s->lists_maxl;
    return 0;
}

// totals over the rows of the live structure (diagnostics): queued own / ext evaluations, sensing candidates,
// listed rods; rows of owned rods only
__global__ void k_list_totals(const uint32_t* __restrict__ pnq, const uint32_t* __restrict__ sorted, uint32_t n, uint32_t n_own,
                              int lp, unsigned long long* __restrict__ out) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long o = 0, e = 0, q = 0, c = 0;
    if (g < n && sorted[g] < n_own) {
        for (int l = 0; l < lp; ++l) {
            const uint32_t w = pnq[(size_t)g * lp + l];
            if ((w >> 24) == 0xffu) continue;
            o += w & 0xffu; e += (w >> 8) & 0xffu; q += (w >> 16) & 0xffu;
            if (l == 0) c += w >> 24;
        }
    }
    for (int k = 16; k > 0; k >>= 1) {
        o += __shfl_xor_sync(0xffffffffu, o, k); e += __shfl_xor_sync(0xffffffffu, e, k);
        q += __shfl_xor_sync(0xffffffffu, q, k); c += __shfl_xor_sync(0xffffffffu, c, k);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(out, o); atomicAdd(out + 1, e); atomicAdd(out + 2, q); atomicAdd(out + 3, c); }
}

int crm_sim_list_totals(crm_sim* s, uint64_t out[6]) {
    if (!s || !out) return fail(CRM_ERR_INVALID, "null argument");
    out[0] = out[1] = out[2] = out[3] = 0;
    out[4] = s->have_lists ? s->n_long : 0;
    out[5] = s->n_two_class_steps;
    if (!s->have_lists || s->n == 0) return 0;
    CK(cudaSetDevice(s->device));
    unsigned long long* d_out = reinterpret_cast<unsigned long long*>(s->stage);  // (scratch: nothing else is in flight)
    CK(cudaStreamSynchronize(s->stream));
    CK(cudaMemsetAsync(d_out, 0, 4 * sizeof(unsigned long long), s->stream));
    const uint32_t n = s->n + s->n_ghost;
    k_list_totals<<<blocks_for(n, 256), 256, 0, s->stream>>>(s->d.pnq, s->d.sorted, n, s->n, s->lp, d_out);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, d_out, 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    return 0;
}

int crm_sim_rebuild_reasons(const crm_sim* s, uint64_t out[8]) {
    if (!s || !out) return fail(CRM_ERR_INVALID, "null argument");
    for (int k = 0; k < 8; ++k) out[k] = s->why_rebuild[k];
    return 0;
}

int crm_sim_list_stats(const crm_sim* s, uint64_t* n_builds, uint64_t* n_eval_steps) {
    if (n_builds) *n_builds = s->n_builds;
    if (n_eval_steps) *n_eval_steps = s->n_eval_steps;
    return 0;
}

// Divisions absorbed by patching the live structure (no new search), and the switch for it.
uint64_t crm_sim_patch_count(const crm_sim* s) { return s ? s->n_patches : 0; }
int crm_sim_set_patching(crm_sim* s, int on) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    s->patching = on != 0;
    return 0;
}

// Open restatement decisions as run-time parameters (the same switches exist in the oracle).
// stored_increments: 0 Euler, 1 Adams-Bashforth 2, 2 Adams-Bashforth 3 (cellular_raza's
// `mechanics_solver_order` kwarg of run_main!, not passed at reference src/simulation.rs:283-292).
int crm_sim_set_solver_order(crm_sim* s, int stored_increments) {
    if (!s || stored_increments < 0 || stored_increments > 2) return fail(CRM_ERR_INVALID, "stored increments must be 0, 1 or 2");
    if (s->n) return fail(CRM_ERR_INVALID, "set the solver order before the agents");
    s->d.ab_hist = stored_increments;
    return 0;
}
// weight of the spring force on each end of a segment: 1 (docs/source/model.rst:32-36) or 1/2
int crm_sim_set_spring_end_weight(crm_sim* s, float w) {
    if (!s || !(w > 0.0f)) return fail(CRM_ERR_INVALID, "spring end weight must be > 0");
    s->d.spring_w = w;
    return 0;
}

int crm_sim_set_exact_math(crm_sim* s, int exact) {
    s->fast_math = !exact;
    return 0;
}

int crm_sim_set_agents(crm_sim* s, const float* pos, const float* vel, const crm_rod_params* par, uint32_t n,
                       uint32_t nv) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    if (!nv_supported((int)nv)) return fail(CRM_ERR_INVALID, "n_vertices=%u unsupported (2..16)", nv);
    if (s->nv != 0 && s->nv != (int)nv) return fail(CRM_ERR_INVALID, "n_vertices cannot change on a live sim");
    CK(cudaSetDevice(s->device));
    s->nv = (int)nv;
    s->lp = lanes_per_rod((int)nv);
    s->n = 0;
    Trace tr("set_agents");
    int rc = crm_reserve(s, std::max<uint32_t>(n + n / 4 + 16, 1024));
    if (rc) return rc;
    tr.mark("reserve");
    s->iteration = 0;
    s->snaps.clear();
    std::fill(s->vox_counter.begin(), s->vox_counter.end(), 0);
    s->base.resize(n);
    s->ids.resize(n);
    s->parents.resize(n);
    s->gslot.resize(n);
    s->n_ghost = 0;
    s->pending_div = 0;
    s->sim_of.clear();
    s->sim_rods.clear();
    s->n_sims = 1;
    s->d.tile = nullptr;
    s->d.sim_fail = nullptr;
    s->rkey_on = false;
    s->d.rkey = nullptr;
    s->vox_counter.assign((size_t)s->cfg.n_voxels[0] * s->cfg.n_voxels[1], 0);
    const size_t m = (size_t)n * nv * 3;
    if (n) {
        // positions and velocities go first: from pinned caller buffers these copies run while the host
        // threads below convert the parameter records
        CK(cudaMemcpyAsync(s->stage, pos, sizeof(float) * m, cudaMemcpyHostToDevice, s->stream));
        launch_unpack(s->nv, s->stage, s->d.pos, n, s->stream);
        CK(cudaMemcpyAsync(s->stage2, vel, sizeof(float) * m, cudaMemcpyHostToDevice, s->stream));
        launch_unpack(s->nv, s->stage2, s->d.vel, n, s->stream);
    }
    tr.mark("host_vectors+enqueue_pos_vel");
    // converted records are staged page-locked and leave in chunks: the copy of chunk k runs while the host
    // threads convert chunk k+1 (and behind the positions / velocities enqueued above)
    PinnedLease lease(PinnedLease::part<PotParams>(n) + PinnedLease::part<MechParams>(n) + PinnedLease::part<GrowState>(n) +
                      PinnedLease::part<CycleParams>(n) + PinnedLease::part<uint32_t>(n));
    size_t off = 0;
    PotParams* const pot = lease.take<PotParams>(off, n);
    MechParams* const mech = lease.take<MechParams>(off, n);
    GrowState* const grow = lease.take<GrowState>(off, n);
    CycleParams* const cyc = lease.take<CycleParams>(off, n);
    uint32_t* const nn = lease.take<uint32_t>(off, n);
    std::atomic<uint32_t> bad{UINT32_MAX}, rbits{0}, cbits{0};
    const size_t CH = (size_t)1 << 18;
    for (size_t c0 = 0; c0 < n; c0 += CH) {
        const size_t c1 = std::min<size_t>(n, c0 + CH);
        parallel_for(c1 - c0, [&](size_t b0, size_t b1) {
            const size_t a0 = c0 + b0, a1 = c0 + b1;
            ParamCache k;
            float rmax = 0.0f, cmax = 0.0f;
            for (size_t a = a0; a < a1; ++a) {
                s->base[a] = par[a];
                s->ids[a] = crm_ident{CRM_ID_INITIAL, 0, (uint64_t)a};
                s->gslot[a] = a;
                s->parents[a] = crm_ident{-1, 0, 0};
                if (par[a].potential != CRM_POT_MORSE && par[a].potential != CRM_POT_MIE) {
                    uint32_t cur = bad.load();
                    while ((uint32_t)a < cur && !bad.compare_exchange_weak(cur, (uint32_t)a)) {}
                    continue;
                }
                fill_device_params_cached(par[a], k, pot[a], mech[a], grow[a], cyc[a]);
                nn[a] = par[a].n_neighbors;
                rmax = std::max(rmax, par[a].radius);
                cmax = std::max(cmax, par[a].cutoff);
            }
            // non-negative floats order like their bit patterns
            auto amax = [](std::atomic<uint32_t>& t, float v) {
                if (!(v > 0.0f)) return;
                const uint32_t b = __builtin_bit_cast(uint32_t, v);
                uint32_t cur = t.load();
                while (b > cur && !t.compare_exchange_weak(cur, b)) {}
            };
            amax(rbits, rmax);
            amax(cbits, cmax);
        });
        if (bad.load() != UINT32_MAX) break;
        const size_t m_ = c1 - c0;
        CK(cudaMemcpyAsync(s->d.pot + c0, pot + c0, sizeof(PotParams) * m_, cudaMemcpyHostToDevice, s->stream));
        CK(cudaMemcpyAsync(s->d.mech + c0, mech + c0, sizeof(MechParams) * m_, cudaMemcpyHostToDevice, s->stream));
        CK(cudaMemcpyAsync(s->d.grow + c0, grow + c0, sizeof(GrowState) * m_, cudaMemcpyHostToDevice, s->stream));
        CK(cudaMemcpyAsync(s->d.cyc + c0, cyc + c0, sizeof(CycleParams) * m_, cudaMemcpyHostToDevice, s->stream));
        CK(cudaMemcpyAsync(s->d.nn + c0, nn + c0, sizeof(uint32_t) * m_, cudaMemcpyHostToDevice, s->stream));
    }
    tr.mark("convert+param_copies");
    if (bad.load() != UINT32_MAX) {
        CK(cudaStreamSynchronize(s->stream));
        return fail(CRM_ERR_INVALID, "rod %u: unknown potential %d", bad.load(), par[bad.load()].potential);
    }
    const float rmax = __builtin_bit_cast(float, rbits.load()), cmax = __builtin_bit_cast(float, cbits.load());
    s->rsearch = std::max(cmax, 2.0f * rmax * 0.70710678f) * 1.0001f;
    s->d.rsearch = search_radius(s);  // the bounds kernels bake it into the cull thresholds
    s->d.skin = s->skin;
    s->need_rebuild = true;
    s->have_lists = false;
    s->holdoff = 0;
    s->backoff = 8;
    s->evals_since_build = 0;
    s->steps_since_build = 0;
    if (n) {
        CK(cudaMemsetAsync(s->d.nhist, 0, n, s->stream));
        CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
        s->n = n;
        s->d.n = s->d.n_own = n;
        launch_bounds_kernel(s->nv, s->d, s->stream);
        s->launches += 3;
        CK(cudaGetLastError());
        if ((rc = crm_read_status(s))) return rc;
        tr.mark("bounds+wait");
        s->emax = __builtin_bit_cast(float, s->h_status->emax_bits);
        s->bounds_fresh = true;
        if (!std::isfinite(s->emax)) return fail(CRM_ERR_NUMERIC, "CalcError: non-finite initial positions");
    }
    return 0;
}

// Batched independent sims (crm_fit ensembles, reference src/crm_fit/optimize.rs:75-95 runs one
// run_simulation_with_agents per parameter point): all rods live in one device state, every sim is
// a tile of the cell list, so rods of different sims never interact; identifiers, voxel counters
// and snapshots are kept per sim.
int crm_sim_set_agents_batched(crm_sim* s, const float* pos, const float* vel, const crm_rod_params* par,
                               const uint32_t* sim_of, uint32_t n, uint32_t nv, uint32_t n_sims) {
    if (!s || !sim_of || n_sims == 0) return fail(CRM_ERR_INVALID, "null argument");
    if (s->sharded) return fail(CRM_ERR_INVALID, "batched sims cannot be sharded (replicas only)");
    for (uint32_t a = 0; a < n; ++a)
        if (sim_of[a] >= n_sims) return fail(CRM_ERR_INVALID, "rod %u: sim index %u out of range", a, sim_of[a]);
    int rc = crm_sim_set_agents(s, pos, vel, par, n, nv);
    if (rc) return rc;
    s->n_sims = n_sims;
    s->sim_of.assign(sim_of, sim_of + n);
    s->sim_rods.assign(n_sims, 0);
    s->sim_rods0.assign(n_sims, 0);
    for (uint32_t a = 0; a < n; ++a) s->sim_rods0[sim_of[a]]++;
    s->vox_counter.assign((size_t)s->cfg.n_voxels[0] * s->cfg.n_voxels[1] * n_sims, 0);
    for (uint32_t a = 0; a < n; ++a) {  // Initial(index inside the own sim)
        const uint64_t k = s->sim_rods[sim_of[a]]++;
        s->ids[a] = crm_ident{CRM_ID_INITIAL, 0, k};
        s->gslot[a] = k;
    }
    if (n) CK(cudaMemcpyAsync(s->tile_buf, sim_of, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    s->d.tile = s->tile_buf;
    s->rkey_on = true;  // a member draws the numbers of its stand-alone run (key = slot inside the member)
    return crm_upload_rkeys(s, 0, n);
}

// Error isolation of batched sims: a BoundaryError / CalcError stops only the sim it happens in
// (the reference evaluates every parameter point in its own run and maps an error to the error cost of
// that member alone, src/crm_fit/predict.rs:224-228,249-256).
int crm_sim_isolate_sim_errors(crm_sim* s, int on) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    if (s->sim_of.empty()) return fail(CRM_ERR_INVALID, "not a batched sim (crm_sim_set_agents_batched)");
    CK(cudaSetDevice(s->device));
    if (on) {
        if (s->sim_fail) CK(cudaFree(s->sim_fail));
        CK(cudaMalloc(&s->sim_fail, sizeof(int) * s->n_sims));
        CK(cudaMemsetAsync(s->sim_fail, 0, sizeof(int) * s->n_sims, s->stream));
        CK(cudaStreamSynchronize(s->stream));
    }
    s->d.sim_fail = on ? s->sim_fail : nullptr;
    return 0;
}

// Per sim: 0 | CRM_ERR_BOUNDARY | CRM_ERR_NUMERIC (isolated errors) | CRM_ERR_DIVISION (the sim has
// divided; only reported, divisions are legal outside the cost evaluation).
int crm_sim_get_sim_status(crm_sim* s, int32_t* out) {
    if (!s || !out) return fail(CRM_ERR_INVALID, "null argument");
    CK(cudaSetDevice(s->device));
    for (uint32_t m = 0; m < s->n_sims; ++m) out[m] = 0;
    if (s->d.sim_fail) {
        CK(cudaMemcpyAsync(out, s->sim_fail, sizeof(int) * s->n_sims, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
    }
    for (uint32_t m = 0; m < s->n_sims && m < s->sim_rods0.size(); ++m)
        if (out[m] == 0 && s->sim_rods[m] != s->sim_rods0[m]) out[m] = CRM_ERR_DIVISION;
    return 0;
}

// ---- cost of a batched ensemble, reduced on the device -------------------------------------------
// One block per sim, fixed-order tree: cost[m] += sum over the sim's rods and vertices of the squared xy
// distance to targets[r][rod index inside the sim][vertex][0..1] (reference src/crm_fit/predict.rs:258-285).
__global__ void __launch_bounds__(128) k_cost(const float* __restrict__ pos, int nv, int lp, const uint32_t* __restrict__ start,
                                              const float* __restrict__ target, uint32_t n_agents, float* __restrict__ cost) {
    __shared__ float red[128];
    const uint32_t m = blockIdx.x;
    const uint32_t s0 = start[m], s1 = start[m + 1];
    float acc = 0.0f;
    for (uint32_t i = s0 * nv + threadIdx.x; i < s1 * nv; i += 128) {
        const uint32_t rod = i / nv, v = i % nv;
        const float x = pos[(size_t)rod * 3 * lp + v], y = pos[(size_t)rod * 3 * lp + lp + v];
        const float* t = target + ((size_t)(rod - s0) * nv + v) * 2;
        const float dx = x - t[0], dy = y - t[1];
        acc += dx * dx + dy * dy;
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) cost[m] += red[0];
    (void)n_agents;
}

// Steps to t_max without snapshots and accumulates, at the given iterations, every sim's cost against
// targets[n_results][n_agents][n_vertices][2] (host, f32).  Errors are isolated per sim; out_status as
// crm_sim_get_sim_status.  Rods must be grouped by sim in the order of set_agents_batched.
int crm_sim_run_costs(crm_sim* s, const float* targets, uint32_t n_results, uint32_t n_agents,
                      const uint64_t* result_iterations, float* out_costs, int32_t* out_status) {
    if (!s || !targets || !result_iterations || !out_costs) return fail(CRM_ERR_INVALID, "null argument");
    if (s->sim_of.empty()) return fail(CRM_ERR_INVALID, "not a batched sim (crm_sim_set_agents_batched)");
    CK(cudaSetDevice(s->device));
    std::vector<uint32_t> start(s->n_sims + 1, 0);
    for (uint32_t a = 0; a < s->n; ++a) {
        if (a && s->sim_of[a] < s->sim_of[a - 1]) return fail(CRM_ERR_INVALID, "rods are not grouped by sim");
        start[s->sim_of[a] + 1]++;
    }
    for (uint32_t m = 0; m < s->n_sims; ++m) {
        if (start[m + 1] > n_agents) return fail(CRM_ERR_INVALID, "sim %u has %u rods, the data %u", m, start[m + 1], n_agents);
        start[m + 1] += start[m];
    }
    for (uint32_t r = 1; r < n_results; ++r)
        if (result_iterations[r] <= result_iterations[r - 1]) return fail(CRM_ERR_INVALID, "result iterations must increase");
    int rc = crm_sim_isolate_sim_errors(s, 1);
    if (rc) return rc;
    float *d_target = nullptr, *d_cost = nullptr;
    uint32_t* d_start = nullptr;
    const size_t per_result = (size_t)n_agents * s->nv * 2;
    auto cleanup = [&]() {
        cudaFree(d_target);
        cudaFree(d_cost);
        cudaFree(d_start);
    };
    if (cudaMalloc(&d_target, sizeof(float) * per_result * std::max(1u, n_results)) != cudaSuccess ||
        cudaMalloc(&d_cost, sizeof(float) * s->n_sims) != cudaSuccess ||
        cudaMalloc(&d_start, sizeof(uint32_t) * (s->n_sims + 1)) != cudaSuccess) {
        cleanup();
        cudaGetLastError();
        return fail(CRM_ERR_NOMEM, "cost evaluation: out of device memory");
    }
    cudaMemcpyAsync(d_target, targets, sizeof(float) * per_result * n_results, cudaMemcpyHostToDevice, s->stream);
    cudaMemcpyAsync(d_start, start.data(), sizeof(uint32_t) * (s->n_sims + 1), cudaMemcpyHostToDevice, s->stream);
    cudaMemsetAsync(d_cost, 0, sizeof(float) * s->n_sims, s->stream);
    uint32_t next = 0;
    auto accumulate = [&]() {
        while (next < n_results && result_iterations[next] < s->iteration) next++;
        if (next < n_results && result_iterations[next] == s->iteration) {
            k_cost<<<s->n_sims, 128, 0, s->stream>>>(s->d.pos, s->nv, s->lp, d_start, d_target + per_result * next, n_agents, d_cost);
            s->launches += 1;
            next++;
        }
    };
    accumulate();
    while (s->iteration < s->max_iter) {  // to t_max like the reference: a late error still costs the member
        while (next < n_results && result_iterations[next] <= s->iteration) next++;
        const uint64_t target = next < n_results ? std::min<uint64_t>(result_iterations[next], s->max_iter) : s->max_iter;
        if ((rc = advance(s, std::max<uint64_t>(1, target - s->iteration)))) {
            cleanup();
            return rc;
        }
        accumulate();
    }
    cudaMemcpyAsync(out_costs, d_cost, sizeof(float) * s->n_sims, cudaMemcpyDeviceToHost, s->stream);
    cudaError_t e = cudaStreamSynchronize(s->stream);
    cleanup();
    if (e != cudaSuccess) return fail(CRM_ERR_CUDA, "cost evaluation: %s", cudaGetErrorString(e));
    if (out_status && (rc = crm_sim_get_sim_status(s, out_status))) return rc;
    return 0;
}

int crm_sim_get_sim_index(const crm_sim* s, uint32_t* out) {
    for (uint32_t a = 0; a < s->n; ++a) out[a] = s->sim_of.empty() ? 0u : s->sim_of[a];
    return 0;
}

int crm_sim_snapshot_sim_index(const crm_sim* s, uint32_t k, uint32_t* out) {
    if (k >= s->snaps.size()) return fail(CRM_ERR_INVALID, "snapshot %u out of range", k);
    const HostSnap& sn = s->snaps[k];
    for (uint32_t a = 0; a < sn.n; ++a) out[a] = sn.sim_of.empty() ? 0u : sn.sim_of[a];
    return 0;
}

int crm_sim_step(crm_sim* s, uint32_t n_steps) {
    CK(cudaSetDevice(s->device));
    return advance(s, n_steps);
}

// The step in two halves for callers that order their own work (halo transport, collectives) on
// the sim's stream: begin enqueues the step and writes [n_div, reason to rebuild] to dev_flags on
// the stream; end waits for the stream once and does the host's part of the step.
void* crm_sim_stream(const crm_sim* s) { return s ? (void*)s->stream : nullptr; }

// Run on a stream the caller owns (e.g. one of torch's pool streams, so that tensors used with it
// never outlive it); the sim's own stream is released.
int crm_sim_set_stream(crm_sim* s, void* stream) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    if (s->step_open) return fail(CRM_ERR_INVALID, "a step is in flight");
    CK(cudaSetDevice(s->device));
    CK(cudaStreamSynchronize(s->stream));
    if (s->own_stream && s->stream) CK(cudaStreamDestroy(s->stream));
    s->stream = (cudaStream_t)stream;
    s->own_stream = false;
    return 0;
}

int crm_sim_step_begin(crm_sim* s, uint32_t* dev_flags) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    CK(cudaSetDevice(s->device));
    return step_enqueue(s, dev_flags);
}

int crm_sim_step_end(crm_sim* s) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    CK(cudaSetDevice(s->device));
    return step_finish(s);
}

// Up to n_max sharded list steps with their halo exchanges (see k_latch_from_flags above).  The four buffers are the
// caller's device message buffers of crm_sim_halo_pack_msgs / crm_sim_set_ghosts_msgs (counts as of the last
// rebuild step).  *done = steps advanced; flags_out [4 * world] = {divisions, rebuild wish, -error, 0} of every slab
// for the last of them.  Collective: every slab calls it with the same n_max.  A model error of THIS slab is
// returned after flags_out has been filled, so that the caller can raise in every slab.
int crm_sim_steps_sharded(crm_sim* s, uint32_t n_max, void* send_lo, void* send_hi, void* recv_lo, void* recv_hi,
                          uint32_t* done, uint32_t* flags_out) {
    if (!s || !done || !flags_out) return fail(CRM_ERR_INVALID, "null argument");
    if (!s->comm) return fail(CRM_ERR_INVALID, "no communicator (crm_sim_set_comm)");
    if (!s->sharded || s->pending_div || s->step_open) return fail(CRM_ERR_INVALID, "not a slab ready to step");
    if (!persistent(s) || !s->have_lists || s->want_rebuild || s->n_dirty)
        return fail(CRM_ERR_INVALID, "no live neighbour structure: a collective rebuild step comes first");
    CK(cudaSetDevice(s->device));
    const int world = crm_comm_world(s->comm), rank = crm_comm_rank(s->comm);
    if (!s->flags_dev) {
        CK(cudaMalloc(&s->flags_dev, sizeof(uint32_t) * 4 * (world + 1)));
        CK(cudaMemset(s->flags_dev, 0, sizeof(uint32_t) * 4 * (world + 1)));
        CK(cudaMallocHost(&s->flags_host, sizeof(uint32_t) * 4 * world));
    }
    *done = 0;
    if (n_max == 0) return 0;
    s->need_rebuild = false;  // (crm_sim_keep_lists: all slabs agreed after the last step)
    const uint64_t it0 = s->iteration;
    const uint32_t B = std::max<uint32_t>(1u, std::min<uint32_t>(n_max, s->batch_len));
    const size_t rb = crm_sim_halo_record_bytes(s);
    const size_t sl = rank > 0 ? 64 + s->halo_sel_n[0] * rb : 0, sh = rank + 1 < world ? 64 + s->halo_sel_n[1] * rb : 0;
    const size_t rl = rank > 0 ? 64 + s->ghost_n[0] * rb : 0, rh = rank + 1 < world ? 64 + s->ghost_n[1] * rb : 0;
    if ((sl && !send_lo) || (sh && !send_hi) || (rl && !recv_lo) || (rh && !recv_hi)) return fail(CRM_ERR_INVALID, "missing message buffer");
    s->d.n = s->n + s->n_ghost;
    s->d.n_own = s->n;
    CK(cudaMemsetAsync(s->d.status, 0, sizeof(Status), s->stream));
    CK(cudaMemsetAsync(s->latch_buf, 0, sizeof(uint32_t), s->stream));
    s->d.latch = s->latch_buf;
    uint32_t* const h_latch = reinterpret_cast<uint32_t*>(s->h_status + 1) + 24;
    const bool uniform = !s->lists_slow;
    const int pt = s->lists_morse == s->lists_mie ? -1 : (s->lists_mie ? CRM_POT_MIE : CRM_POT_MORSE);
    set_staging(s, false);
    const bool fuse = fusable(s);
    uint32_t swaps = 0;
    // Per step: ONE pack launch (both messages + the flags of the step just integrated), ONE NCCL group (halo
    // send/recv + all-gather of those flags), ONE unpack launch (ghosts + the latch from the gathered flags of the
    // previous step), force, integrate.  The flags of the last step are gathered behind the loop.
    int rc = 0;
    uint32_t* const my_flags = s->flags_dev;
    uint32_t* const all_flags = s->flags_dev + 4;
    void* const slo = rank > 0 ? send_lo : nullptr;
    void* const shi = rank + 1 < world ? send_hi : nullptr;
    void* const rlo = rank > 0 ? recv_lo : nullptr;
    void* const rhi = rank + 1 < world ? recv_hi : nullptr;
    for (uint32_t k = 0; k < B && !rc; ++k) {
        s->d.iteration = it0 + k + 1;
        // (k == 0: the status block has just been zeroed, the flags read "nothing happened")
        if ((rc = crm_shard_pack_all(s, slo, shi, my_flags))) break;
        if ((rc = crm_comm_exchange(s->comm, send_lo, sl, recv_lo, rl, send_hi, sh, recv_hi, rh, my_flags, all_flags, 16, s->stream))) break;
        if ((rc = crm_shard_unpack_all(s, rlo, rhi, all_flags, world, s->latch_buf, (uint32_t)(it0 + k)))) break;
        if (s->timers && k == 0) CK(cudaEventRecord(s->ev0, s->stream));
        if (fuse) s->d.vrec_out = s->vrec_alt;
        launch_list_step(s, uniform, pt, fuse, false);
        if (s->timers && k == 0) CK(cudaEventRecord(s->ev1, s->stream));
        if (fuse) {
            swap_vrec(s);
            ++swaps;
        } else {
            if (s->timers && k == 0) CK(cudaEventRecord(s->ev2, s->stream));
            launch_integrate(s->nv, s->d, s->stream);
            if (s->timers && k == 0) CK(cudaEventRecord(s->ev3, s->stream));
        }
        s->launches += fuse ? 0 : 1;
    }
    if (!rc) {  // the flags of the last step (the host's copy; the latch of a full batch stays 0 unless it fired)
        k_status_flags<<<1, 1, 0, s->stream>>>(s->d.status, s->skin, 0u, my_flags);
        rc = crm_comm_allgather(s->comm, my_flags, all_flags, 16, s->stream);
        k_latch_from_flags<<<1, 32, 0, s->stream>>>(all_flags, world, s->latch_buf, (uint32_t)(it0 + B));
        s->launches += 2;
    }
    s->d.latch = nullptr;
    s->bounds_fresh = false;
    if (rc) return rc;  // (a transport failure: nothing sensible can follow)
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(s->h_status, s->d.status, sizeof(Status), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaMemcpyAsync(h_latch, s->latch_buf, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaMemcpyAsync(s->flags_host, s->flags_dev + 4, sizeof(uint32_t) * 4 * world, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaEventRecord(s->ev_status, s->stream));
    CK(cudaEventSynchronize(s->ev_status));
    const uint32_t stop = *h_latch;
    const uint32_t n_done = stop ? (uint32_t)(stop - (uint32_t)it0) : B;
    if (n_done == 0 || n_done > B) return fail(CRM_ERR_CUDA, "sharded steps: latch out of range (%u of %u)", n_done, B);
    if (fuse) {
        if ((swaps - n_done) & 1u) swap_vrec(s);
        s->n_fused_steps += n_done;
    }
    s->iteration = it0 + n_done;
    s->d.iteration = s->iteration;
    s->rod_steps += (uint64_t)s->n * n_done;
    s->n_eval_steps += n_done;
    s->evals_since_build += n_done;
    s->steps_since_build += n_done;
    s->n_batches += 1;
    s->batch_len = stop ? std::max<uint32_t>(1u, n_done) : std::min<uint32_t>(2u * B, 64u);
    if (s->timers) {
        float ms = 0.0f;
        CK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
        s->force_ms += ms;
        s->force_launches += 1;
        if (!fuse) {
            CK(cudaEventElapsedTime(&ms, s->ev2, s->ev3));
            s->integrate_ms += ms;
            s->integrate_launches += 1;
        }
    }
    *done = n_done;
    memcpy(flags_out, s->flags_host, sizeof(uint32_t) * 4 * world);
    return step_conclude(s);
}

int crm_sim_step_timed(crm_sim* s, uint32_t n_steps, double* ms_out) {
    CK(cudaSetDevice(s->device));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    CK(cudaEventRecord(a, s->stream));
    const int rc = advance(s, n_steps);
    CK(cudaEventRecord(b, s->stream));
    CK(cudaEventSynchronize(b));
    float ms = 0.0f;
    CK(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    if (ms_out) *ms_out = ms;
    return rc;
}

// Device-side stopwatch on the sim's stream for step sequences driven from above the ABI
// (sharded runs): start records an event, stop records another one and returns the elapsed ms.
int crm_sim_timer_start(crm_sim* s) {
    CK(cudaSetDevice(s->device));
    if (!s->ev_t0) {
        CK(cudaEventCreate(&s->ev_t0));
        CK(cudaEventCreate(&s->ev_t1));
    }
    CK(cudaEventRecord(s->ev_t0, s->stream));
    return 0;
}

int crm_sim_timer_stop(crm_sim* s, double* ms_out) {
    CK(cudaSetDevice(s->device));
    if (!s->ev_t0) return fail(CRM_ERR_INVALID, "timer not started");
    CK(cudaEventRecord(s->ev_t1, s->stream));
    CK(cudaEventSynchronize(s->ev_t1));
    float ms = 0.0f;
    CK(cudaEventElapsedTime(&ms, s->ev_t0, s->ev_t1));
    if (ms_out) *ms_out = ms;
    return 0;
}

int crm_sim_run(crm_sim* s) {
    CK(cudaSetDevice(s->device));
    size_t next = 0;
    int rc;
    while (next < s->save_iter.size() && s->save_iter[next] < s->iteration) next++;
    if (next < s->save_iter.size() && s->save_iter[next] == s->iteration) {
        if ((rc = crm_take_snapshot(s))) return rc;
        next++;
    }
    while (s->iteration < s->max_iter) {
        // to the next save point (or the end) in as few host round trips as the colony allows
        const uint64_t target = next < s->save_iter.size() ? std::min(s->save_iter[next], s->max_iter) : s->max_iter;
        if ((rc = advance(s, std::max<uint64_t>(1, target - s->iteration)))) return rc;
        while (next < s->save_iter.size() && s->save_iter[next] < s->iteration) next++;
        if (next < s->save_iter.size() && s->save_iter[next] == s->iteration) {
            if ((rc = crm_take_snapshot(s))) return rc;
            next++;
        }
    }
    return 0;
}

int crm_sim_compute_forces(crm_sim* s, float* out_force, uint32_t* out_nn) {
    CK(cudaSetDevice(s->device));
    if (s->n == 0) return 0;
    int rc = launch_neighbour_and_force(s, true);
    if (rc) return rc;
    s->need_rebuild = true;  // the cell order the persistent structure is indexed by has been rewritten
    s->have_lists = false;
    CK(cudaGetLastError());
    if ((rc = crm_read_status(s))) return rc;
    s->last_pairs_cut = s->h_status->pairs_in_cutoff;
    s->last_pairs_tested = s->h_status->pairs_tested;
    if (out_force) {
        launch_pack(s->nv, s->d.force, s->stage, s->n, s->stream);
        CK(cudaMemcpyAsync(out_force, s->stage, sizeof(float) * (size_t)s->n * s->nv * 3, cudaMemcpyDeviceToHost, s->stream));
    }
    if (out_nn) CK(cudaMemcpyAsync(out_nn, s->d.nn, sizeof(uint32_t) * s->n, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    return 0;
}

// Interaction forces the last step used ([n][n_vertices][3]); tests compare them with a fresh search.
int crm_sim_last_forces(crm_sim* s, float* out_force) {
    if (!s || !out_force) return fail(CRM_ERR_INVALID, "null argument");
    if (s->fused && s->n_fused_steps)
        return fail(CRM_ERR_INVALID, "fused list steps do not store their forces: crm_sim_set_fused(sim, 0) before stepping");
    CK(cudaSetDevice(s->device));
    if (s->n == 0) return 0;
    launch_pack(s->nv, s->d.force, s->stage, s->n, s->stream);
    CK(cudaMemcpyAsync(out_force, s->stage, sizeof(float) * (size_t)s->n * s->nv * 3, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    return 0;
}

uint32_t crm_sim_n_rods(const crm_sim* s) { return s->n; }
uint32_t crm_sim_n_vertices(const crm_sim* s) { return (uint32_t)s->nv; }
uint64_t crm_sim_iteration(const crm_sim* s) { return s->iteration; }
uint64_t crm_sim_launch_count(const crm_sim* s) { return s->launches; }
uint64_t crm_sim_rod_steps(const crm_sim* s) { return s->rod_steps; }

void* crm_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        fail(CRM_ERR_CUDA, "cudaHostAlloc of %zu bytes failed", bytes);
        return nullptr;
    }
    return p;
}
void crm_host_free(void* ptr) {
    if (ptr) cudaFreeHost(ptr);
}

int crm_sim_get_state(crm_sim* s, float* pos, float* vel, crm_rod_params* par, crm_ident* ids, crm_ident* parents) {
    CK(cudaSetDevice(s->device));
    if (s->n == 0) return 0;
    return download_state(s, pos, vel, par, ids, parents);
}

uint32_t crm_sim_n_snapshots(const crm_sim* s) { return (uint32_t)s->snaps.size(); }

int crm_sim_snapshot_info(const crm_sim* s, uint32_t k, uint64_t* iteration, uint32_t* n_rods) {
    if (k >= s->snaps.size()) return fail(CRM_ERR_INVALID, "snapshot %u out of range", k);
    if (iteration) *iteration = s->snaps[k].iteration;
    if (n_rods) *n_rods = s->snaps[k].n;
    return 0;
}

int crm_sim_snapshot_copy(const crm_sim* s, uint32_t k, float* pos, float* vel, crm_rod_params* par, crm_ident* ids,
                          crm_ident* parents) {
    if (k >= s->snaps.size()) return fail(CRM_ERR_INVALID, "snapshot %u out of range", k);
    const HostSnap& sn = s->snaps[k];
    if (pos) memcpy(pos, sn.pos.data(), sizeof(float) * sn.pos.size());
    if (vel) memcpy(vel, sn.vel.data(), sizeof(float) * sn.vel.size());
    if (par) memcpy(par, sn.par.data(), sizeof(crm_rod_params) * sn.n);
    if (ids) memcpy(ids, sn.ids.data(), sizeof(crm_ident) * sn.n);
    if (parents) memcpy(parents, sn.parents.data(), sizeof(crm_ident) * sn.n);
    return 0;
}

int crm_sim_enable_timers(crm_sim* s, int enable) {
    s->timers = enable != 0;
    s->force_ms = 0.0;
    s->force_launches = 0;
    s->integrate_ms = 0.0;
    s->integrate_launches = 0;
    return 0;
}

// Integrator launches timed next to the force launches (first step of every batch), and how the steps were
// driven: batches of list steps (one status read-back each) vs single steps.
int crm_sim_read_step_stats(crm_sim* s, double* integrate_ms, uint64_t* integrate_launches, uint64_t* n_batches,
                            uint32_t* batch_len) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    if (integrate_ms) *integrate_ms = s->integrate_ms;
    if (integrate_launches) *integrate_launches = s->integrate_launches;
    if (n_batches) *n_batches = s->n_batches;
    if (batch_len) *batch_len = s->batch_len;
    return 0;
}

// 1 (default): a list step is ONE kernel (forces + mechanics, the force never goes to memory); 0: k_force_eval and
// k_integrate as two launches (crm_sim_last_forces then returns the forces of the last step).
int crm_sim_set_fused(crm_sim* s, int on) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    s->fused = on != 0;
    return 0;
}
uint64_t crm_sim_fused_steps(const crm_sim* s) { return s ? s->n_fused_steps : 0; }

// 1 (default): list steps are enqueued in batches with a device-side latch; 0: one host round trip per step.
int crm_sim_set_batching(crm_sim* s, int on) {
    if (!s) return fail(CRM_ERR_INVALID, "null sim");
    s->batching = on != 0;
    return 0;
}

int crm_sim_read_timers(crm_sim* s, double* force_ms, uint64_t* force_launches, uint64_t* pairs_cut,
                        uint64_t* pairs_tested) {
    if (force_ms) *force_ms = s->force_ms;
    if (force_launches) *force_launches = s->force_launches;
    if (pairs_cut) *pairs_cut = s->last_pairs_cut;
    if (pairs_tested) *pairs_tested = s->last_pairs_tested;
    return 0;
}

}  // extern "C"
