// crm_sim.h — private definition of the simulation object shared by the host translation units
// (crm_sim.cu: stepping + C ABI; crm_shard.cu: voxel-slab sharding).  Not part of the C ABI.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <string>
#include <thread>
#include <memory>
#include <utility>
#include <vector>

#include "crm_device.cuh"

// std::vector whose resize() leaves trivially constructible elements uninitialised: the 1e6-rod host
// mirrors are written right after they are sized, a zero fill first would only cost bus time.
// Large blocks come from (and go back to) a process-wide cache (crm_host.cpp): a simulation of 1e6 rods sizes ~260 MB of
// host mirrors, and fresh pages cost a fault each when they are first written (~40 core-ms per upload, all of it in
// the timed region of every repeated run_simulation call and multiplied by the ranks sharing a host).
void* crm_host_block_alloc(size_t bytes);
void crm_host_block_free(void* p, size_t bytes) noexcept;
template <class T>
struct crm_default_init_alloc : std::allocator<T> {
    template <class U> struct rebind { using other = crm_default_init_alloc<U>; };
    using std::allocator<T>::allocator;
    T* allocate(size_t n) { return static_cast<T*>(crm_host_block_alloc(n * sizeof(T))); }
    void deallocate(T* p, size_t n) noexcept { crm_host_block_free(p, n * sizeof(T)); }
    template <class U> void construct(U* p) noexcept(std::is_nothrow_default_constructible<U>::value) { ::new (static_cast<void*>(p)) U; }
    template <class U, class... A> void construct(U* p, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A>(a)...); }
};
template <class T> using crm_hvec = std::vector<T, crm_default_init_alloc<T>>;

namespace crm {
// launchers (crm_mech.cu / crm_force.cu)
bool nv_supported(int nv);
int lanes_per_rod(int nv);
void launch_unpack(int nv, const float* src, float* dst, uint32_t n, cudaStream_t st);
void launch_pack(int nv, const float* src, float* dst, uint32_t n, cudaStream_t st);
void launch_bounds_kernel(int nv, const Dev& d, cudaStream_t st);
int launch_binning(const Dev& d, uint32_t n_tiles, cudaStream_t st);
void launch_force(int nv, const Dev& d, bool fast, bool count, bool build, cudaStream_t st);
bool launch_force_eval(int nv, const Dev& d, bool fast, bool uniform, int pot_type, bool fused, cudaStream_t st);
size_t list_kernel_smem(int lp, int maxl, bool fused);
void launch_force_lists(int nv, const Dev& d, bool fast, cudaStream_t st);
int launch_patch(const Dev& d, uint32_t n_div, uint32_t n0, int lp, int maxp, cudaStream_t st);
void launch_integrate(int nv, const Dev& d, cudaStream_t st);
void launch_divide(int nv, const Dev& d, uint32_t n_div, uint32_t n0, cudaStream_t st);
// host helpers (crm_host.cpp)
int build_save_schedule(const crm_config* c, std::vector<uint64_t>& out, uint64_t& max_iter, std::string& err);
}  // namespace crm

// records `fmt...` as the calling thread's last error and returns `code`
int crm_fail(int code, const char* fmt, ...);

#define CK(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return crm_fail(CRM_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                            __FILE__, __LINE__);                                                  \
    } while (0)

struct HostSnap {
    uint64_t iteration;
    uint32_t n;
    std::vector<float> pos, vel;
    std::vector<crm_rod_params> par;
    std::vector<crm_ident> ids, parents;
    std::vector<uint64_t> gslot;
    std::vector<uint32_t> sim_of;
};

// NCCL transport of the sharded step loop (crm_comm.cu)
struct crm_comm;
int crm_comm_rank(const crm_comm* c);
int crm_comm_world(const crm_comm* c);
int crm_comm_sendrecv(crm_comm* c, const void* send_lo, size_t n_send_lo, void* recv_lo, size_t n_recv_lo,
                      const void* send_hi, size_t n_send_hi, void* recv_hi, size_t n_recv_hi, cudaStream_t st);
int crm_comm_allgather(crm_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st);
// both halo messages and the all-gather of the flags in ONE NCCL group (one launch)
int crm_comm_exchange(crm_comm* c, const void* send_lo, size_t n_send_lo, void* recv_lo, size_t n_recv_lo, const void* send_hi,
                      size_t n_send_hi, void* recv_hi, size_t n_recv_hi, const void* flags, void* flags_all,
                      size_t flag_bytes, cudaStream_t st);
// fused pack / unpack kernels of a sharded step between rebuilds (crm_shard.cu)
int crm_shard_pack_all(crm_sim* s, void* msg_lo, void* msg_hi, uint32_t* flags);
int crm_shard_unpack_all(crm_sim* s, const void* msg_lo, const void* msg_hi, const uint32_t* flags_all, int world,
                         uint32_t* latch, uint32_t iteration_prev);
void crm_comm_free(crm_comm* c);

struct crm_sim {
    crm_config cfg{};
    int device = 0;
    cudaStream_t stream = nullptr;
    int nv = 0, lp = 0;
    uint32_t n = 0, cap = 0;  // owned rods, slot capacity (owned + ghosts)
    uint32_t n_ghost = 0;     // ghost rods in slots [n, n + n_ghost)
    uint32_t ncells_cap = 0;
    crm::Dev d{};
    float4* vrec_alt = nullptr;  // second vertex-record buffer of the fused list step (swapped with d.vrec per step)
    bool step_fused = false;     // the list step in flight runs the fused kernel
    bool fused = true;           // list steps run force + mechanics in one kernel (CRM_B200_FUSED=0 / crm_sim_set_fused)
    uint64_t n_fused_steps = 0;
    char* arena = nullptr;   // the one device allocation all per-rod arrays are parts of (crm_reserve)
    float* stage = nullptr;  // boundary-layout staging buffer [cap][nv][3]
    float* stage2 = nullptr; // second one: positions and velocities cross the bus back to back
    crm::Status* h_status = nullptr;  // pinned
    float emax = 0.0f;        // max centroid->vertex extent over owned rods
    float emax_ghost = 0.0f;  // ... over the current ghost rods
    float rsearch = 0.0f;
    bool fast_math = true;
    bool bounds_fresh = false;  // bsphere / chord / cull thresholds / emax describe the current positions
    // persistent neighbour structure (Verlet lists): valid until a vertex moved more than skin/2
    float skin = 0.5f;          // um; 0 disables the structure (every step searches from scratch)
    bool need_rebuild = true;   // the next step must search (and rebuild the structure)
    bool have_lists = false;    // the structure on the device describes the current slots
    bool lists_morse = true, lists_mie = true;  // potential types among the rods of the structure (same protocol)
    uint32_t overflow_reasons = 0;  // why builds were abandoned (bits of Status::plist_overflow), how often
    uint64_t n_overflows = 0;
    // why steps asked for a new search (crm_sim_rebuild_reasons): 0 displacement beyond skin/2, 1 rod left the slab,
    // 2 a rod did not fit the structure, 3 a patched row is full, 4 divisions that could not be patched,
    // 5 no live structure (hold-off / first step / reallocation), 6 search repeated on a finer grid, 7 steps in total
    uint64_t why_rebuild[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t lists_maxl = 32;   // longest list of the structure (sizes the staging area of the list kernel)
    bool lists_slow = true;     // ... and queues pairs with mixed potentials (known after the build step's status)
    bool want_rebuild = true;   // outcome of the last step: a local reason to rebuild (sharded runs: the
                                // coordinator combines the slabs' wishes, crm_sim_keep_lists)
    uint32_t* halo_sel[2] = {nullptr, nullptr};  // frozen halo selections (slots) per side
    uint32_t halo_sel_n[2] = {0, 0};
    uint32_t halo_sel_cap = 0;
    uint32_t holdoff = 0;       // steps for which rebuilding is not attempted (overflow / frequent rebuilds)
    uint32_t steps_since_build = 0;
    uint32_t short_builds = 0;  // consecutive rebuilds whose lists served fewer than 2 steps
    uint32_t evals_since_build = 0;
    uint32_t n_dirty = 0;       // rows queued for k_force<LISTS> before the next list step (divisions)
    // rows in two classes: lists of up to `short_maxl` rods are staged by the main launch of a list step (sized for
    // 7+ resident blocks per SM), the n_long longer ones (long_rows, from k_classify_rows) by a second, small launch
    uint32_t* long_rows = nullptr;
    uint32_t n_long = 0;
    bool classified = false;    // this step's status block carries a fresh Status::n_long
    uint64_t n_two_class_steps = 0;
    uint32_t n_patches = 0;     // division steps absorbed without a new search
    bool patching = true;       // CRM_B200_PATCH=0 / crm_sim_set_patching: search again after every division
    bool patch_divisions = false;  // set by the step that saw divisions and no other reason to search again
    uint32_t backoff = 8;       // next holdoff (doubles while the lists keep being useless)
    uint64_t n_builds = 0, n_eval_steps = 0;
    // host-side bookkeeping
    crm_hvec<crm_rod_params> base;  // cold fields per slot
    // (default-initialised storage: every element is written before it is read, and a 1e6-rod upload must not
    // zero-fill and page-fault 56 MB on one thread first)
    crm_hvec<crm_ident> ids, parents;
    crm_hvec<uint64_t> gslot;          // global slot (= slot of the unsharded run) per owned slot
    std::vector<uint64_t> vox_counter;   // [sim][user voxel]
    std::vector<uint32_t> sim_of;        // batched sims: sim index per slot (empty: one colony)
    uint32_t n_sims = 1;
    uint32_t* tile_buf = nullptr;        // device copy of sim_of (Dev::tile points here when batched)
    uint32_t* rkey_buf = nullptr;        // device copy of gslot (Dev::rkey points here when gslot != slot)
    bool rkey_on = false;                // sharded / batched runs: random streams are keyed by the global slot
    std::vector<uint64_t> sim_rods;      // batched sims: rods ever created per sim (next local slot)
    std::vector<uint32_t> sim_rods0;     // batched sims: rods at set_agents (a sim that divided has more now)
    int* sim_fail = nullptr;             // device [n_sims], Dev::sim_fail points here while isolation is on
    std::vector<crm::DivRecord> h_div;
    crm::DivRecord* h_div_pinned = nullptr;  // page-locked landing zone of the device-ordered division list
    uint32_t h_div_pinned_cap = 0;
    uint64_t iteration = 0, rod_steps = 0, launches = 0;
    std::vector<uint64_t> save_iter;
    uint64_t max_iter = 0;
    std::vector<HostSnap> snaps;
    // sharding (crm_shard.cu)
    bool sharded = false;      // divisions are deferred to crm_sim_apply_divisions
    bool own_stream = true;    // false after crm_sim_set_stream
    bool step_open = false;    // crm_sim_step_begin enqueued a step that crm_sim_step_end has not finished
    uint32_t ghost_n[2] = {0, 0};  // ghosts installed from the lower / upper message (crm_sim_ghosts_refresh)
    float slab_lo = -INFINITY, slab_hi = INFINITY;
    uint32_t pending_div = 0;  // divisions of the last step awaiting crm_sim_apply_divisions
    crm_comm* comm = nullptr;        // communicator of all slabs (crm_sim_comm_init)
    uint32_t* flags_dev = nullptr;   // [4] this slab's flags of a step, [4 * world] the gathered ones behind them
    uint32_t* flags_host = nullptr;  // pinned [4 * world]
    void* select_tmp = nullptr;  // cub temp storage
    size_t select_tmp_bytes = 0;
    uint32_t* select_out = nullptr;  // [select_cap] selected slots
    uint32_t select_cap = 0;
    uint32_t* select_count = nullptr;
    // timers
    bool timers = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev2 = nullptr, ev3 = nullptr;  // around the integrator launch of the first step of a batch
    double integrate_ms = 0.0;
    uint64_t integrate_launches = 0;
    // batches of list steps (crm_sim.cu step_batch)
    bool async_bounds = true;   // search steps do not wait for the bounds kernel (CRM_B200_ASYNC_BOUNDS=0: they do)
    bool async_search = false;  // the step in flight is such a search step
    uint64_t n_grid_redos = 0;  // search steps repeated because a rod outgrew the grid's head-room
    uint32_t* latch_buf = nullptr;  // device word (own cache line) the latching step writes its iteration to
    bool batching = true;      // CRM_B200_BATCH=0 / crm_sim_set_batching: one host round trip per step
    uint32_t batch_len = 4;    // steps of the next batch: doubles after a full batch (<= 64), shrinks to what a latched one served
    uint64_t n_batches = 0;
    cudaEvent_t ev_status = nullptr;  // recorded behind the status read-back of a step: what step_finish waits for
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;  // crm_sim_timer_start / stop
    double force_ms = 0.0;
    uint64_t force_launches = 0;
    uint64_t last_pairs_cut = 0, last_pairs_tested = 0;
};

// shared host helpers (crm_sim.cu)
int crm_take_snapshot(crm_sim* s);
int crm_refresh_bounds(crm_sim* s);
int crm_apply_divisions(crm_sim* s, const crm_ident* id_mother, const crm_ident* id_daughter, const uint64_t* gslot_daughter);
void crm_fill_device_params(const crm_rod_params& p, crm::PotParams& pot, crm::MechParams& mech, crm::GrowState& grow,
                            crm::CycleParams& cyc);
int crm_reserve(crm_sim* s, uint32_t need);
int crm_read_status(crm_sim* s);
int crm_upload_rkeys(crm_sim* s, uint32_t first, uint32_t count);
int crm_report_device_error(crm_sim* s);
