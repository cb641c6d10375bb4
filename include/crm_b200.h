/*
 * crm_b200.h — C ABI of the B200-native colony-mechanics path.
 *
 * This is the drop-in boundary beneath the Python surface of cr_mech_coli
 * (crm.Configuration / crm.AgentSettings / crm.generate_agents /
 * crm.run_simulation_with_agents -> CellContainer).  The reference has no
 * C-ABI of its own: its boundary is the PyO3 module `cr_mech_coli.cr_mech_coli`
 * (reference src/lib.rs:46-116).  Every entry point below names the reference
 * interface it replaces (file:line relative to the reference tree).
 *
 * Conventions: plain pointers and sizes only; the caller owns host buffers,
 * the library owns device buffers; every function returns 0 on success or a
 * negative crm_status; crm_last_error() returns the message of the last
 * failure on the calling thread.  No exceptions cross the ABI.
 *
 * Array layouts at the boundary are those of the Python API: a rod's
 * position is an (n_vertices, 3) row-major f32 matrix (numpy layout of
 * RodAgent.pos, reference src/agent.rs:520-525); N rods are N such matrices
 * back to back.
 */
#ifndef CRM_B200_H
#define CRM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRM_ABI_VERSION 3

typedef enum crm_status {
    CRM_OK = 0,
    CRM_ERR_INVALID = -1,   /* bad argument / configuration                    */
    CRM_ERR_CUDA = -2,      /* CUDA runtime failure                            */
    CRM_ERR_BOUNDARY = -3,  /* "Particle is out of domain" (BoundaryError)     */
    CRM_ERR_NUMERIC = -4,   /* NaN/Inf in the state (CalcError)                */
    CRM_ERR_DIVISION = -5,  /* division failed (DivisionError)                 */
    CRM_ERR_TIME = -6,      /* invalid time configuration (TimeError)          */
    CRM_ERR_NOMEM = -7
} crm_status;

/* Interaction potential variants: reference src/agent.rs:272-277 (enum PhysInt). */
enum { CRM_POT_MORSE = 0, CRM_POT_MIE = 1 };
/* Setter variants: reference src/agent.rs:11-35 / :104-112
 * (GrowthRateSetter / SpringLengthThresholdSetter: NormalDistr{mean,std} | Explicit{a,b}). */
enum { CRM_SETTER_NORMAL = 0, CRM_SETTER_EXPLICIT = 1 };
/* CellIdentifier variants (cellular_raza CellIdentifier; reference
 * cr_mech_coli/cr_mech_coli.pyi:81-95, src/crm_fit/predict.rs:267,
 * src/crm_divide.rs:219).  The numeric order is the sort order. */
enum { CRM_ID_DIVISION = 0, CRM_ID_INITIAL = 1, CRM_ID_INSERTED = 2 };

/* Replaces `Configuration` (reference src/config.rs:395-465). */
typedef struct crm_config {
    float t0;
    float dt;
    float t_max;
    uint32_t n_saves;
    float domain_size[2];
    float domain_height;
    uint32_t n_voxels[2];
    uint64_t rng_seed;
    float gel_pressure;
    float surface_friction;
    float surface_friction_distance;
    uint32_t n_threads; /* accepted, ignored: the device replaces the rayon pool */
} crm_config;

/* Per-rod scalar state: the non-matrix fields of `RodAgent`
 * (reference src/agent.rs:212-239; RodMechanics fields src/config.rs:63-84;
 * potentials src/agent.rs:323-339). */
typedef struct crm_rod_params {
    float diffusion_constant;
    float spring_tension;
    float rigidity;
    float spring_length;
    float damping;
    int32_t potential;          /* CRM_POT_* */
    float radius;
    float strength;
    float cutoff;
    float potential_stiffness;  /* Morse only */
    float bound;                /* Mie only   */
    float en;                   /* Mie only   */
    float em;                   /* Mie only   */
    uint32_t n_neighbors;       /* PhysicalInteraction.1 */
    float growth_rate;
    int32_t growth_setter_kind; /* CRM_SETTER_* */
    float growth_setter_a;      /* mean | g1 */
    float growth_setter_b;      /* std  | g2 */
    float spring_length_threshold;
    int32_t threshold_setter_kind;
    float threshold_setter_a;   /* mean | l1 */
    float threshold_setter_b;   /* std  | l2 */
    int32_t has_neighbor_reduction;
    uint32_t neighbor_reduction_max;
    float neighbor_reduction_exp;
} crm_rod_params;

/* CellIdentifier: Initial(n) -> {CRM_ID_INITIAL, 0, n};
 * Division(voxel, counter) -> {CRM_ID_DIVISION, voxel, counter}.
 * kind < 0 encodes `None` (no parent). */
typedef struct crm_ident {
    int64_t kind;
    uint64_t voxel;
    uint64_t counter;
} crm_ident;

typedef struct crm_sim crm_sim;

/* ---- library ---------------------------------------------------------- */
int crm_abi_version(void);
const char* crm_last_error(void);
/* Number of visible CUDA devices (<0: CUDA unavailable). */
int crm_device_count(void);
/* Free / total device memory in bytes (leak checks of callers that open a sim per cost evaluation,
 * reference src/crm_fit/predict.rs:231-287). */
int crm_device_mem_info(int device, size_t* free_bytes, size_t* total_bytes);

/* ---- host helpers on the boundary ------------------------------------ */
/* Replaces `counter_to_color` / `color_to_counter` (reference src/imaging.rs:28-66). */
void crm_counter_to_color(uint32_t counter, uint8_t rgb[3]);
uint32_t crm_color_to_counter(const uint8_t rgb[3]);

/* Replaces `_generate_positions` (reference src/simulation.rs:107-166):
 * ChaCha8 + rand-0.9 stream-exact placement.  out_pos: [n_agents][n_vertices][3]. */
int crm_generate_positions(uint32_t n_agents, float spring_length, const crm_config* config,
                           uint64_t rng_seed, const float dx[2], float randomize_positions,
                           uint32_t n_vertices, float* out_pos);

/* Save schedule of `run_simulation_with_agents` (reference src/simulation.rs:239-249;
 * cellular_raza FixedStepsize::from_partial_save_points).  Writes at most `cap`
 * iteration indices, returns the number of distinct save iterations (or <0);
 * *n_iterations receives the index of the last iteration. */
int crm_save_schedule(const crm_config* config, uint64_t* out_iterations, uint32_t cap,
                      uint64_t* n_iterations);

/* ---- simulation object ------------------------------------------------ */
/* Replaces the setup half of `run_simulation_with_agents`
 * (reference src/simulation.rs:234-292): domain, stepper, storage. */
int crm_sim_create(const crm_config* config, int device, crm_sim** out);
void crm_sim_destroy(crm_sim* sim);

/* Agents are copied in (reference moves `Vec<RodAgent>` by value,
 * src/simulation.rs:234-237).  pos/vel: [n_rods][n_vertices][3] host f32. */
int crm_sim_set_agents(crm_sim* sim, const float* pos, const float* vel,
                       const crm_rod_params* params, uint32_t n_rods, uint32_t n_vertices);

/* Batched independent simulations in one device state (parameter-fitting ensembles: the reference
 * runs one run_simulation_with_agents per parameter point, src/crm_fit/predict.rs:231-287, fanned out
 * by rayon / scipy workers, src/crm_fit/optimize.rs:75-95,235-252).  sim_of[a] = simulation of rod a;
 * rods of different simulations never interact; identifiers, per-voxel counters and snapshots are
 * kept per simulation (crm_sim_snapshot_sim_index splits a snapshot).  All simulations share the
 * Configuration. */
int crm_sim_set_agents_batched(crm_sim* sim, const float* pos, const float* vel, const crm_rod_params* params,
                               const uint32_t* sim_of, uint32_t n_rods, uint32_t n_vertices, uint32_t n_sims);
int crm_sim_get_sim_index(const crm_sim* sim, uint32_t* out);
/* Error isolation: a BoundaryError / CalcError stops only the simulation it happens in (its rods keep
 * their last valid state), like the reference's one run per parameter point where an error costs that
 * member alone (src/crm_fit/predict.rs:224-228,249-256).  Status per simulation: 0, CRM_ERR_BOUNDARY,
 * CRM_ERR_NUMERIC, or CRM_ERR_DIVISION if the simulation has divided (reported, not an error). */
int crm_sim_isolate_sim_errors(crm_sim* sim, int on);
int crm_sim_get_sim_status(crm_sim* sim, int32_t* out);
/* `predict_calculate_cost_rs` for a whole population (reference src/crm_fit/predict.rs:231-287): steps
 * to the last result iteration without snapshots and accumulates on the device, at result_iterations[r]
 * (increasing), every simulation's sum of squared xy distances to targets[r][rod][vertex][0..1]
 * (host f32 [n_results][n_agents][n_vertices][2]; rod = index inside its simulation).  Errors are
 * isolated; out_status as crm_sim_get_sim_status (may be NULL). */
int crm_sim_run_costs(crm_sim* sim, const float* targets, uint32_t n_results, uint32_t n_agents,
                      const uint64_t* result_iterations, float* out_costs, int32_t* out_status);
int crm_sim_snapshot_sim_index(const crm_sim* sim, uint32_t k, uint32_t* out);

/* The whole of `run_main!` (reference src/simulation.rs:283-292): steps to
 * t_max, snapshots at the save iterations into host memory owned by `sim`. */
int crm_sim_run(crm_sim* sim);

/* Advance `n_steps` mechanics steps without taking snapshots (bench / tests). */
int crm_sim_step(crm_sim* sim, uint32_t n_steps);

/* Same as crm_sim_step, timed on the device: CUDA events recorded on the sim's own stream
 * before the first and after the last kernel of the n_steps (ms). */
int crm_sim_step_timed(crm_sim* sim, uint32_t n_steps, double* ms);

/* Device stopwatch on the sim's own stream around step sequences driven from above the ABI
 * (sharded runs: halo exchange + crm_sim_step per step): CUDA events, elapsed ms. */
int crm_sim_timer_start(crm_sim* sim);
int crm_sim_timer_stop(crm_sim* sim, double* ms);

/* Force accumulators the next step would use, i.e. the reference's
 * interaction + domain force for the current state (cellular_raza
 * update_mechanics_interaction_step_1..3 + calculate_custom_domain_force).
 * out_force: [n_rods][n_vertices][3]; out_n_neighbors: [n_rods] (either may be NULL). */
int crm_sim_compute_forces(crm_sim* sim, float* out_force, uint32_t* out_n_neighbors);

/* 1: IEEE sqrt/div + libdevice expf/powf in the force kernel; 0 (default): MUFU
 * approximations (rsqrt/ex2/lg2/rcp).  Env CRM_B200_EXACT_MATH=1 sets the default. */
int crm_sim_set_exact_math(crm_sim* sim, int exact);

/* Two decisions the un-vendored cellular_raza source would settle, as run-time
 * parameters; the oracle has the same switches.  Both must precede crm_sim_set_agents.
 * stored_increments: history depth of the Adams-Bashforth mechanics solver, 0 Euler, 1 AB2,
 * 2 AB3 (default; cellular_raza's `mechanics_solver_order` kwarg of run_main!, which reference
 * src/simulation.rs:283-292 does not pass).  Env CRM_B200_SOLVER_ORDER sets the default. */
int crm_sim_set_solver_order(crm_sim* sim, int stored_increments);
/* Weight of a segment's spring force on each of its two ends: 1 (default, reference
 * docs/source/model.rst:32-36) or 1/2.  Env CRM_B200_SPRING_END_WEIGHT sets the default. */
int crm_sim_set_spring_end_weight(crm_sim* sim, float weight);

/* How the steps are driven.  Between two searches a step needs the host only if it divides a rod, fails or moves
 * a vertex beyond skin/2; crm_sim_step / _run / _run_costs therefore enqueue list steps in batches (the first
 * such step latches on the device, later steps of the batch return at once) and read ONE status block per
 * batch.  Results equal one-by-one stepping.  crm_sim_set_batching(0) / env CRM_B200_BATCH=0 restores one host
 * round trip per step.  crm_sim_read_step_stats: integrator time next to crm_sim_read_timers' force time (both
 * sampled on the first step of every batch), batches run, current batch length. */
int crm_sim_set_batching(crm_sim* sim, int on);
/* A list step (between two searches) is ONE kernel by default: the interaction forces are summed in registers and
 * the mechanics of the rod (RodMechanics::calculate_increment, Adams-Bashforth, update_cycle, boundary) follow at
 * once, so the force array is neither written nor read.  0 restores the two launches (needed for
 * crm_sim_last_forces).  Env CRM_B200_FUSED=0 sets the default.  Mixed-potential colonies and exact-math runs
 * always use two launches. */
int crm_sim_set_fused(crm_sim* sim, int on);
uint64_t crm_sim_fused_steps(const crm_sim* sim);
int crm_sim_read_step_stats(crm_sim* sim, double* integrate_ms, uint64_t* integrate_launches, uint64_t* n_batches,
                            uint32_t* batch_len);

/* Skin (um) of the persistent neighbour structure: rods and point-segment evaluations within
 * range + skin are listed once and reused until a vertex has moved more than skin/2 (Verlet
 * lists; results are identical because skipped evaluations are exact zeros).  0 = search from
 * scratch every step.  Default 0.5 or env CRM_B200_SKIN.  Must precede crm_sim_set_agents. */
int crm_sim_set_skin(crm_sim* sim, float skin);
/* Steps that rebuilt the structure / steps that reused it. */
int crm_sim_list_stats(const crm_sim* sim, uint64_t* n_builds, uint64_t* n_eval_steps);
/* Builds abandoned because a rod did not fit the structure's capacities: reason bits (1: more than 32 neighbour
 * rods (16 for 9..16 vertices), 2 / 4: more than 128 queued own / ext evaluations per vertex, 8: more than 8
 * sensing candidates per vertex), their number, and the longest list of the live structure. */
int crm_sim_list_overflows(const crm_sim* sim, uint32_t* reasons, uint64_t* count, uint32_t* longest_list);
/* Why steps asked for a new neighbour search (diagnostics of the persistent structure; the reference re-sorts its
 * voxels every step, cellular_raza chili sort_cells_in_subdomains): counts of steps since the upload with
 * [0] a vertex displaced beyond skin/2, [1] a rod that left its slab, [2] a rod that did not fit the structure,
 * [3] a patched row that is full, [4] divisions that could not be patched, [5] no live structure (first step,
 * hold-off, reallocation), [6] a search repeated on a finer grid, [7] steps in total. */
int crm_sim_rebuild_reasons(const crm_sim* sim, uint64_t out[8]);
/* Totals over the live structure (diagnostics; no counterpart in the reference, whose voxel stencil holds no such
 * state): [0] queued own evaluations, [1] queued ext evaluations, [2] sensing candidates, [3] listed rods, [4] rows
 * whose list is too long for the main launch of a list step (served by a second, small launch), [5] list steps so far
 * that ran as those two launches.  What a list step evaluates is out[0] + out[1]; [0..4] are zero when no structure
 * is live. */
int crm_sim_list_totals(crm_sim* sim, uint64_t out[6]);
/* Divisions while the structure is live do not force a new search: the rows of the daughters and of the
 * mothers' neighbours are rebuilt from candidate lists (the reference re-sorts its voxels every step;
 * cellular_raza chili update_cell_cycle + sort_cells_in_subdomains).  Count of division steps absorbed
 * that way, and the switch (default on; CRM_B200_PATCH=0). */
uint64_t crm_sim_patch_count(const crm_sim* sim);
/* Interaction forces the last step used ([n_rods][n_vertices][3]; rods created by that step's divisions
 * read zero): lets a test compare a list step with a fresh search of the same positions. */
int crm_sim_last_forces(crm_sim* sim, float* out_force);
int crm_sim_set_patching(crm_sim* sim, int on);

uint32_t crm_sim_n_rods(const crm_sim* sim);
uint32_t crm_sim_n_vertices(const crm_sim* sim);
uint64_t crm_sim_iteration(const crm_sim* sim);

/* Page-locked host memory for the buffers handed to crm_sim_set_agents / crm_sim_get_state: from such
 * buffers positions and velocities cross the bus at link speed and overlap the host-side parameter
 * conversion (pageable buffers work too, through the driver's bounce buffers).  The reference copies its
 * inputs out of numpy arrays (src/agent.rs:485-489) and hands back fresh copies (:522-525); this pair is
 * what a binding uses to allocate those arrays.  crm_host_alloc returns NULL on failure. */
void* crm_host_alloc(size_t bytes);
void crm_host_free(void* ptr);

/* Current state in slot order (any pointer may be NULL). */
int crm_sim_get_state(crm_sim* sim, float* pos, float* vel, crm_rod_params* params,
                      crm_ident* ids, crm_ident* parents);

/* Snapshots taken by crm_sim_run: what `storage.cells.load_all_elements()`
 * yields in the reference (src/simulation.rs:293-306). */
uint32_t crm_sim_n_snapshots(const crm_sim* sim);
int crm_sim_snapshot_info(const crm_sim* sim, uint32_t k, uint64_t* iteration, uint32_t* n_rods);
int crm_sim_snapshot_copy(const crm_sim* sim, uint32_t k, float* pos, float* vel,
                          crm_rod_params* params, crm_ident* ids, crm_ident* parents);

/* Counters for the bench harness: kernels launched by this sim so far and
 * the rod-steps it has advanced. */
uint64_t crm_sim_launch_count(const crm_sim* sim);
uint64_t crm_sim_rod_steps(const crm_sim* sim);
/* Device time (ms, CUDA events on the sim's stream) spent in the force kernel
 * and in all kernels since the last call to crm_sim_reset_timers; the number of
 * interacting ordered rod pairs of the last force launch (for the roofline). */
int crm_sim_enable_timers(crm_sim* sim, int enable);
int crm_sim_read_timers(crm_sim* sim, double* force_ms, uint64_t* force_launches,
                        uint64_t* last_pairs_in_cutoff, uint64_t* last_pairs_tested);

/* ---- voxel-slab sharding over several GPUs ------------------------------
 * Replaces the sub-domain machinery of the reference backend: `Configuration.n_threads`
 * sub-domains of voxels that exchange boundary-voxel agents and returned forces every step
 * (reference src/simulation.rs:251-256,283-292; cellular_raza chili SubDomainBox).  One crm_sim
 * per GPU owns the rods whose centroid x lies in [x_lo, x_hi).  Forces are gathered by the owner,
 * so one exchange of halo records per step replaces the reference's position + force round trip.
 * The transport (NCCL send/recv, or a plain copy between two sims of one device in the tests)
 * lives above the ABI; `dev_*` arguments are DEVICE pointers owned by the caller. */
/* Declare the slab and switch the sim to sharded mode: divisions are deferred to
 * crm_sim_apply_divisions because identifiers depend on the divisions of all slabs. */
int crm_sim_set_slab(crm_sim* sim, float x_lo, float x_hi);
/* Global slot of every owned rod = its slot in the unsharded run (index in the input list for
 * initial rods): the partition-independent ordering key of divisions and snapshots. */
int crm_sim_set_global_slots(crm_sim* sim, const uint64_t* gslots, uint32_t n);
int crm_sim_get_global_slots(const crm_sim* sim, uint64_t* out);
uint32_t crm_sim_halo_record_bytes(const crm_sim* sim);
uint32_t crm_sim_migrant_record_bytes(const crm_sim* sim);
float crm_sim_max_extent(const crm_sim* sim);
/* The integrator maintains vertex records only; bounding data (centroids, extents) is recomputed
 * before every search.  Call this before reading crm_sim_max_extent for a new halo selection. */
int crm_sim_refresh_bounds(crm_sim* sim);    /* max centroid->vertex distance over owned rods */
float crm_sim_search_radius(const crm_sim* sim); /* max interaction range over owned rods */
/* Halo records of the owned rods whose centroid lies within `width` of the lower (side 0) /
 * upper (side 1) slab edge, in slot order, into dev_buf (capacity `cap` records). */
int crm_sim_halo_pack(crm_sim* sim, int side, float width, void* dev_buf, uint32_t cap, uint32_t* n);
/* Replace the ghost rods by the records received from the lower / upper neighbour;
 * max_extent = the largest extent among them (as reported by their owners). */
int crm_sim_set_ghosts(crm_sim* sim, const void* dev_lo, uint32_t n_lo, const void* dev_hi, uint32_t n_hi,
                       float max_extent);
/* The same per step with less host traffic: both messages [64-byte header {n_records, max_extent,
 * flags} | records] packed in one call (NULL = no neighbour on that side) ... */
int crm_sim_halo_pack_msgs(crm_sim* sim, float width, void* dev_msg_lo, void* dev_msg_hi, uint32_t cap,
                           uint32_t flags, int reselect, uint32_t* n_lo, uint32_t* n_hi);
/* reselect = 0 re-sends the rods chosen at the last reselect (frozen halo: ghost slots stay put, which
 * is what lets the receiver keep its persistent neighbour structure between rebuilds).
 * Rebuilds are collective: after a step every slab reports crm_sim_wants_rebuild; if none does, the
 * coordinator calls crm_sim_keep_lists on all of them, otherwise the next step reselects the halos
 * and searches again everywhere. */
int crm_sim_wants_rebuild(const crm_sim* sim);
int crm_sim_keep_lists(crm_sim* sim);
float crm_sim_skin(const crm_sim* sim);
/* ... halo records of the owned rods that are about to migrate (centroid outside the slab): the
 * old owner keeps them as ghosts during the hand-over step ... */
int crm_sim_pack_leavers(crm_sim* sim, void* dev_buf, uint32_t cap, uint32_t* n);
/* ... and the ghosts installed from the two received messages plus those leavers; returns the
 * neighbours' record counts [2], flags [2] and the largest extent they reported. */
int crm_sim_set_ghosts_msgs(crm_sim* sim, const void* dev_msg_lo, const void* dev_msg_hi, const void* dev_self,
                            uint32_t n_self, uint32_t* counts, uint32_t* flags, float* max_extent);
/* Stream-ordered variants for the steps between two collective rebuilds (the reference's sub-domain
 * threads exchange forces and positions through channels without a global barrier as well:
 * cellular_raza chili run_main update_mechanics_interaction_step_1..3).  None of them waits for the
 * device; the caller's transport must be ordered on crm_sim_stream (a cudaStream_t).
 *   crm_sim_halo_repack     the frozen selections packed again into the two messages
 *   crm_sim_ghosts_refresh  received records scattered to the ghost slots of the last rebuild step
 *   crm_sim_step_begin      the step enqueued; dev_flags[0..3] = {divisions, reason to rebuild, -error code
 *                           (0: none), 0} written on the stream (device memory, 16 bytes, may be NULL)
 *   crm_sim_step_end        one wait + the host part of the step.  It waits for the step's own work (an event
 *                           behind its status read-back), not for what the caller enqueued on the stream after
 *                           crm_sim_step_begin: the next halo exchange keeps running while the host decides
 * crm_sim_step(sim, 1) == crm_sim_step_begin(sim, NULL) + crm_sim_step_end(sim). */
void* crm_sim_stream(const crm_sim* sim);
/* ... or make the sim run on a stream the caller owns (it must outlive the sim). */
int crm_sim_set_stream(crm_sim* sim, void* stream);
int crm_sim_halo_repack(crm_sim* sim, void* dev_msg_lo, void* dev_msg_hi);
int crm_sim_ghosts_refresh(crm_sim* sim, const void* dev_msg_lo, const void* dev_msg_hi);
int crm_sim_step_begin(crm_sim* sim, uint32_t* dev_flags);
int crm_sim_step_end(crm_sim* sim);
/* Migration (the reference re-sorts agents into voxels / sub-domains after every step): remove the
 * owned rods whose centroid left the slab and write their full records (device state, parameters,
 * identifier, parent, global slot) into HOST buffers; append received records as owned rods. */
/* Owned rods whose centroid was outside the slab after the last step (counted by the integrator). */
uint32_t crm_sim_n_leaving(const crm_sim* sim);
int crm_sim_migrate_out(crm_sim* sim, void* host_lo, uint32_t cap_lo, uint32_t* n_lo, void* host_hi, uint32_t cap_hi,
                        uint32_t* n_hi);
int crm_sim_migrate_in(crm_sim* sim, const void* host_buf, uint32_t n);
/* Divisions of the last step in sharded mode: (global slot of the mother, user voxel) ordered by
 * global slot; the coordinator merges the lists of all slabs, derives the identifiers
 * (cellular_raza update_cell_cycle: per-voxel counter) and hands them back. */
uint32_t crm_sim_pending_divisions(const crm_sim* sim, uint64_t* gslot_mother, uint32_t* user_voxel, uint32_t cap);
int crm_sim_apply_divisions(crm_sim* sim, const crm_ident* id_mother, const crm_ident* id_daughter,
                            const uint64_t* gslot_daughter, uint32_t n);
/* Snapshot of the owned rods now (sharded runs are stepped by the coordinator, not crm_sim_run). */
int crm_sim_take_snapshot(crm_sim* sim);
int crm_sim_snapshot_global_slots(const crm_sim* sim, uint32_t k, uint64_t* out);

/* Non-tensor peaks of the device the force kernel is bounded by (not in the driver's
 * MEASURED_PEAKS.json): dense FP32 FMA TFLOP/s, MUFU.EX2 Top/s, and the SM clock the FMA
 * figure implies at 128 lanes/SM/clk. */
int crm_microbench(int device, double* fp32_tflops, double* mufu_tops, double* sm_mhz_est);

/* ---- sharded step loop in the C++ host (NCCL over NVLink) ------------------------------------------------
 * The reference's sub-domain threads exchange boundary agents without a global barrier per aspect (reference
 * src/simulation.rs:251-256,283-292).  Here one process per GPU joins a communicator; between two collective
 * rebuilds crm_sim_steps_sharded enqueues whole batches of steps with their halo send/recv and the all-gather of
 * the per-step flags on the sim's stream and reads ONE status block per batch (csrc/crm_sim.cu, crm_comm.cu).
 * NCCL is bound at run time (dlopen libnccl.so.2); without it these calls fail, nothing else is affected. */
/* 128-byte unique id made by one rank; the caller hands it to all ranks (any side channel). */
int crm_comm_unique_id(void* out128);
/* Collective: the communicator of all slabs (rank r owns slab r) on CUDA device `device`.  It outlives the sims
 * that borrow it through crm_sim_set_comm (NCCL sets its connections up lazily, once per job). */
typedef struct crm_comm crm_comm;
int crm_comm_create(const void* id128, int rank, int world, int device, crm_comm** out);
void crm_comm_destroy(crm_comm* comm);
int crm_sim_set_comm(crm_sim* sim, crm_comm* comm);
/* Collective: up to n_max list steps incl. their exchanges; the four device buffers are those of
 * crm_sim_halo_pack_msgs / crm_sim_set_ghosts_msgs.  *done = steps advanced, flags_out[4 * world] = {divisions,
 * rebuild wish, -error code, 0} per slab for the last of them. */
int crm_sim_steps_sharded(crm_sim* sim, uint32_t n_max, void* send_lo, void* send_hi, void* recv_lo, void* recv_hi,
                          uint32_t* done, uint32_t* flags_out);

/* ---- 2-D mask path: the step after the mechanics in the crm_divide objective ------------
 * Host buffers in and out; kernels in csrc/crm_raster.cu.  No graphics API is involved. */
/* Replaces `calculate_polygon_hull` (reference src/fitting.rs:514-620): the capsule-chain outline of each of
 * the n rods (pos [n][n_vertices][3], xy used).  out_xy [n][cap][2], out_count [n]. */
int crm_polygon_hulls(const float* pos, const float* radii, uint32_t n, uint32_t n_vertices, float delta_angle,
                      float epsilon, float* out_xy, uint32_t* out_count, uint32_t cap);
/* Replaces `render_mask_2d` (reference src/fitting.rs:121-244): rods in draw order (the reference iterates a
 * BTreeMap over CellIdentifier) with their colours [n][3] -> instance-colour bitmap and overlap bitmap,
 * both [resolution[1]][resolution[0]][3] u8, row 0 = largest y.  kernel_ms (may be NULL) receives the
 * device time of the kernels. */
int crm_render_mask_2d(const float* pos, const float* radii, const uint8_t* colors, uint32_t n, uint32_t n_vertices,
                       const float domain_size[2], const uint32_t resolution[2], float delta_angle, float epsilon,
                       uint8_t* out_mask, uint8_t* out_overlap, double* kernel_ms);
/* Replaces `overlap` (reference src/fitting.rs:299-345): summed intersection area of all pairs of outlines. */
int crm_overlap(const float* pos, const float* radii, uint32_t n, uint32_t n_vertices, float delta_angle, float epsilon,
                float* out_area);
/* Replaces `parents_diff_mask` (reference src/fitting.rs:42-109): 0 where the masks agree, parent_penalty where
 * one cell is the parent of the other, 1 otherwise.  Colours are decoded with color_to_counter
 * (src/imaging.rs:42-66): cell_of_counter[k] = index of the cell coloured counter_to_color(k) or -1;
 * parent_of[cell] = index of its parent or -1.  *missing_pixel = first pixel whose colour is unknown (the
 * reference raises KeyError) or -1. */
int crm_parents_diff_mask(const uint8_t* mask1, const uint8_t* mask2, uint64_t n_pixels, const int32_t* cell_of_counter,
                          uint32_t n_counters, const int32_t* parent_of, uint32_t n_cells, float parent_penalty,
                          float* out, int64_t* missing_pixel);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  `crm_overlap` without
 * any allocation: d_pos [n][n_vertices][3] f32, d_radii [n] f32, d_out_area (1 f32) and d_out_status (1 i32: 0,
 * CRM_ERR_NUMERIC "lines should be intersecting" or CRM_ERR_NOMEM "outline too long") are DEVICE pointers of the
 * current device; the kernels are enqueued on `stream` (a cudaStream_t, NULL = default stream) and the call returns
 * without waiting.  The scratch (outlines, boxes, accumulator: 2.1 MB) is a fixed device array inside the library that
 * every GPU loading it holds until exit; n > CRM_NOALLOC_MAX_RODS is rejected with CRM_ERR_INVALID.  Calls neither
 * cudaMalloc / cudaFree nor any allocating library routine.  All calls share the scratch: calls that overlap on
 * the GPU (different streams, no ordering between them) race. */
#define CRM_NOALLOC_MAX_RODS 1024u
int crm_overlap_noalloc(const float* d_pos, const float* d_radii, uint32_t n, uint32_t n_vertices, float delta_angle,
                        float epsilon, float* d_out_area, int32_t* d_out_status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CRM_B200_H */
