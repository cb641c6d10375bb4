/*
 * crm_oracle.c — CPU restatement of cr_mech_coli's per-step colony mechanics.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the checker for the CUDA path: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it.  The product (cr_mech_coli_b200) never links or calls it.
 *
 * PARITY UNPINNED for the arithmetic: the reference delegates the whole step to
 * the un-vendored git dependency cellular_raza 0.5.0 @
 * dbb0896bdedc0324fa1dea9a4739703a9bb8b512 (reference Cargo.toml:23,
 * Cargo.lock:210-300), whose source is not on this machine and which cannot be
 * built here (no Rust toolchain).  The reference's own tests hold no golden
 * vectors for forces, increments or trajectories.  What follows
 * restates the published algorithm of that crate (RodMechanics, RodInteraction,
 * MorsePotentialF32, MiePotentialF32, CartesianCuboidRods, the chili backend's
 * step order and Adams-Bashforth solver) and anchors on every in-tree witness:
 *   - call sites            reference src/simulation.rs:234-319
 *   - potential dispatch    reference src/agent.rs:382-405
 *   - neighbour sensing     reference src/agent.rs:407-434
 *   - growth / division     reference src/agent.rs:573-629
 *   - model equations       reference docs/source/model.rst:18-124
 *   - potential shapes      reference cr_mech_coli/crm_fit/plotting.py:285-309,386-391
 *   - closest point         reference plots/mechanics_schematic.py:103-131
 *   - trait signatures      reference src/crm_amir.rs:88-144,232-256
 * The decisions taken where the upstream source would be needed are listed in
 * DESIGN.md ("restatement decisions").
 *
 * All arithmetic is f32 like the reference (src/agent.rs:218,380).  Layout of a
 * rod: [n_vertices][3] row-major.  Plain C99; optional OpenMP over rods in the
 * force gather (the reference parallelises over voxel sub-domains with rayon,
 * src/simulation.rs:252,291; results do not depend on the thread count because
 * each rod's force is accumulated by one thread in a fixed order).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/crm_b200.h"

#define ORACLE_API __attribute__((visibility("default")))

typedef struct snapshot {
    uint64_t iteration;
    uint32_t n;
    float* pos;
    float* vel;
    crm_rod_params* par;
    crm_ident* id;
    crm_ident* parent;
} snapshot;

typedef struct oracle_sim {
    crm_config cfg;
    uint32_t nv, n, cap;
    float *pos, *vel;               /* [n][nv][3] */
    float *force;                   /* [n][nv][3] accumulator */
    float *dx1, *dv1, *dx2, *dv2;   /* Adams-Bashforth increment history */
    uint8_t* nhist;
    crm_rod_params* par;
    crm_ident *id, *parent;
    uint8_t* div_flag;
    uint32_t* vox_of;               /* voxel (plain index) the rod is sorted into */
    uint64_t* vox_counter;          /* per-voxel id counter */
    uint64_t iteration;
    uint64_t rod_steps;
    /* voxel lists (rebuilt each step) */
    uint32_t *vox_start, *vox_items;
    /* save schedule */
    uint64_t* save_iter;
    uint32_t n_save;
    uint64_t max_iter;
    snapshot* snaps;
    uint32_t n_snaps;
    uint64_t pairs_in_cutoff;
    /* open restatement decisions as run-time switches (same in the CUDA path) */
    int ab_hist;       /* stored increments of the Adams-Bashforth solver: 0 Euler, 1 AB2, 2 AB3 */
    float spring_w;    /* weight of a segment's spring force on each end: 1 or 1/2 */
    int mie_textbook;  /* 1: Mie constant with the textbook exponent m/(n-m) (probe only) */
    int pair_once;     /* 1: compute_forces evaluates each unordered pair once and scatters to both rods
                          (the reference backend's schedule; timing path of bench.py) */
    float* force_abs;  /* [n][nv][3] sum of |half-weighted pair contributions| per component (tolerance scale) */
    float* scratch;    /* pair_once: per-thread force accumulators */
    int scratch_threads;
    uint32_t scratch_cap;
    char err[512];
} oracle_sim;

static __thread char g_err[512];

/* ------------------------------------------------------------------------ */
/* counter-based RNG shared (by definition, not by code) with the CUDA path:  */
/* splitmix64 finaliser chain -> two 24-bit uniforms -> Box-Muller.           */
/* The reference draws from per-voxel ChaCha8 streams (cellular_raza Voxel    */
/* rng); that stream order is not reproducible on a GPU, so stochastic runs   */
/* are compared statistically against the reference and exactly against this. */
static inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
static inline void normal_pair(uint64_t seed, uint64_t a, uint64_t b, uint64_t c, float out[2]) {
    uint64_t h = mix64(mix64(mix64(mix64(seed) ^ a) ^ b) ^ c);
    float u1 = ((float)(uint32_t)(h >> 40) + 0.5f) * (1.0f / 16777216.0f);
    float u2 = ((float)(uint32_t)((h >> 16) & 0xFFFFFFu) + 0.5f) * (1.0f / 16777216.0f);
    float r = sqrtf(-2.0f * logf(u1));
    float ang = 6.28318530717958647692f * u2;
    out[0] = r * cosf(ang);
    out[1] = r * sinf(ang);
}

/* the same deviates through f64 log / sin / cos rounded to f32 (division-time draws: the rounded results of
 * glibc and of the device library agree, so drawn growth rates and thresholds are bit-identical) */
static inline void normal_pair_precise(uint64_t seed, uint64_t a, uint64_t b, uint64_t c, float out[2]) {
    uint64_t h = mix64(mix64(mix64(mix64(seed) ^ a) ^ b) ^ c);
    double u1 = ((double)(uint32_t)(h >> 40) + 0.5) * (1.0 / 16777216.0);
    double u2 = ((double)(uint32_t)((h >> 16) & 0xFFFFFFu) + 0.5) * (1.0 / 16777216.0);
    double r = sqrt(-2.0 * log(u1));
    double ang = 6.28318530717958647692 * u2;
    out[0] = (float)(r * cos(ang));
    out[1] = (float)(r * sin(ang));
}

/* ------------------------------------------------------------------------ */
/* save schedule: reference src/simulation.rs:239-249 + cellular_raza          */
/* FixedStepsize::from_partial_save_points (iteration = round((t - t0)/dt)).  */
static int build_schedule(const crm_config* c, uint64_t** out, uint32_t* n_out, uint64_t* max_iter,
                          char* err) {
    uint32_t n = c->n_saves + 2;
    uint64_t* it = (uint64_t*)malloc(sizeof(uint64_t) * n);
    float t0 = c->t0, dt = c->dt, t_max = c->t_max;
    if (!(dt > 0.0f) || !(t_max >= t0)) {
        snprintf(err, 512, "TimeError: invalid time configuration t0=%g dt=%g t_max=%g", t0, dt, t_max);
        free(it);
        return CRM_ERR_TIME;
    }
    uint32_t k = 0;
    it[k++] = (uint64_t)roundf((t0 - t0) / dt);
    if (c->n_saves > 0) {
        float dtsave = (t_max - t0) / (float)(c->n_saves + 1);
        for (uint32_t s = 1; s < c->n_saves + 1; ++s) {
            float t = t0 + (float)s * dtsave;
            it[k++] = (uint64_t)roundf((t - t0) / dt);
        }
    }
    it[k++] = (uint64_t)roundf((t_max - t0) / dt);
    /* sort + unique (BTreeMap keyed by iteration in the storage) */
    for (uint32_t i = 1; i < k; ++i) {
        uint64_t v = it[i];
        uint32_t j = i;
        while (j > 0 && it[j - 1] > v) { it[j] = it[j - 1]; --j; }
        it[j] = v;
    }
    uint32_t u = 0;
    for (uint32_t i = 0; i < k; ++i)
        if (u == 0 || it[u - 1] != it[i]) it[u++] = it[i];
    *out = it;
    *n_out = u;
    *max_iter = it[u - 1];
    return CRM_OK;
}

/* ------------------------------------------------------------------------ */
/* potentials                                                                */

/* Force on `own` located at p from a partner point q (force on the partner is
 * the negative).  Morse: V(x)=strength*(1-exp(-stiffness*(x-r)))^2, r=radius_own+radius_ext,
 * zero force beyond cutoff (reference cr_mech_coli/crm_fit/plotting.py:386-391;
 * ext_info is the partner's radius, src/agent.rs:436-447).
 * Mie: V(x)=strength*C*((sigma/x)^n-(sigma/x)^m), sigma=(m/n)^(1/(n-m))*r,
 * C=n/(n-m)*(n/m)^(n/(n-m)), repulsive magnitude clamped by `bound`, zero beyond
 * cutoff (reference cr_mech_coli/crm_fit/plotting.py:285-309). */
static int g_mie_textbook = 0; /* probe switch (oracle_set_mie_textbook): SURVEY 9.4 */
static inline void point_force(const crm_rod_params* own, float ext_radius, const float p[3],
                               const float q[3], float f[3]) {
    float zx = p[0] - q[0], zy = p[1] - q[1], zz = p[2] - q[2];
    float r = sqrtf(zx * zx + zy * zy + zz * zz);
    float dx = zx / r, dy = zy / r, dz = zz / r;
    float mag;
    if (own->potential == CRM_POT_MORSE) {
        float R = own->radius + ext_radius;
        float e = expf(-own->potential_stiffness * (r - R));
        float co = (r > own->cutoff) ? 0.0f : 1.0f;
        /* -dV/dx along dir */
        mag = -(2.0f * own->potential_stiffness * own->strength * e * (1.0f - e)) * co;
    } else {
        float R = own->radius + ext_radius;
        float en = own->en, em = own->em;
        float sigma = R * powf(em / en, 1.0f / (en - em));
        float C = en / (en - em) * powf(en / em, (g_mie_textbook ? em : en) / (en - em));
        float part = en / sigma * powf(sigma / r, en + 1.0f) - em / sigma * powf(sigma / r, em + 1.0f);
        float fm = fminf(own->strength * C * part, own->bound);
        float co = (r > own->cutoff) ? 0.0f : 1.0f;
        mag = fm * co;
    }
    f[0] = dx * mag;
    f[1] = dy * mag;
    f[2] = dz * mag;
}

/* RodInteraction::calculate_force_between (cellular_raza bacterial_rods; model
 * text reference docs/source/model.rst:76-90): every own vertex against every
 * partner segment; the partner's share is split (1-q, q) onto the segment ends. */
static void rod_force_between(const crm_rod_params* own, uint32_t nv, const float* own_pos,
                              const float* ext_pos, float ext_radius, float* f_own, float* f_ext) {
    memset(f_own, 0, sizeof(float) * nv * 3);
    memset(f_ext, 0, sizeof(float) * nv * 3);
    for (uint32_t i = 0; i < nv; ++i) {
        const float* p = own_pos + 3 * i;
        for (uint32_t j = 0; j + 1 < nv; ++j) {
            const float* a = ext_pos + 3 * j;
            const float* b = ext_pos + 3 * (j + 1);
            /* nearest point on segment [a,b] (reference plots/mechanics_schematic.py:103-131) */
            float abx = b[0] - a[0], aby = b[1] - a[1], abz = b[2] - a[2];
            float apx = p[0] - a[0], apy = p[1] - a[1], apz = p[2] - a[2];
            float t = (abx * apx + aby * apy + abz * apz) / (abx * abx + aby * aby + abz * abz);
            t = fminf(fmaxf(t, 0.0f), 1.0f);
            float near[3] = {a[0] * (1.0f - t) + b[0] * t, a[1] * (1.0f - t) + b[1] * t,
                             a[2] * (1.0f - t) + b[2] * t};
            float f[3];
            point_force(own, ext_radius, p, near, f);
            for (int c = 0; c < 3; ++c) {
                f_own[3 * i + c] += f[c];
                f_ext[3 * j + c] += -f[c] * (1.0f - t);
                f_ext[3 * ((j + 1) % nv) + c] += -f[c] * t;
            }
        }
    }
}

/* NeighborSensing::accumulate_information (reference src/agent.rs:407-424). */
static int senses(uint32_t nv, const float* own_pos, const float* ext_pos, float r_own, float r_ext) {
    float thr = (r_own + r_ext) / 1.41421356237309504880f;
    for (uint32_t i = 0; i < nv; ++i)
        for (uint32_t j = 0; j < nv; ++j) {
            float dx = own_pos[3 * i] - ext_pos[3 * j];
            float dy = own_pos[3 * i + 1] - ext_pos[3 * j + 1];
            float dz = own_pos[3 * i + 2] - ext_pos[3 * j + 2];
            if (sqrtf(dx * dx + dy * dy + dz * dz) < thr) return 1;
        }
    return 0;
}

/* is the minimum vertex-segment distance of the ordered pair below the cutoff?
 * (only used to count the algorithmic work for the roofline) */
static int within_cutoff(uint32_t nv, const float* A, const float* B, float cutoff) {
    for (uint32_t i = 0; i < nv; ++i)
        for (uint32_t j = 0; j + 1 < nv; ++j) {
            const float* p = A + 3 * i;
            const float* a = B + 3 * j;
            const float* b = B + 3 * (j + 1);
            float abx = b[0] - a[0], aby = b[1] - a[1], abz = b[2] - a[2];
            float apx = p[0] - a[0], apy = p[1] - a[1], apz = p[2] - a[2];
            float t = (abx * apx + aby * apy + abz * apz) / (abx * abx + aby * aby + abz * abz);
            t = fminf(fmaxf(t, 0.0f), 1.0f);
            float dx = p[0] - (a[0] * (1.0f - t) + b[0] * t);
            float dy = p[1] - (a[1] * (1.0f - t) + b[1] * t);
            float dz = p[2] - (a[2] * (1.0f - t) + b[2] * t);
            if (sqrtf(dx * dx + dy * dy + dz * dz) <= cutoff) return 1;
        }
    return 0;
}

/* ------------------------------------------------------------------------ */
/* voxels: CartesianCuboid::from_boundaries_and_n_voxels([0;3],[Lx,Ly,H],[nx,ny,1])
 * (reference src/simulation.rs:258-269); a rod lives in the voxel of the mean
 * of its vertices; plain index = ix*ny + iy (nz = 1). */
static uint32_t voxel_of(const oracle_sim* s, const float* pos) {
    float mx = 0.0f, my = 0.0f;
    for (uint32_t i = 0; i < s->nv; ++i) {
        mx += pos[3 * i];
        my += pos[3 * i + 1];
    }
    mx /= (float)s->nv;
    my /= (float)s->nv;
    float dxv = s->cfg.domain_size[0] / (float)s->cfg.n_voxels[0];
    float dyv = s->cfg.domain_size[1] / (float)s->cfg.n_voxels[1];
    int64_t ix = (int64_t)floorf(mx / dxv);
    int64_t iy = (int64_t)floorf(my / dyv);
    if (ix < 0) ix = 0;
    if (iy < 0) iy = 0;
    if (ix >= (int64_t)s->cfg.n_voxels[0]) ix = s->cfg.n_voxels[0] - 1;
    if (iy >= (int64_t)s->cfg.n_voxels[1]) iy = s->cfg.n_voxels[1] - 1;
    return (uint32_t)(ix * s->cfg.n_voxels[1] + iy);
}

static void sort_into_voxels(oracle_sim* s) {
    uint32_t nvox = s->cfg.n_voxels[0] * s->cfg.n_voxels[1];
    for (uint32_t a = 0; a < s->n; ++a) s->vox_of[a] = voxel_of(s, s->pos + (size_t)a * s->nv * 3);
    memset(s->vox_start, 0, sizeof(uint32_t) * (nvox + 1));
    for (uint32_t a = 0; a < s->n; ++a) s->vox_start[s->vox_of[a] + 1]++;
    for (uint32_t v = 0; v < nvox; ++v) s->vox_start[v + 1] += s->vox_start[v];
    uint32_t* fill = (uint32_t*)calloc(nvox, sizeof(uint32_t));
    for (uint32_t a = 0; a < s->n; ++a) {
        uint32_t v = s->vox_of[a];
        s->vox_items[s->vox_start[v] + fill[v]++] = a; /* slot order inside a voxel */
    }
    free(fill);
}

/* ------------------------------------------------------------------------ */
/* Interaction + neighbour sensing + domain force for the current state.
 * Restates the chili backend's update_mechanics_interaction_step_1..3: inside a
 * voxel and across the 3x3 neighbour voxels every ordered pair is evaluated
 * twice, calculate_force_between(A,B) and (B,A), each added with weight 1/2
 * (SURVEY §9.1).  Written as a gather: rod A sums 1/2 F_own(A,B) + 1/2 F_ext(B,A). */
static void compute_forces(oracle_sim* s, int count_pairs) {
    const uint32_t nv = s->nv;
    const int nx = (int)s->cfg.n_voxels[0], ny = (int)s->cfg.n_voxels[1];
    sort_into_voxels(s);
    uint64_t pairs = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : pairs)
#endif
    for (int64_t a = 0; a < (int64_t)s->n; ++a) {
        float fo[3 * 64], fe[3 * 64], fo2[3 * 64], fe2[3 * 64];
        float* F = s->force + (size_t)a * nv * 3;
        const float* A = s->pos + (size_t)a * nv * 3;
        const crm_rod_params* pa = &s->par[a];
        memset(F, 0, sizeof(float) * nv * 3);
        float* FA = s->force_abs ? s->force_abs + (size_t)a * nv * 3 : NULL;
        if (FA) memset(FA, 0, sizeof(float) * nv * 3);
        uint32_t count = 0;
        int vx = (int)(s->vox_of[a] / (uint32_t)ny), vy = (int)(s->vox_of[a] % (uint32_t)ny);
        for (int ix = vx - 1; ix <= vx + 1; ++ix) {
            if (ix < 0 || ix >= nx) continue;
            for (int iy = vy - 1; iy <= vy + 1; ++iy) {
                if (iy < 0 || iy >= ny) continue;
                uint32_t v = (uint32_t)(ix * ny + iy);
                for (uint32_t k = s->vox_start[v]; k < s->vox_start[v + 1]; ++k) {
                    uint32_t b = s->vox_items[k];
                    if (b == (uint32_t)a) continue;
                    const float* B = s->pos + (size_t)b * nv * 3;
                    const crm_rod_params* pb = &s->par[b];
                    rod_force_between(pa, nv, A, B, pb->radius, fo, fe);   /* A's own share  */
                    rod_force_between(pb, nv, B, A, pa->radius, fo2, fe2); /* B's call, ext=A */
                    for (uint32_t t = 0; t < nv * 3; ++t) {
                        F[t] += 0.5f * fo[t];
                        F[t] += 0.5f * fe2[t];
                        if (FA) FA[t] += 0.5f * fabsf(fo[t]) + 0.5f * fabsf(fe2[t]);
                    }
                    count += (uint32_t)senses(nv, A, B, pa->radius, pb->radius);
                    if (count_pairs) pairs += (uint64_t)within_cutoff(nv, A, B, pa->cutoff);
                }
            }
        }
        s->par[a].n_neighbors = count; /* react_to_neighbors, reference src/agent.rs:430-433 */
        /* DomainForce: CartesianSubDomainRods::calculate_custom_force(pos, vel, radius).
         * gel pressure pushes every vertex towards -z; vertices within
         * surface_friction_distance of the bottom plate feel -surface_friction*vel
         * (reference src/config.rs:426-431, src/simulation.rs:270-275). */
        const float* V = s->vel + (size_t)a * nv * 3;
        for (uint32_t i = 0; i < nv; ++i) {
            F[3 * i + 2] -= s->cfg.gel_pressure;
            if (s->cfg.surface_friction != 0.0f && A[3 * i + 2] <= s->cfg.surface_friction_distance) {
                F[3 * i + 0] -= s->cfg.surface_friction * V[3 * i + 0];
                F[3 * i + 1] -= s->cfg.surface_friction * V[3 * i + 1];
                F[3 * i + 2] -= s->cfg.surface_friction * V[3 * i + 2];
            }
        }
    }
    if (count_pairs) s->pairs_in_cutoff = pairs;
}

/* The same forces with the reference backend's schedule: every unordered pair of the
 * 3x3 voxel stencil is visited once, calculate_force_between is called for (A,B) and for (B,A), and
 * the four half-weighted results are scattered to both rods.  Half the point-segment evaluations of
 * the gather form above; the f32 summation order differs (per-thread accumulators are reduced at the
 * end), so this is the TIMING path of bench.py's CPU baseline, the gather form stays the checker. */
static void compute_forces_pair_once(oracle_sim* s) {
    const uint32_t nv = s->nv;
    const size_t m = (size_t)nv * 3;
    const int nx = (int)s->cfg.n_voxels[0], ny = (int)s->cfg.n_voxels[1];
    sort_into_voxels(s);
    int nt = 1;
#ifdef _OPENMP
    nt = omp_get_max_threads();
#endif
    if (s->scratch_threads != nt || s->scratch_cap < s->n) {
        free(s->scratch);
        s->scratch = (float*)malloc(sizeof(float) * (size_t)nt * s->n * m);
        s->scratch_threads = nt;
        s->scratch_cap = s->n;
    }
    uint32_t* counts = (uint32_t*)calloc((size_t)nt * s->n, sizeof(uint32_t));
#ifdef _OPENMP
#pragma omp parallel
#endif
    {
        int tid = 0;
#ifdef _OPENMP
        tid = omp_get_thread_num();
#endif
        float* acc = s->scratch + (size_t)tid * s->n * m;
        uint32_t* cnt = counts + (size_t)tid * s->n;
        memset(acc, 0, sizeof(float) * (size_t)s->n * m);
        float fo[3 * 64], fe[3 * 64], fo2[3 * 64], fe2[3 * 64];
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 16)
#endif
        for (int64_t a = 0; a < (int64_t)s->n; ++a) {
            const float* A = s->pos + (size_t)a * m;
            const crm_rod_params* pa = &s->par[a];
            int vx = (int)(s->vox_of[a] / (uint32_t)ny), vy = (int)(s->vox_of[a] % (uint32_t)ny);
            for (int ix = vx - 1; ix <= vx + 1; ++ix) {
                if (ix < 0 || ix >= nx) continue;
                for (int iy = vy - 1; iy <= vy + 1; ++iy) {
                    if (iy < 0 || iy >= ny) continue;
                    uint32_t v = (uint32_t)(ix * ny + iy);
                    for (uint32_t k = s->vox_start[v]; k < s->vox_start[v + 1]; ++k) {
                        uint32_t b = s->vox_items[k];
                        if (b <= (uint32_t)a) continue; /* each unordered pair once */
                        const float* B = s->pos + (size_t)b * m;
                        const crm_rod_params* pb = &s->par[b];
                        rod_force_between(pa, nv, A, B, pb->radius, fo, fe);
                        rod_force_between(pb, nv, B, A, pa->radius, fo2, fe2);
                        float* FA = acc + (size_t)a * m;
                        float* FB = acc + (size_t)b * m;
                        for (uint32_t t = 0; t < m; ++t) {
                            FA[t] += 0.5f * fo[t] + 0.5f * fe2[t];
                            FB[t] += 0.5f * fe[t] + 0.5f * fo2[t];
                        }
                        const uint32_t sn = (uint32_t)senses(nv, A, B, pa->radius, pb->radius); /* symmetric */
                        cnt[a] += sn;
                        cnt[b] += sn;
                    }
                }
            }
        }
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
        for (int64_t a = 0; a < (int64_t)s->n; ++a) {
            float* F = s->force + (size_t)a * m;
            const float* A = s->pos + (size_t)a * m;
            const float* V = s->vel + (size_t)a * m;
            uint32_t c = 0;
            memset(F, 0, sizeof(float) * m);
            for (int t = 0; t < nt; ++t) {
                const float* src = s->scratch + ((size_t)t * s->n + a) * m;
                for (uint32_t q = 0; q < m; ++q) F[q] += src[q];
                c += counts[(size_t)t * s->n + a];
            }
            s->par[a].n_neighbors = c;
            for (uint32_t i = 0; i < nv; ++i) {
                F[3 * i + 2] -= s->cfg.gel_pressure;
                if (s->cfg.surface_friction != 0.0f && A[3 * i + 2] <= s->cfg.surface_friction_distance) {
                    F[3 * i + 0] -= s->cfg.surface_friction * V[3 * i + 0];
                    F[3 * i + 1] -= s->cfg.surface_friction * V[3 * i + 1];
                    F[3 * i + 2] -= s->cfg.surface_friction * V[3 * i + 2];
                }
            }
        }
    }
    free(counts);
}

/* RodMechanics::calculate_increment (reference docs/source/model.rst:28-57;
 * delegation witnessed at src/crm_amir.rs:88-91): springs between consecutive
 * vertices, restoring bending force eta*(pi-alpha) on interior vertices with
 * -1/2 on both neighbours, damping; returns (dx, dv) = (vel, total force). */
static void increment(const oracle_sim* s, uint32_t a, float* dv) {
    const uint32_t nv = s->nv;
    const float* P = s->pos + (size_t)a * nv * 3;
    const float* V = s->vel + (size_t)a * nv * 3;
    const float* F = s->force + (size_t)a * nv * 3;
    const crm_rod_params* p = &s->par[a];
    memcpy(dv, F, sizeof(float) * nv * 3);
    for (uint32_t i = 0; i + 1 < nv; ++i) {
        float c[3] = {P[3 * i] - P[3 * i + 3], P[3 * i + 1] - P[3 * i + 4], P[3 * i + 2] - P[3 * i + 5]};
        float len = sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
        if (len != 0.0f) {
            float k = (p->spring_tension * (len - p->spring_length)) * s->spring_w;
            for (int d = 0; d < 3; ++d) {
                float f = -(c[d] / len) * k;
                dv[3 * i + d] += f;
                dv[3 * i + 3 + d] -= f;
            }
        }
    }
    for (uint32_t i = 1; i + 1 < nv; ++i) {
        float c1[3], c2[3];
        for (int d = 0; d < 3; ++d) {
            c1[d] = P[3 * (i - 1) + d] - P[3 * i + d];
            c2[d] = P[3 * i + d] - P[3 * (i + 1) + d];
        }
        float n1 = sqrtf(c1[0] * c1[0] + c1[1] * c1[1] + c1[2] * c1[2]);
        float n2 = sqrtf(c2[0] * c2[0] + c2[1] * c2[1] + c2[2] * c2[2]);
        /* angle between c1 and -c2 (nalgebra Matrix::angle: 0 if a norm vanishes) */
        float angle;
        if (n1 == 0.0f || n2 == 0.0f) {
            angle = 0.0f;
        } else {
            float cs = -(c1[0] * c2[0] + c1[1] * c2[1] + c1[2] * c2[2]) / (n1 * n2);
            cs = fminf(fmaxf(cs, -1.0f), 1.0f);
            angle = acosf(cs);
        }
        float fd[3] = {c1[0] / n1 - c2[0] / n2, c1[1] / n1 - c2[1] / n2, c1[2] / n1 - c2[2] / n2};
        float nfd = sqrtf(fd[0] * fd[0] + fd[1] * fd[1] + fd[2] * fd[2]);
        float scale = p->rigidity * (3.14159265358979323846f - angle);
        for (int d = 0; d < 3; ++d) {
            float dirn = (nfd == 0.0f) ? fd[d] : fd[d] / nfd;
            float f = dirn * scale;
            dv[3 * (i - 1) + d] -= 0.5f * f;
            dv[3 * i + d] += f;
            dv[3 * (i + 1) + d] -= 0.5f * f;
        }
    }
    for (uint32_t t = 0; t < nv * 3; ++t) dv[t] -= p->damping * V[t];
}

/* chili mechanics solver: Adams-Bashforth of order up to 3 over the stored
 * increments, Euler / AB2 during start-up; the Wiener term of
 * RodMechanics::get_random_contribution is added outside the multistep sum. */
static void integrate(oracle_sim* s, uint32_t a, const float* dv) {
    const uint32_t m = s->nv * 3;
    const float dt = s->cfg.dt;
    float* P = s->pos + (size_t)a * m;
    float* V = s->vel + (size_t)a * m;
    float* dx1 = s->dx1 + (size_t)a * m;
    float* dv1 = s->dv1 + (size_t)a * m;
    float* dx2 = s->dx2 + (size_t)a * m;
    float* dv2 = s->dv2 + (size_t)a * m;
    const float c23 = 23.0f / 12.0f, c16 = 16.0f / 12.0f, c5 = 5.0f / 12.0f;
    const float D = s->par[a].diffusion_constant;
    const float amp = (D > 0.0f) ? sqrtf(2.0f * D * dt) : 0.0f;
    for (uint32_t t = 0; t < m; ++t) {
        float dx = V[t];
        float dvv = dv[t];
        float ip, iv;
        if (s->nhist[a] >= 2) {
            ip = (c23 * dx - c16 * dx1[t] + c5 * dx2[t]) * dt;
            iv = (c23 * dvv - c16 * dv1[t] + c5 * dv2[t]) * dt;
        } else if (s->nhist[a] == 1) {
            ip = (1.5f * dx - 0.5f * dx1[t]) * dt;
            iv = (1.5f * dvv - 0.5f * dv1[t]) * dt;
        } else {
            ip = dx * dt;
            iv = dvv * dt;
        }
        if (amp != 0.0f) {
            float nrm[2];
            normal_pair(s->cfg.rng_seed, s->iteration, a, t, nrm);
            ip += amp * nrm[0];
        }
        dx2[t] = dx1[t];
        dv2[t] = dv1[t];
        dx1[t] = dx;
        dv1[t] = dvv;
        P[t] += ip;
        V[t] += iv;
    }
    if (s->nhist[a] < s->ab_hist) s->nhist[a]++;
}

/* Cycle::update_cycle (reference src/agent.rs:573-595). */
static int update_cycle(oracle_sim* s, uint32_t a) {
    crm_rod_params* p = &s->par[a];
    float rate;
    if (p->has_neighbor_reduction) {
        float m = (float)p->neighbor_reduction_max;
        float n = (float)p->n_neighbors;
        rate = p->growth_rate * powf(fmaxf((m - n) / m, 0.0f), p->neighbor_reduction_exp);
    } else {
        rate = p->growth_rate;
    }
    p->spring_length += rate * s->cfg.dt * p->spring_length;
    return p->spring_length > p->spring_length_threshold;
}

static int ensure_capacity(oracle_sim* s, uint32_t need) {
    if (need <= s->cap) return 0;
    uint32_t cap = s->cap ? s->cap : 16;
    while (cap < need) cap *= 2;
    size_t m = (size_t)s->nv * 3;
#define GROW(ptr, type, count) ptr = (type*)realloc(ptr, sizeof(type) * (count))
    GROW(s->pos, float, cap * m);
    GROW(s->vel, float, cap * m);
    GROW(s->force, float, cap * m);
    GROW(s->dx1, float, cap * m);
    GROW(s->dv1, float, cap * m);
    GROW(s->dx2, float, cap * m);
    GROW(s->dv2, float, cap * m);
    GROW(s->nhist, uint8_t, cap);
    GROW(s->par, crm_rod_params, cap);
    GROW(s->id, crm_ident, cap);
    GROW(s->parent, crm_ident, cap);
    GROW(s->div_flag, uint8_t, cap);
    GROW(s->vox_of, uint32_t, cap);
    GROW(s->vox_items, uint32_t, cap);
#undef GROW
    s->cap = cap;
    return 0;
}

/* RodMechanics::divide(radius) (geometry described at reference
 * docs/source/model.rst:112-124) + Cycle::divide (reference
 * src/agent.rs:597-629).  The mother becomes daughter 1 (first half of the
 * polyline), the returned clone daughter 2 (second half); both keep the
 * mother's velocities; spring_length shrinks by 1/2 - radius/(n_vertices*spring_length). */
static void divide_positions(uint32_t nv, const float* P, float radius, float* d1, float* d2) {
    float seg[64], cum[65];
    float total = 0.0f;
    cum[0] = 0.0f;
    for (uint32_t k = 0; k + 1 < nv; ++k) {
        float dx = P[3 * k] - P[3 * k + 3], dy = P[3 * k + 1] - P[3 * k + 4], dz = P[3 * k + 2] - P[3 * k + 5];
        seg[k] = sqrtf(dx * dx + dy * dy + dz * dz);
        total += seg[k];
        cum[k + 1] = total;
    }
    float seg_len = (total - 2.0f * radius) / (float)(nv - 1) / 2.0f;
    for (uint32_t n = 0; n < nv; ++n) {
        float sarc = (float)n * seg_len;
        /* forward walk */
        uint32_t k = 0;
        for (uint32_t j = 0; j + 1 < nv; ++j)
            if (cum[j] <= sarc) k = j;
        float q = (sarc - cum[k]) / seg[k];
        for (int d = 0; d < 3; ++d) d1[3 * n + d] = P[3 * k + d] * (1.0f - q) + P[3 * (k + 1) + d] * q;
        /* backward walk */
        float cr = 0.0f, cm = 0.0f;
        uint32_t m = 0;
        for (uint32_t j = 0; j + 1 < nv; ++j) {
            if (cr <= sarc) { m = j; cm = cr; }
            cr += seg[nv - 2 - j];
        }
        float p = (sarc - cm) / seg[nv - 2 - m];
        for (int d = 0; d < 3; ++d)
            d2[3 * (nv - 1 - n) + d] = P[3 * (nv - 1 - m) + d] * (1.0f - p) + P[3 * (nv - 2 - m) + d] * p;
    }
}

static float setter_sample(int kind, float a, float b, int which, uint64_t seed, uint64_t it,
                           uint64_t slot, uint64_t stream) {
    if (kind == CRM_SETTER_EXPLICIT) return which == 0 ? a : b;
    if (b == 0.0f) return a; /* Normal(mean, 0) */
    float nrm[2];
    normal_pair_precise(seed ^ 0xD1B54A32D192ED03ULL, it, slot, stream, nrm);
    return a + b * nrm[which];
}

/* update_cell_cycle_4 of the chili backend restated in slot order: inside a
 * voxel mothers are renamed first (pre-incremented per-voxel counter), then the
 * new cells are appended and named (two new identifiers per
 * division with parent = mother, reference src/crm_divide.rs:49-59,204-216). */
static int do_divisions(oracle_sim* s) {
    uint32_t n0 = s->n, nd = 0;
    const uint32_t nv = s->nv;
    for (uint32_t a = 0; a < n0; ++a) nd += s->div_flag[a];
    if (nd == 0) return 0;
    ensure_capacity(s, n0 + nd);
    uint32_t nvox = s->cfg.n_voxels[0] * s->cfg.n_voxels[1];
    uint32_t* per_vox = (uint32_t*)calloc(nvox, sizeof(uint32_t));
    uint32_t* rank_vox = (uint32_t*)calloc(nvox, sizeof(uint32_t));
    for (uint32_t a = 0; a < n0; ++a)
        if (s->div_flag[a]) per_vox[s->vox_of[a]]++;
    uint32_t k = 0;
    float d1[3 * 64], d2[3 * 64];
    for (uint32_t a = 0; a < n0; ++a) {
        if (!s->div_flag[a]) continue;
        uint32_t b = n0 + k++;
        uint32_t v = s->vox_of[a];
        size_t m = (size_t)nv * 3;
        crm_rod_params* pa = &s->par[a];
        if (!(pa->radius > 0.0f) || !isfinite(pa->spring_length)) {
            snprintf(s->err, sizeof(s->err), "DivisionError: cannot divide rod %u", a);
            free(per_vox);
            free(rank_vox);
            return CRM_ERR_DIVISION;
        }
        divide_positions(nv, s->pos + a * m, pa->radius, d1, d2);
        float div_factor = 0.5f - pa->radius / ((float)nv * pa->spring_length);
        pa->spring_length = div_factor * pa->spring_length;
        memcpy(s->pos + a * m, d1, sizeof(float) * m);
        /* daughter 2 = clone of the (already updated) mother + second half */
        s->par[b] = *pa;
        memcpy(s->pos + b * m, d2, sizeof(float) * m);
        memcpy(s->vel + b * m, s->vel + a * m, sizeof(float) * m);
        float g1 = setter_sample(pa->growth_setter_kind, pa->growth_setter_a, pa->growth_setter_b, 0,
                                 s->cfg.rng_seed, s->iteration, a, 0);
        float g2 = setter_sample(pa->growth_setter_kind, pa->growth_setter_a, pa->growth_setter_b, 1,
                                 s->cfg.rng_seed, s->iteration, a, 0);
        float l1 = setter_sample(pa->threshold_setter_kind, pa->threshold_setter_a,
                                 pa->threshold_setter_b, 0, s->cfg.rng_seed, s->iteration, a, 1);
        float l2 = setter_sample(pa->threshold_setter_kind, pa->threshold_setter_a,
                                 pa->threshold_setter_b, 1, s->cfg.rng_seed, s->iteration, a, 1);
        pa->growth_rate = g1;
        pa->spring_length_threshold = l1;
        s->par[b].growth_rate = g2;
        s->par[b].spring_length_threshold = l2;
        /* aux storage: the mother's slot keeps its increment history, the new cell starts empty */
        s->nhist[b] = 0;
        memset(s->dx1 + b * m, 0, sizeof(float) * m);
        memset(s->dv1 + b * m, 0, sizeof(float) * m);
        memset(s->dx2 + b * m, 0, sizeof(float) * m);
        memset(s->dv2 + b * m, 0, sizeof(float) * m);
        /* identifiers */
        crm_ident mother = s->id[a];
        uint32_t r = rank_vox[v]++;
        s->id[a].kind = CRM_ID_DIVISION;
        s->id[a].voxel = v;
        s->id[a].counter = s->vox_counter[v] + 1 + r;
        s->parent[a] = mother;
        s->id[b].kind = CRM_ID_DIVISION;
        s->id[b].voxel = v;
        s->id[b].counter = s->vox_counter[v] + per_vox[v] + 1 + r;
        s->parent[b] = mother;
        s->vox_of[b] = v;
        s->div_flag[a] = 0;
        s->div_flag[b] = 0;
    }
    for (uint32_t v = 0; v < nvox; ++v) s->vox_counter[v] += 2ull * per_vox[v];
    s->n = n0 + nd;
    free(per_vox);
    free(rank_vox);
    return 0;
}

/* SubDomainMechanics::apply_boundary of CartesianSubDomainRods: reflect each
 * vertex coordinate at the cuboid faces, error if still outside (error text
 * witnessed at reference docs/source/_static/2025-04-30-presentation-short/
 * figures/crm_fit/0012/logs.txt:2). */
static int apply_boundary(oracle_sim* s) {
    const float lo[3] = {0.0f, 0.0f, 0.0f};
    const float hi[3] = {s->cfg.domain_size[0], s->cfg.domain_size[1], s->cfg.domain_height};
    for (uint32_t a = 0; a < s->n; ++a) {
        float* P = s->pos + (size_t)a * s->nv * 3;
        float* V = s->vel + (size_t)a * s->nv * 3;
        for (uint32_t i = 0; i < s->nv; ++i)
            for (int d = 0; d < 3; ++d) {
                float* p = &P[3 * i + d];
                float* v = &V[3 * i + d];
                if (*p < lo[d]) { *p = 2.0f * lo[d] - *p; *v = fabsf(*v); }
                if (*p > hi[d]) { *p = 2.0f * hi[d] - *p; *v = -fabsf(*v); }
                if (!(*p >= lo[d] && *p <= hi[d])) {
                    if (isnan(*p))
                        snprintf(s->err, sizeof(s->err), "CalcError: NaN in position of rod %u", a);
                    else
                        snprintf(s->err, sizeof(s->err),
                                 "BoundaryError: Particle is out of domain at pos [%g, %g, %g] (rod %u)",
                                 P[3 * i], P[3 * i + 1], P[3 * i + 2], a);
                    return isnan(*p) ? CRM_ERR_NUMERIC : CRM_ERR_BOUNDARY;
                }
            }
    }
    return 0;
}

static int step_once(oracle_sim* s) {
    s->iteration += 1;
    s->rod_steps += s->n;
    if (s->pair_once) compute_forces_pair_once(s);
    else compute_forces(s, 0);
    float dv[3 * 64];
    for (uint32_t a = 0; a < s->n; ++a) {
        increment(s, a, dv);
        integrate(s, a, dv);
        s->div_flag[a] = (uint8_t)update_cycle(s, a);
    }
    int rc = do_divisions(s);
    if (rc) return rc;
    return apply_boundary(s);
}

static void take_snapshot(oracle_sim* s) {
    s->snaps = (snapshot*)realloc(s->snaps, sizeof(snapshot) * (s->n_snaps + 1));
    snapshot* sn = &s->snaps[s->n_snaps++];
    size_t m = (size_t)s->n * s->nv * 3;
    sn->iteration = s->iteration;
    sn->n = s->n;
    sn->pos = (float*)malloc(sizeof(float) * m);
    sn->vel = (float*)malloc(sizeof(float) * m);
    sn->par = (crm_rod_params*)malloc(sizeof(crm_rod_params) * s->n);
    sn->id = (crm_ident*)malloc(sizeof(crm_ident) * s->n);
    sn->parent = (crm_ident*)malloc(sizeof(crm_ident) * s->n);
    memcpy(sn->pos, s->pos, sizeof(float) * m);
    memcpy(sn->vel, s->vel, sizeof(float) * m);
    memcpy(sn->par, s->par, sizeof(crm_rod_params) * s->n);
    memcpy(sn->id, s->id, sizeof(crm_ident) * s->n);
    memcpy(sn->parent, s->parent, sizeof(crm_ident) * s->n);
}

/* ------------------------------------------------------------------------ */
/* exported API: mirrors include/crm_b200.h with the prefix oracle_           */

ORACLE_API const char* oracle_last_error(void) { return g_err; }

ORACLE_API int oracle_save_schedule(const crm_config* c, uint64_t* out, uint32_t cap, uint64_t* n_it) {
    uint64_t* it;
    uint32_t n;
    uint64_t mx;
    int rc = build_schedule(c, &it, &n, &mx, g_err);
    if (rc) return rc;
    for (uint32_t i = 0; i < n && i < cap; ++i) out[i] = it[i];
    if (n_it) *n_it = mx;
    free(it);
    return (int)n;
}

ORACLE_API int oracle_sim_create(const crm_config* c, int device, oracle_sim** out) {
    (void)device;
    oracle_sim* s = (oracle_sim*)calloc(1, sizeof(oracle_sim));
    s->cfg = *c;
    s->ab_hist = 2;
    s->spring_w = 1.0f;
    if (c->n_voxels[0] == 0 || c->n_voxels[1] == 0 || !(c->domain_size[0] > 0) || !(c->domain_size[1] > 0) ||
        !(c->domain_height > 0)) {
        snprintf(g_err, sizeof(g_err), "invalid domain configuration");
        free(s);
        return CRM_ERR_INVALID;
    }
    int rc = build_schedule(c, &s->save_iter, &s->n_save, &s->max_iter, g_err);
    if (rc) { free(s); return rc; }
    uint32_t nvox = c->n_voxels[0] * c->n_voxels[1];
    s->vox_counter = (uint64_t*)calloc(nvox, sizeof(uint64_t));
    s->vox_start = (uint32_t*)calloc(nvox + 1, sizeof(uint32_t));
    *out = s;
    return 0;
}

ORACLE_API void oracle_sim_destroy(oracle_sim* s) {
    if (!s) return;
    for (uint32_t k = 0; k < s->n_snaps; ++k) {
        free(s->snaps[k].pos); free(s->snaps[k].vel); free(s->snaps[k].par);
        free(s->snaps[k].id); free(s->snaps[k].parent);
    }
    free(s->snaps);
    free(s->pos); free(s->vel); free(s->force); free(s->dx1); free(s->dv1); free(s->dx2); free(s->dv2);
    free(s->nhist); free(s->par); free(s->id); free(s->parent); free(s->div_flag); free(s->vox_of);
    free(s->vox_items); free(s->vox_counter); free(s->vox_start); free(s->save_iter);
    free(s->scratch);
    free(s);
}

ORACLE_API int oracle_sim_set_agents(oracle_sim* s, const float* pos, const float* vel,
                                     const crm_rod_params* par, uint32_t n, uint32_t nv) {
    if (nv < 2 || nv > 64) {
        snprintf(g_err, sizeof(g_err), "n_vertices must be in [2, 64]");
        return CRM_ERR_INVALID;
    }
    s->nv = nv;
    s->n = 0;
    ensure_capacity(s, n ? n : 1);
    size_t m = (size_t)n * nv * 3;
    memcpy(s->pos, pos, sizeof(float) * m);
    memcpy(s->vel, vel, sizeof(float) * m);
    memcpy(s->par, par, sizeof(crm_rod_params) * n);
    memset(s->nhist, 0, n);
    memset(s->div_flag, 0, n);
    for (uint32_t a = 0; a < n; ++a) {
        s->id[a].kind = CRM_ID_INITIAL;
        s->id[a].voxel = 0;
        s->id[a].counter = a;
        s->parent[a].kind = -1;
        s->parent[a].voxel = 0;
        s->parent[a].counter = 0;
    }
    s->n = n;
    s->iteration = 0;
    return 0;
}

ORACLE_API int oracle_sim_step(oracle_sim* s, uint32_t n_steps) {
    for (uint32_t k = 0; k < n_steps; ++k) {
        int rc = step_once(s);
        if (rc) { memcpy(g_err, s->err, sizeof(g_err)); return rc; }
    }
    return 0;
}

ORACLE_API int oracle_sim_run(oracle_sim* s) {
    uint32_t next = 0;
    if (next < s->n_save && s->save_iter[next] == s->iteration) { take_snapshot(s); next++; }
    while (s->iteration < s->max_iter) {
        int rc = step_once(s);
        if (rc) { memcpy(g_err, s->err, sizeof(g_err)); return rc; }
        while (next < s->n_save && s->save_iter[next] < s->iteration) next++;
        if (next < s->n_save && s->save_iter[next] == s->iteration) { take_snapshot(s); next++; }
    }
    return 0;
}

ORACLE_API int oracle_sim_compute_forces(oracle_sim* s, float* out_force, uint32_t* out_nn) {
    compute_forces(s, 1);
    if (out_force) memcpy(out_force, s->force, sizeof(float) * (size_t)s->n * s->nv * 3);
    if (out_nn)
        for (uint32_t a = 0; a < s->n; ++a) out_nn[a] = s->par[a].n_neighbors;
    return 0;
}

/* switches for the open restatement decisions and the timing schedule */
ORACLE_API int oracle_sim_set_solver_order(oracle_sim* s, int stored_increments) {
    if (stored_increments < 0 || stored_increments > 2) return CRM_ERR_INVALID;
    s->ab_hist = stored_increments;
    return 0;
}
ORACLE_API int oracle_sim_set_spring_end_weight(oracle_sim* s, float w) {
    s->spring_w = w;
    return 0;
}
ORACLE_API void oracle_set_mie_textbook(int on) { g_mie_textbook = on; }
ORACLE_API int oracle_sim_set_pair_once(oracle_sim* s, int on) {
    s->pair_once = on;
    return 0;
}
ORACLE_API int oracle_sim_last_forces(oracle_sim* s, float* out_force) {
    memcpy(out_force, s->force, sizeof(float) * (size_t)s->n * s->nv * 3);
    return 0;
}

/* Per component: sum of the absolute values of the half-weighted pair contributions that make up the
 * interaction force of the current state (the scale of the f32 summation noise of that component). */
ORACLE_API int oracle_sim_force_abs_sum(oracle_sim* s, float* out) {
    size_t m = (size_t)s->n * s->nv * 3;
    s->force_abs = (float*)malloc(sizeof(float) * (m ? m : 1));
    compute_forces(s, 0);
    memcpy(out, s->force_abs, sizeof(float) * m);
    free(s->force_abs);
    s->force_abs = NULL;
    return 0;
}

ORACLE_API uint64_t oracle_sim_pairs_in_cutoff(const oracle_sim* s) { return s->pairs_in_cutoff; }
ORACLE_API uint32_t oracle_sim_n_rods(const oracle_sim* s) { return s->n; }
ORACLE_API uint32_t oracle_sim_n_vertices(const oracle_sim* s) { return s->nv; }
ORACLE_API uint64_t oracle_sim_iteration(const oracle_sim* s) { return s->iteration; }
ORACLE_API uint64_t oracle_sim_rod_steps(const oracle_sim* s) { return s->rod_steps; }

ORACLE_API int oracle_sim_get_state(oracle_sim* s, float* pos, float* vel, crm_rod_params* par,
                                    crm_ident* ids, crm_ident* parents) {
    size_t m = (size_t)s->n * s->nv * 3;
    if (pos) memcpy(pos, s->pos, sizeof(float) * m);
    if (vel) memcpy(vel, s->vel, sizeof(float) * m);
    if (par) memcpy(par, s->par, sizeof(crm_rod_params) * s->n);
    if (ids) memcpy(ids, s->id, sizeof(crm_ident) * s->n);
    if (parents) memcpy(parents, s->parent, sizeof(crm_ident) * s->n);
    return 0;
}

ORACLE_API uint32_t oracle_sim_n_snapshots(const oracle_sim* s) { return s->n_snaps; }

ORACLE_API int oracle_sim_snapshot_info(const oracle_sim* s, uint32_t k, uint64_t* it, uint32_t* n) {
    if (k >= s->n_snaps) return CRM_ERR_INVALID;
    if (it) *it = s->snaps[k].iteration;
    if (n) *n = s->snaps[k].n;
    return 0;
}

ORACLE_API int oracle_sim_snapshot_copy(const oracle_sim* s, uint32_t k, float* pos, float* vel,
                                        crm_rod_params* par, crm_ident* ids, crm_ident* parents) {
    if (k >= s->n_snaps) return CRM_ERR_INVALID;
    const snapshot* sn = &s->snaps[k];
    size_t m = (size_t)sn->n * s->nv * 3;
    if (pos) memcpy(pos, sn->pos, sizeof(float) * m);
    if (vel) memcpy(vel, sn->vel, sizeof(float) * m);
    if (par) memcpy(par, sn->par, sizeof(crm_rod_params) * sn->n);
    if (ids) memcpy(ids, sn->id, sizeof(crm_ident) * sn->n);
    if (parents) memcpy(parents, sn->parent, sizeof(crm_ident) * sn->n);
    return 0;
}

ORACLE_API int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

ORACLE_API void oracle_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* counter_to_color / color_to_counter: reference src/imaging.rs:28-66. */
ORACLE_API void oracle_counter_to_color(uint32_t counter, uint8_t rgb[3]) {
    uint64_t c = ((uint64_t)counter * 157ull * 163ull * 173ull) % (251ull * 251ull * 251ull);
    rgb[0] = (uint8_t)(c / (251ull * 251ull));
    c %= 251ull * 251ull;
    rgb[1] = (uint8_t)(c / 251ull);
    rgb[2] = (uint8_t)(c % 251ull);
}

ORACLE_API uint32_t oracle_color_to_counter(const uint8_t rgb[3]) {
    const uint64_t M = 251ull * 251ull * 251ull;
    uint64_t c = (uint64_t)rgb[0] * 251ull * 251ull + (uint64_t)rgb[1] * 251ull + rgb[2];
    c = (c * 12590168ull) % M;
    c = (c * 13775961ull) % M;
    c = (c * 12157008ull) % M;
    return (uint32_t)c;
}
