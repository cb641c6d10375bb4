// crm_integrate.cuh — RodMechanics::calculate_increment + Adams-Bashforth + Cycle::update_cycle + boundary for one
// rod, as a device function (see integrate_rod).  Reference: delegation src/crm_amir.rs:88-100, equations
// docs/source/model.rst:28-69, growth src/agent.rs:573-595.  Rust never contracts a*b+c and the bending force of a
// nearly straight rod is rounding noise that only reproduces op for op: include this only from translation units
// compiled with --fmad=false.
#pragma once
#include "crm_device.cuh"

namespace crm {

// ------------------------------------------------------------------------------------
// boundary reflection (CartesianSubDomainRods::apply_boundary)
__device__ __forceinline__ bool reflect(float& p, float& v, float hi) {
    if (p < 0.0f) { p = -p; v = fabsf(v); }
    if (p > hi) { p = 2.0f * hi - p; v = -fabsf(v); }
    return p >= 0.0f && p <= hi;
}

// What a rod's mechanics read of its own state: loaded from the arrays by k_integrate (load_mech_global), staged
// through shared memory by the fused list step (crm_force.cu: cp.async at kernel start, so that the round trip to
// HBM runs under the force loops instead of stalling the tail).
struct MechIn {
    float V[3], X1[3], V1[3], X2[3], V2[3];  // velocity, newer and older stored increment (dx, dv)
    MechParams mp;
    GrowState gs;
    float4 vb;  // vertex record at the last build of the neighbour structure (zero if there is none)
    int nh;     // stored increments so far
};

// increment history as a two-slot ring: `newx/newv` were written at step it-1, `oldx/oldv` at step it-2 (read as
// the older increment, then rewritten by this step)
struct HistPtrs { float *newx, *newv, *oldx, *oldv; };
__device__ __forceinline__ HistPtrs hist_ptrs(const Dev& d) {
    const bool odd = (d.iteration & 1ull) != 0;
    return HistPtrs{odd ? d.dx2 : d.dx1, odd ? d.dv2 : d.dv1, odd ? d.dx1 : d.dx2, odd ? d.dv1 : d.dv2};
}

template <int NV>
__device__ __forceinline__ void load_mech_global(const Dev& d, const uint32_t a, const int l, MechIn& m) {
    constexpr int LP = Lay<NV>::LP;
    constexpr int STRIDE = Lay<NV>::STRIDE;
    const size_t o = (size_t)a * STRIDE;
    const int li = l < NV ? l : 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) m.V[c] = d.vel[o + c * LP + li];
    m.mp = d.mech[a];
    m.gs = d.grow[a];
    const HistPtrs h = hist_ptrs(d);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const size_t i = o + c * LP + li;
        m.X1[c] = h.newx[i]; m.V1[c] = h.newv[i]; m.X2[c] = h.oldx[i]; m.V2[c] = h.oldv[i];
    }
    m.nh = d.nhist[a];
    m.vb = make_float4(0.f, 0.f, 0.f, 0.f);
    if (d.vbuild) m.vb = d.vbuild[(size_t)a * LP + li];
}

// fused mechanics: RodMechanics::calculate_increment + Adams-Bashforth(3) + Cycle::update_cycle
// + boundary + bounding sphere of the new state.
// Body shared by k_integrate (crm_mech.cu: P and F come from the pos / force arrays) and by the fused list step
// (crm_force.cu k_force_eval<.., FUSE>: P is the rod's own vertex record, F the force the kernel has just summed
// in registers).  One lane per vertex of rod `a`; every lane named in the warp-uniform mask `gm` calls it
// (shuffles run over gm), `valid` lanes store.  `skip_step`: a step of a batch behind the one that latched,
// tested behind the loads so that they are all in flight first.  `nn_count` < 0: the neighbour count is read from
// d.nn.  The new vertex record goes to `vrec_out` (the fused step must not overwrite records other blocks
// are still reading as neighbours).  `m`: the rod's own state (every load of the step is issued before the first
// store: the stores below may alias as far as the compiler knows, and one DRAM round trip per vector component is
// what bounds k_integrate).  Every TU that includes this is compiled with --fmad=false.
template <int NV>
__device__ __forceinline__ void integrate_rod(const Dev& d, const uint32_t a, const int l, const bool valid, const unsigned gm,
                                              const float (&P)[3], const float (&F)[3], const bool skip_step,
                                              const int nn_count, float4* __restrict__ vrec_out, const MechIn& m) {
    constexpr int LP = Lay<NV>::LP;
    constexpr int STRIDE = Lay<NV>::STRIDE;
    const size_t o = (size_t)a * STRIDE;
    const bool hv = valid && l < NV;
    const float (&V)[3] = m.V;
    const float (&X1)[3] = m.X1;
    const float (&V1)[3] = m.V1;
    const float (&X2)[3] = m.X2;
    const float (&V2)[3] = m.V2;
    const MechParams mp = m.mp;
    GrowState gs = m.gs;
    const HistPtrs hp = hist_ptrs(d);
    float* const oldx = hp.oldx;
    float* const oldv = hp.oldv;
    const int nh = m.nh;
    // the cold cycle record (48 B) is read only by rods whose growth depends on their neighbour count
    // (GrowState.w flags them); the setters in it matter at divisions only (k_divide)
    const bool has_nr = gs.w != 0.0f;
    float nr_max = 1.0f, nr_exp = 0.0f, nnb = 0.0f;
    if (has_nr) {
        const CycleParams cp = d.cyc[a];
        nr_max = (float)cp.nr_max;
        nr_exp = cp.nr_exp;
        nnb = nn_count >= 0 ? (float)nn_count : (float)d.nn[a];
    }
    const float4 vb = m.vb;
    const bool iso = d.sim_fail != nullptr;
    const uint32_t sim_a = iso ? d.tile[a] : 0u;
    const bool frozen = iso && d.sim_fail[sim_a] != 0;
    const float tension = mp.x, rigidity = mp.y, damping = mp.z, diffusion = mp.w;
    const float L = gs.z;
    if (skip_step) return;

    // next vertex (the connection to the previous one is the previous lane's connection to this one, below)
    float Pn[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) Pn[c] = __shfl_down_sync(gm, P[c], 1, LP);
    // spring of segment l (between l and l+1): force on vertex l
    float fs[3] = {0.f, 0.f, 0.f};
    float c2[3] = {P[0] - Pn[0], P[1] - Pn[1], P[2] - Pn[2]};  // conn_l = v_l - v_{l+1}
    float n2 = sqrtf(c2[0] * c2[0] + c2[1] * c2[1] + c2[2] * c2[2]);
    // unit connections: correctly rounded quotients from one reciprocal per length (div_rn)
    const float y2 = rcp_refined(n2);
    const float u2[3] = {div_rn(c2[0], n2, y2), div_rn(c2[1], n2, y2), div_rn(c2[2], n2, y2)};
    if (l < NV - 1 && n2 != 0.0f) {
        float k = (tension * (n2 - L)) * d.spring_w;
#pragma unroll
        for (int c = 0; c < 3; ++c) fs[c] = -u2[c] * k;
    }
    // bending at interior vertex l
    float fb[3] = {0.f, 0.f, 0.f};
    {
        // conn_{l-1} = v_{l-1} - v_l, its length and its unit vector are what lane l-1 has just computed as its
        // (c2, n2, u2): the same operations on the same operands, taken over instead of being repeated
        float c1[3], u1[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            c1[c] = __shfl_up_sync(gm, c2[c], 1, LP);
            u1[c] = __shfl_up_sync(gm, u2[c], 1, LP);
        }
        const float n1 = __shfl_up_sync(gm, n2, 1, LP);
        if (l >= 1 && l < NV - 1) {
            float angle;
            if (n1 == 0.0f || n2 == 0.0f) {
                angle = 0.0f;
            } else {
                const float nn = n1 * n2;
                float cs = div_rn(-(c1[0] * c2[0] + c1[1] * c2[1] + c1[2] * c2[2]), nn, rcp_refined(nn));
                cs = fminf(fmaxf(cs, -1.0f), 1.0f);
                angle = acosf(cs);
            }
            float fd[3] = {u1[0] - u2[0], u1[1] - u2[1], u1[2] - u2[2]};
            float nfd = sqrtf(fd[0] * fd[0] + fd[1] * fd[1] + fd[2] * fd[2]);
            float scale = rigidity * (3.14159265358979323846f - angle);
            const float yf = rcp_refined(nfd);
#pragma unroll
            for (int c = 0; c < 3; ++c) fb[c] = ((nfd == 0.0f) ? fd[c] : div_rn(fd[c], nfd, yf)) * scale;
        }
    }
    float dv[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float fs_prev = __shfl_up_sync(gm, fs[c], 1, LP);
        float fb_prev = __shfl_up_sync(gm, fb[c], 1, LP);
        float fb_next = __shfl_down_sync(gm, fb[c], 1, LP);
        if (l == 0) { fs_prev = 0.f; fb_prev = 0.f; }
        if (l >= NV - 1) fb_next = 0.f;
        float t = F[c];
        t -= fs_prev;  // reaction of segment l-1
        t += fs[c];
        t -= 0.5f * fb_prev;
        t += fb[c];
        t -= 0.5f * fb_next;
        t -= damping * V[c];
        dv[c] = t;
    }
    // Adams-Bashforth over the stored increments
    const float dt = d.dt;
    const float c23 = 23.0f / 12.0f, c16 = 16.0f / 12.0f, c5 = 5.0f / 12.0f;
    const float amp = (diffusion > 0.0f) ? sqrtf(2.0f * diffusion * dt) : 0.0f;
    float Pnew[3], Vnew[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float x1 = X1[c], v1 = V1[c], x2 = X2[c], v2 = V2[c];
        float ip, iv;
        if (nh >= 2) {
            ip = (c23 * V[c] - c16 * x1 + c5 * x2) * dt;
            iv = (c23 * dv[c] - c16 * v1 + c5 * v2) * dt;
        } else if (nh == 1) {
            ip = (1.5f * V[c] - 0.5f * x1) * dt;
            iv = (1.5f * dv[c] - 0.5f * v1) * dt;
        } else {
            ip = V[c] * dt;
            iv = dv[c] * dt;
        }
        if (amp != 0.0f) {
            float n0, n1;
            normal_pair(d.seed, d.iteration, d.rkey ? d.rkey[a] : a, (uint64_t)(l * 3 + c), n0, n1);
            ip += amp * n0;
        }
        Pnew[c] = P[c] + ip;
        Vnew[c] = V[c] + iv;
    }
    // Cycle::update_cycle (bit-exact ops: no FMA contraction, same order as the reference)
    float rate = gs.x;
    if (has_nr) {
        const float mx = nr_max;
        rate = __fmul_rn(gs.x, powf(fmaxf(__fdiv_rn(__fsub_rn(mx, nnb), mx), 0.0f), nr_exp));
    }
    const float newL = __fadd_rn(L, __fmul_rn(__fmul_rn(rate, dt), L));
    bool dividing = newL > gs.y && !frozen;

    bool inside = true;
    if (!dividing) {  // dividing rods are reflected after the division (k_divide)
        inside &= reflect(Pnew[0], Vnew[0], d.lx);
        inside &= reflect(Pnew[1], Vnew[1], d.ly);
        inside &= reflect(Pnew[2], Vnew[2], d.lz);
    }
    const bool finite = isfinite(Pnew[0]) && isfinite(Pnew[1]) && isfinite(Pnew[2]);
    bool keep = false;  // isolation: this rod keeps its old state
    if (iso) {
        const unsigned bad = __ballot_sync(gm, hv && (!finite || !inside)) & group_mask<LP>();
        if (bad && l == 0 && valid && !frozen) d.sim_fail[sim_a] = (bad && !finite) ? CRM_ERR_NUMERIC : CRM_ERR_BOUNDARY;
        keep = bad != 0 || frozen;
        if (keep) {
            dividing = false;
#pragma unroll
            for (int c = 0; c < 3; ++c) Pnew[c] = P[c];
        }
    }
    if (hv && !keep) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            d.pos[o + c * LP + l] = Pnew[c];
            d.vel[o + c * LP + l] = Vnew[c];
            oldx[o + c * LP + l] = V[c];
            oldv[o + c * LP + l] = dv[c];
        }
        if (!finite || !inside) {
            if (atomicCAS(&d.status->err_code, 0, finite ? CRM_ERR_BOUNDARY : CRM_ERR_NUMERIC) == 0)
                d.status->err_slot = a;
            if (d.latch) *d.latch = (unsigned)d.iteration;
        }
    }
    if (d.vbuild) {  // displacement since the persistent neighbour structure was built
        const float ex = Pnew[0] - vb.x, ey = Pnew[1] - vb.y, ez = Pnew[2] - vb.z;
        float e2 = ex * ex + ey * ey + ez * ez;
        if (vb.w > 0.0f) {  // a daughter: part of the budget was used up before its division
            const float e = sqrtf(e2) + vb.w;
            e2 = e * e;
        }
        const float dd = group_max<LP>(hv ? e2 : 0.0f, gm);
        if (l == 0 && valid) {
            atomicMax(&d.status->disp2_bits, __float_as_uint(dd));  // NaN compares high: forces a rebuild
            // the host's rule (step_conclude): the lists serve while every vertex stays within skin/2
            if (d.latch && !(sqrtf(dd) < 0.4999f * d.skin)) *d.latch = (unsigned)d.iteration;
        }
    }
    // The bounds of the search (sphere, chord, cull thresholds) are recomputed by k_bounds only when a
    // search follows (rebuild steps); every step needs just the vertex record of the force kernel.
    if (valid) vrec_out[(size_t)a * LP + l] = make_float4(Pnew[0], Pnew[1], Pnew[2], -1.0f);
    if (d.slab_lo > -INFINITY || d.slab_hi < INFINITY) {  // sharded: rods whose centroid left the slab
        const float cx = group_sum<LP>(hv ? Pnew[0] : 0.0f, gm) * (1.0f / NV);
        if (l == 0 && valid && !(cx >= d.slab_lo && cx < d.slab_hi)) atomicAdd(&d.status->n_leave, 1u);
    }
    if (__any_sync(gm, dividing)) {  // (warp-uniform: the shuffles below need every lane)
        // user voxel of the mother = voxel of the mean vertex at the start of the step,
        // summed sequentially like the reference (nalgebra row_mean)
        float mx = 0.f, my = 0.f;
#pragma unroll
        for (int q = 0; q < NV; ++q) {
            mx = __fadd_rn(mx, __shfl_sync(gm, P[0], q, LP));
            my = __fadd_rn(my, __shfl_sync(gm, P[1], q, LP));
        }
        if (l == 0 && valid && dividing) {
            mx = __fdiv_rn(mx, (float)NV);
            my = __fdiv_rn(my, (float)NV);
            int ix = (int)floorf(__fdiv_rn(mx, d.uvx_w)), iy = (int)floorf(__fdiv_rn(my, d.uvy_w));
            ix = min(max(ix, 0), d.unx - 1);
            iy = min(max(iy, 0), d.uny - 1);
            unsigned k = atomicAdd(&d.status->n_div, 1u);
            if (d.latch) *d.latch = (unsigned)d.iteration;
            if (k < d.divcap) {
                d.divlist[k].slot = a;
                d.divlist[k].uvox = (uint32_t)(ix * d.uny + iy);
            } else {
                d.status->overflow = 1;
            }
        }
    }
    if (l == 0 && valid && !keep) {
        gs.z = newL;
        d.grow[a] = gs;
        if (nh < d.ab_hist) d.nhist[a] = (uint8_t)(nh + 1);
    }
}

}  // namespace crm
