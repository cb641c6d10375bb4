// crm_mech.cu — binning, mechanics/integrator, division and layout kernels of the colony-mechanics
// step (compiled with --fmad=false: the reference's Rust arithmetic never contracts a*b+c, and the
// bending force of a nearly straight rod is rounding noise that only reproduces op for op).
//
//   k_bin_count / k_bin_scan / k_bin_scatter   centroid cell list (counting sort)
//   k_force<NV>         gather-style rod-rod interaction + neighbour sensing + domain force
//   k_integrate<NV>     springs + bending + damping, Adams-Bashforth, growth, boundary
//   k_divide<NV>        division geometry for the (host-ordered) list of dividing rods
//   k_pack / k_unpack   boundary layout <-> device layout
//
// Thread mapping everywhere: LP lanes per rod (lane = vertex), 32/LP rods per warp.
#include "crm_integrate.cuh"

namespace crm {


// ------------------------------------------------------------------------------------
// layout conversion: boundary [n][nv][3]  <->  device [n][3][LP]
template <int NV>
__global__ void k_unpack(const float* __restrict__ src, float* __restrict__ dst, uint32_t n) {
    constexpr int LP = Lay<NV>::LP;
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t rod = tid / (3 * LP);
    uint32_t rem = tid % (3 * LP);
    if (rod >= n) return;
    uint32_t c = rem / LP, i = rem % LP;
    dst[(size_t)rod * 3 * LP + rem] = (i < NV) ? src[((size_t)rod * NV + i) * 3 + c] : 0.0f;
}

template <int NV>
__global__ void k_pack(const float* __restrict__ src, float* __restrict__ dst, uint32_t n) {
    constexpr int LP = Lay<NV>::LP;
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t rod = tid / (3 * NV);
    uint32_t rem = tid % (3 * NV);
    if (rod >= n) return;
    uint32_t i = rem / 3, c = rem % 3;
    dst[(size_t)rod * 3 * NV + rem] = src[(size_t)rod * 3 * LP + c * LP + i];
}

// ------------------------------------------------------------------------------------
// bounding spheres of the current state + max extent (used after set_agents)
template <int NV>
__global__ void k_bounds(Dev d) {
    constexpr int LP = Lay<NV>::LP;
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a = tid / LP;
    int l = threadIdx.x % LP;
    unsigned m = group_mask<LP>();
    if (a >= d.n_own) return;  // whole groups exit together
    const float* p = d.pos + (size_t)a * 3 * LP;
    float x = p[l], y = p[LP + l], z = p[2 * LP + l];
    const float ext = store_bounds<NV>(d, a, x, y, z, l, m);
    if (l == 0) atomicMax(&d.status->emax_bits, __float_as_uint(ext));
}

// ------------------------------------------------------------------------------------
// cell list over rod centroids
__device__ __forceinline__ int cell_index(const Dev& d, float cx, float cy) {
    int ix = (int)floorf(cx * d.inv_wx), iy = (int)floorf(cy * d.inv_wy);
    ix = min(max(ix, 0), d.ncx - 1);
    iy = min(max(iy, 0), d.ncy - 1);
    return iy * d.ncx + ix;
}

__global__ void k_bin_count(Dev d) {
    uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= d.n) return;
    float4 bs = d.bsphere[a];
    int c = cell_index(d, bs.x, bs.y);
    if (d.tile) c += (int)(d.tile[a] * d.cpt);
    d.cell_of[a] = (uint32_t)c;
    d.cell_rank[a] = atomicAdd(&d.cell_count[c], 1u);
}

// single-block exclusive scan of cell_count -> cell_start (ncells + 1 entries); every thread owns
// ITEMS consecutive cells per round (vector loads), so 1e5 cells take a handful of rounds
__global__ void __launch_bounds__(1024) k_bin_scan(Dev d, uint32_t ncells) {
    constexpr int ITEMS = 16;
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry_s;
    const uint32_t T = blockDim.x;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < ncells; base += T * ITEMS) {
        const uint32_t i0 = base + threadIdx.x * ITEMS;
        uint32_t v[ITEMS];
        if (i0 + ITEMS <= ncells) {
            const uint4* src = reinterpret_cast<const uint4*>(d.cell_count + i0);
#pragma unroll
            for (int q = 0; q < ITEMS / 4; ++q) {
                const uint4 w = src[q];
                v[4 * q] = w.x; v[4 * q + 1] = w.y; v[4 * q + 2] = w.z; v[4 * q + 3] = w.w;
            }
        } else {
#pragma unroll
            for (int q = 0; q < ITEMS; ++q) v[q] = (i0 + q < ncells) ? d.cell_count[i0 + q] : 0u;
        }
        uint32_t tot = 0;
#pragma unroll
        for (int q = 0; q < ITEMS; ++q) {
            const uint32_t t = v[q];
            v[q] = tot;  // exclusive within the thread
            tot += t;
        }
        uint32_t s = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(FULL, s, o);
            if ((threadIdx.x & 31) >= (unsigned)o) s += t;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = (threadIdx.x < (T >> 5)) ? warp_tot[threadIdx.x] : 0u;
            uint32_t ws = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(FULL, ws, o);
                if (threadIdx.x >= (unsigned)o) ws += t;
            }
            warp_tot[threadIdx.x] = ws - w;  // exclusive
        }
        __syncthreads();
        const uint32_t excl = carry_s + warp_tot[threadIdx.x >> 5] + s - tot;
        if (i0 + ITEMS <= ncells) {
            uint4* dst = reinterpret_cast<uint4*>(d.cell_start + i0);
#pragma unroll
            for (int q = 0; q < ITEMS / 4; ++q)
                dst[q] = make_uint4(excl + v[4 * q], excl + v[4 * q + 1], excl + v[4 * q + 2], excl + v[4 * q + 3]);
        } else {
#pragma unroll
            for (int q = 0; q < ITEMS; ++q)
                if (i0 + q < ncells) d.cell_start[i0 + q] = excl + v[q];
        }
        __syncthreads();
        if (threadIdx.x == T - 1) carry_s = excl + tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) d.cell_start[ncells] = carry_s;
}

__global__ void k_bin_scatter(Dev d) {
    uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= d.n) return;
    d.sorted[d.cell_start[d.cell_of[a]] + d.cell_rank[a]] = a;
}

// The atomic ranks are arrival-ordered; sorting each cell's entries by slot makes
// the candidate order (and with it the f32 summation order) reproducible.
__global__ void k_bin_sort(Dev d, uint32_t ncells) {
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    uint32_t s = d.cell_start[c], e = d.cell_start[c + 1];
    for (uint32_t i = s + 1; i < e; ++i) {
        uint32_t v = d.sorted[i];
        uint32_t j = i;
        while (j > s && d.sorted[j - 1] > v) {
            d.sorted[j] = d.sorted[j - 1];
            --j;
        }
        d.sorted[j] = v;
    }
    for (uint32_t i = s; i < e; ++i) {  // cell-sorted copies of the bounds for the force kernel's candidate scan
        const uint32_t v = d.sorted[i];
        d.gof[v] = i;
        d.bs_sorted[i] = d.bsphere[v];
        d.chord_sorted[2 * i] = d.chord[2 * v];
        d.chord_sorted[2 * i + 1] = d.chord[2 * v + 1];
    }
}


// ------------------------------------------------------------------------------------
// fused mechanics: RodMechanics::calculate_increment + Adams-Bashforth(3) + Cycle::update_cycle
// + boundary + vertex record of the new state (body: crm_integrate.cuh).
template <int NV>
__global__ void __launch_bounds__(128) k_integrate(Dev d) {
    constexpr int LP = Lay<NV>::LP;
    constexpr int STRIDE = Lay<NV>::STRIDE;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int l = threadIdx.x % LP;
    // the whole warp stays converged (full-mask shuffles: a per-group mask costs a uniform-register
    // waterfall per shuffle); groups past the last rod recompute it and store nothing
    const bool valid = tid / LP < d.n_own;
    const uint32_t a = valid ? tid / LP : d.n_own - 1;
    const size_t o = (size_t)a * STRIDE;
    const int li = l < NV ? l : 0;
    // (batches: the latch is loaded with the first wave of loads and tested behind them; a step of a batch behind
    // the one that latched returns, blocks of the latching step itself read their own iteration or 0 and go on.
    // d.skip: a search step whose cell grid turned out too fine for the rods, the host repeats it)
    const unsigned stop = d.latch ? *d.latch : 0u;
    const unsigned skip = d.skip ? *d.skip : 0u;
    float P[3], F[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        P[c] = d.pos[o + c * LP + li];
        F[c] = d.force[o + c * LP + li];
    }
    MechIn m;
    load_mech_global<NV>(d, a, l, m);
    integrate_rod<NV>(d, a, l, valid, 0xffffffffu, P, F, (stop != 0u && stop < (unsigned)d.iteration) || skip != 0u, -1, d.vrec, m);
}

// ------------------------------------------------------------------------------------
// division: RodMechanics::divide + Cycle::divide for the slot-ordered list of mothers.
// Group r handles mother divlist[r]; daughter 1 stays in the mother's slot, daughter 2
// goes to slot n0 + r.
__device__ __forceinline__ float setter_sample(int kind, float a, float b, int which, uint64_t seed,
                                               uint64_t it, uint64_t slot, uint64_t stream) {
    if (kind == CRM_SETTER_EXPLICIT) return which == 0 ? a : b;
    if (b == 0.0f) return a;
    float n0, n1;
    normal_pair_precise(seed ^ 0xD1B54A32D192ED03ULL, it, slot, stream, n0, n1);
    return __fadd_rn(a, __fmul_rn(b, which == 0 ? n0 : n1));
}

template <int NV>
__global__ void __launch_bounds__(128) k_divide(Dev d, uint32_t n_div, uint32_t n0) {
    constexpr int LP = Lay<NV>::LP;
    constexpr int STRIDE = Lay<NV>::STRIDE;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t r = tid / LP;
    const int l = threadIdx.x % LP;
    const unsigned gm = group_mask<LP>();
    if (r >= n_div) return;
    const uint32_t a = d.divlist[r].slot;
    const uint32_t b = n0 + r;
    const size_t oa = (size_t)a * STRIDE, ob = (size_t)b * STRIDE;
    const bool hv = l < NV;
    const int li = hv ? l : 0;
    float P[3] = {d.pos[oa + li], d.pos[oa + LP + li], d.pos[oa + 2 * LP + li]};
    float V[3] = {d.vel[oa + li], d.vel[oa + LP + li], d.vel[oa + 2 * LP + li]};
    const PotParams pot = d.pot[a];
    GrowState gs = d.grow[a];
    const CycleParams cp = d.cyc[a];
    // all vertices and segment lengths, replicated in every lane
    float X[NV], Y[NV], Z[NV], seg[NV], cum[NV + 1];
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        X[q] = __shfl_sync(gm, P[0], q, LP);
        Y[q] = __shfl_sync(gm, P[1], q, LP);
        Z[q] = __shfl_sync(gm, P[2], q, LP);
    }
    float total = 0.0f;
    cum[0] = 0.0f;
#pragma unroll
    for (int q = 0; q < NV - 1; ++q) {
        float dx = X[q] - X[q + 1], dy = Y[q] - Y[q + 1], dz = Z[q] - Z[q + 1];
        seg[q] = sqrtf(dx * dx + dy * dy + dz * dz);
        total += seg[q];
        cum[q + 1] = total;
    }
    const float seg_len = (total - 2.0f * pot.radius) / (float)(NV - 1) / 2.0f;
    float D1[3] = {0.f, 0.f, 0.f}, D2[3] = {0.f, 0.f, 0.f};
    {
        // lane l computes vertex l of daughter 1 and vertex NV-1-n (n = NV-1-l) of daughter 2
        const float s1 = (float)li * seg_len;
        int k = 0;
#pragma unroll
        for (int j = 0; j < NV - 1; ++j)
            if (cum[j] <= s1) k = j;
        float ck = 0.f, sk = 1.f, x0 = 0.f, y0 = 0.f, z0 = 0.f, x1 = 0.f, y1 = 0.f, z1 = 0.f;
#pragma unroll
        for (int j = 0; j < NV - 1; ++j)
            if (j == k) { ck = cum[j]; sk = seg[j]; x0 = X[j]; y0 = Y[j]; z0 = Z[j]; x1 = X[j + 1]; y1 = Y[j + 1]; z1 = Z[j + 1]; }
        float q = (s1 - ck) / sk;
        D1[0] = x0 * (1.0f - q) + x1 * q;
        D1[1] = y0 * (1.0f - q) + y1 * q;
        D1[2] = z0 * (1.0f - q) + z1 * q;
        // daughter 2: vertex index l corresponds to n = NV-1-l steps from the end
        const float s2 = (float)(NV - 1 - li) * seg_len;
        float cr = 0.0f, cm = 0.0f;
        int m = 0;
#pragma unroll
        for (int j = 0; j < NV - 1; ++j) {
            if (cr <= s2) { m = j; cm = cr; }
            cr += seg[NV - 2 - j];
        }
        float sm = 1.f;
#pragma unroll
        for (int j = 0; j < NV - 1; ++j)
            if (j == m) { sm = seg[NV - 2 - j]; x0 = X[NV - 1 - j]; y0 = Y[NV - 1 - j]; z0 = Z[NV - 1 - j]; x1 = X[NV - 2 - j]; y1 = Y[NV - 2 - j]; z1 = Z[NV - 2 - j]; }
        float p = (s2 - cm) / sm;
        D2[0] = x0 * (1.0f - p) + x1 * p;
        D2[1] = y0 * (1.0f - p) + y1 * p;
        D2[2] = z0 * (1.0f - p) + z1 * p;
    }
    float V1[3] = {V[0], V[1], V[2]}, V2[3] = {V[0], V[1], V[2]};
    const float U1[3] = {D1[0], D1[1], D1[2]}, U2[3] = {D2[0], D2[1], D2[2]};  // before the reflection
    bool inside = true;
    inside &= reflect(D1[0], V1[0], d.lx); inside &= reflect(D1[1], V1[1], d.ly); inside &= reflect(D1[2], V1[2], d.lz);
    inside &= reflect(D2[0], V2[0], d.lx); inside &= reflect(D2[1], V2[1], d.ly); inside &= reflect(D2[2], V2[2], d.lz);
    // Budget of the persistent neighbour structure the daughters start with: they lie on the mother's
    // polyline, so whatever is out of the mother's reach is out of theirs, up to (i) the mother's own
    // displacement since the structure was built, (ii) the corner a daughter segment cuts where it spans a
    // mother vertex, (iii) a reflection at the boundary.  The sum goes into vbuild.w (k_integrate adds it).
    float slack = 0.0f;
    if (d.vbuild) {
        const float4 vbm = d.vbuild[(size_t)a * LP + li];
        const float mx = P[0] - vbm.x, my = P[1] - vbm.y, mz = P[2] - vbm.z;
        float far = hv ? sqrtf(mx * mx + my * my + mz * mz) : 0.0f;
        const float r1x = D1[0] - U1[0], r1y = D1[1] - U1[1], r1z = D1[2] - U1[2];
        const float r2x = D2[0] - U2[0], r2y = D2[1] - U2[1], r2z = D2[2] - U2[2];
        float refl = hv ? fmaxf(sqrtf(r1x * r1x + r1y * r1y + r1z * r1z), sqrtf(r2x * r2x + r2y * r2y + r2z * r2z)) : 0.0f;
        far = group_max<LP>(far, gm);
        refl = group_max<LP>(refl, gm);
        float sag = 0.0f;
#pragma unroll
        for (int q = 1; q < NV - 1; ++q) {  // distance of mother vertex q from the chord of its neighbours
            const float cx = X[q + 1] - X[q - 1], cy = Y[q + 1] - Y[q - 1], cz = Z[q + 1] - Z[q - 1];
            const float wx = X[q] - X[q - 1], wy = Y[q] - Y[q - 1], wz = Z[q] - Z[q - 1];
            const float kx = wy * cz - wz * cy, ky = wz * cx - wx * cz, kz = wx * cy - wy * cx;
            const float c2 = cx * cx + cy * cy + cz * cz;
            sag = fmaxf(sag, c2 > 0.0f ? sqrtf((kx * kx + ky * ky + kz * kz) / c2) : sqrtf(wx * wx + wy * wy + wz * wz));
        }
        slack = (far + fmaxf(vbm.w, 0.0f) + sag + refl) * 1.0001f + 1e-6f;
    }
    const bool finite = isfinite(D1[0]) && isfinite(D1[1]) && isfinite(D1[2]) && isfinite(D2[0]) && isfinite(D2[1]) && isfinite(D2[2]);
    if (hv) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            d.pos[oa + c * LP + l] = D1[c];
            d.vel[oa + c * LP + l] = V1[c];
            d.pos[ob + c * LP + l] = D2[c];
            d.vel[ob + c * LP + l] = V2[c];
        }
        if (!finite || !inside) {
            if (d.sim_fail) {  // isolation: flag the sim, park the daughters inside the domain
                d.sim_fail[d.tile[a]] = finite ? CRM_ERR_BOUNDARY : CRM_ERR_NUMERIC;
                const float hi[3] = {d.lx, d.ly, d.lz};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    D1[c] = fminf(fmaxf(isfinite(D1[c]) ? D1[c] : 0.5f * hi[c], 0.0f), hi[c]);
                    D2[c] = fminf(fmaxf(isfinite(D2[c]) ? D2[c] : 0.5f * hi[c], 0.0f), hi[c]);
                    d.pos[oa + c * LP + l] = D1[c];
                    d.pos[ob + c * LP + l] = D2[c];
                }
            } else if (atomicCAS(&d.status->err_code, 0, finite ? CRM_ERR_BOUNDARY : CRM_ERR_NUMERIC) == 0) {
                d.status->err_slot = a;
            }
        }
    }
    if (d.vbuild) {
        d.vbuild[(size_t)a * LP + l] = make_float4(D1[0], D1[1], D1[2], slack);
        d.vbuild[(size_t)b * LP + l] = make_float4(D2[0], D2[1], D2[2], slack);
    }
    // the new cell starts with an empty increment history
    {
        const int lz = threadIdx.x % LP;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            d.dx1[ob + c * LP + lz] = 0.f; d.dv1[ob + c * LP + lz] = 0.f;
            d.dx2[ob + c * LP + lz] = 0.f; d.dv2[ob + c * LP + lz] = 0.f;
            d.force[ob + c * LP + lz] = 0.f;
        }
    }
    const float ext1 = store_bounds<NV>(d, a, D1[0], D1[1], D1[2], l, gm);
    const float ext2 = store_bounds<NV>(d, b, D2[0], D2[1], D2[2], l, gm);
    if (l == 0) {
        // spring_length *= 1/2 - radius/(n_vertices*spring_length)   (bit-exact op order)
        const float L = gs.z;
        const float div_factor = __fsub_rn(0.5f, __fdiv_rn(pot.radius, __fmul_rn((float)NV, L)));
        const float newL = __fmul_rn(div_factor, L);
        const uint64_t key = d.rkey ? d.rkey[a] : a;  // the mother's global slot
        const float g1 = setter_sample(cp.growth_kind, cp.growth_a, cp.growth_b, 0, d.seed, d.iteration, key, 0);
        const float g2 = setter_sample(cp.growth_kind, cp.growth_a, cp.growth_b, 1, d.seed, d.iteration, key, 0);
        const float l1 = setter_sample(cp.thr_kind, cp.thr_a, cp.thr_b, 0, d.seed, d.iteration, key, 1);
        const float l2 = setter_sample(cp.thr_kind, cp.thr_a, cp.thr_b, 1, d.seed, d.iteration, key, 1);
        d.grow[a] = make_float4(g1, l1, newL, gs.w);
        d.grow[b] = make_float4(g2, l2, newL, gs.w);
        d.pot[b] = pot;
        d.mech[b] = d.mech[a];
        d.cyc[b] = cp;
        d.nn[b] = d.nn[a];
        d.nhist[b] = 0;
        if (d.tile) d.tile[b] = d.tile[a];
        atomicMax(&d.status->emax_bits, __float_as_uint(fmaxf(ext1, ext2)));
    }
}

// ------------------------------------------------------------------------------------
// launchers (dispatch on n_vertices)


bool nv_supported(int nv) { return nv >= 2 && nv <= 16; }
int lanes_per_rod(int nv) { return pow2ceil(nv); }

void launch_unpack(int nv, const float* src, float* dst, uint32_t n, cudaStream_t st) {
    if (n == 0) return;
#define X(NVV) if (nv == NVV) k_unpack<NVV><<<blocks_for((uint64_t)n * 3 * Lay<NVV>::LP, 256), 256, 0, st>>>(src, dst, n);
    CRM_NV_LIST(X)
#undef X
}
void launch_pack(int nv, const float* src, float* dst, uint32_t n, cudaStream_t st) {
    if (n == 0) return;
#define X(NVV) if (nv == NVV) k_pack<NVV><<<blocks_for((uint64_t)n * 3 * NVV, 256), 256, 0, st>>>(src, dst, n);
    CRM_NV_LIST(X)
#undef X
}
void launch_bounds_kernel(int nv, const Dev& d, cudaStream_t st) {
    if (d.n_own == 0) return;
#define X(NVV) if (nv == NVV) k_bounds<NVV><<<blocks_for((uint64_t)d.n_own * Lay<NVV>::LP, 128), 128, 0, st>>>(d);
    CRM_NV_LIST(X)
#undef X
}
int launch_binning(const Dev& d, uint32_t n_tiles, cudaStream_t st) {
    if (d.n == 0) return 0;
    uint32_t ncells = (uint32_t)d.ncx * (uint32_t)d.ncy * n_tiles;
    cudaMemsetAsync(d.cell_count, 0, sizeof(uint32_t) * ncells, st);
    k_bin_count<<<blocks_for(d.n, 256), 256, 0, st>>>(d);
    k_bin_scan<<<1, 1024, 0, st>>>(d, ncells);
    k_bin_scatter<<<blocks_for(d.n, 256), 256, 0, st>>>(d);
    k_bin_sort<<<blocks_for(ncells, 128), 128, 0, st>>>(d, ncells);
    return 4;
}
void launch_integrate(int nv, const Dev& d, cudaStream_t st) {
    if (d.n_own == 0) return;
    // Occupancy cap: measured on B200 (1e6 rods, 8 vertices) the kernel runs 0.297 ms at <= 26 resident
    // warps per SM and 0.37-0.40 ms at >= 28 (its ~30 loads per warp are all issued up front; beyond that
    // point the memory system thrashes).  A dynamic shared-memory reservation nobody touches keeps
    // six 128-thread blocks = 24 warps per SM (36000 B: below the 48 KB that needs no opt-in).
    constexpr int CAP_SMEM = 36000;
#define X(NVV) \
    if (nv == NVV) k_integrate<NVV><<<blocks_for((uint64_t)d.n_own * Lay<NVV>::LP, 128), 128, CAP_SMEM, st>>>(d);
    CRM_NV_LIST(X)
#undef X
}
// ------------------------------------------------------------------------------------
// Divisions while the persistent neighbour structure is live.  Phase 1 (one thread per division): the new
// daughter gets a row of its own, a copy of the mother's list; the two daughters become each other's extra
// candidates.  Phase 2 (one thread per (division, listed neighbour)): every rod that lists the mother gets
// the new daughter as an extra candidate, and so does the daughter such a rod split off in the same step.
// All those rows are queued in dirty_list for k_force<.., LISTS>.
__device__ __forceinline__ void patch_add(const Dev& d, uint32_t g, uint32_t slot, int lp, int maxp) {
    const uint32_t p = atomicAdd(&d.dcnt[g], 1u);
    if (p < PATCH_EXTRA) d.dextra[(size_t)g * PATCH_EXTRA + p] = slot;
    // the rebuilt row keeps every candidate: it has to fit (the list step after this one reads it)
    if (p >= PATCH_EXTRA || (d.pnq[(size_t)g * lp] >> 24) + p + 1 > (uint32_t)maxp) d.status->patch_overflow = 1u;
    if (atomicExch(&d.dmark[g], 1u) == 0u) d.dirty_list[atomicAdd(d.dirty_n, 1u)] = g;
}
__global__ void k_patch_rows(Dev d, uint32_t n_div, uint32_t n0, int lp, int maxp) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_div) return;
    const uint32_t a = d.divlist[r].slot, b = n0 + r;
    const uint32_t ga = d.gof[a], gb = n0 + r;  // rows and slots are both appended in division order
    d.sorted[gb] = b;
    d.gof[b] = gb;
    const uint32_t cnt = d.pnq[(size_t)ga * lp] >> 24;
    for (uint32_t i = 0; i < cnt; ++i) d.plist[(size_t)gb * maxp + i] = d.plist[(size_t)ga * maxp + i];
    for (int l = 0; l < lp; ++l) d.pnq[(size_t)gb * lp + l] = cnt << 24;
    d.dsib[ga] = gb;
    patch_add(d, ga, b, lp, maxp);
    patch_add(d, gb, a, lp, maxp);
}
__global__ void k_patch_neighbours(Dev d, uint32_t n_div, uint32_t n0, int lp, int maxp) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t r = t / maxp, i = t % maxp;
    if (r >= n_div) return;
    const uint32_t a = d.divlist[r].slot, b = n0 + r;
    const uint32_t ga = d.gof[a];
    if (i >= (d.pnq[(size_t)ga * lp] >> 24)) return;
    const uint32_t gc = d.gof[d.plist[(size_t)ga * maxp + i]];
    patch_add(d, gc, b, lp, maxp);
    const uint32_t gs = d.dsib[gc];
    if (gs != 0xffffffffu) patch_add(d, gs, b, lp, maxp);
}
int launch_patch(const Dev& d, uint32_t n_div, uint32_t n0, int lp, int maxp, cudaStream_t st) {
    if (n_div == 0) return 0;
    k_patch_rows<<<blocks_for(n_div, 128), 128, 0, st>>>(d, n_div, n0, lp, maxp);
    k_patch_neighbours<<<blocks_for((uint64_t)n_div * maxp, 128), 128, 0, st>>>(d, n_div, n0, lp, maxp);
    return 2;
}

void launch_divide(int nv, const Dev& d, uint32_t n_div, uint32_t n0, cudaStream_t st) {
    if (n_div == 0) return;
#define X(NVV) if (nv == NVV) k_divide<NVV><<<blocks_for((uint64_t)n_div * Lay<NVV>::LP, 128), 128, 0, st>>>(d, n_div, n0);
    CRM_NV_LIST(X)
#undef X
}

}  // namespace crm
