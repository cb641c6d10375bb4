// crm_raster.cu — 2-D mask path as CUDA kernels: capsule-chain outline of every rod ->
// instance-colour bitmap + overlap bitmap, summed outline intersection area, per-pixel mask difference.
// Replaces, behind the C ABI, the reference's CPU code (`plotters` bitmap backend + `geo` boolean ops):
//   calculate_polygon_hull   reference src/fitting.rs:514-620 (+ helpers :347-500)
//   render_mask_2d           reference src/fitting.rs:121-244
//   overlap                  reference src/fitting.rs:299-345
//   parents_diff_mask        reference src/fitting.rs:42-109
// No graphics API: rasterisation is integer scanline arithmetic.  All kernels are HBM / latency bound
// (bitmaps: 8 B of atomics + 6 B of output per pixel); nothing here is a contraction.
// Compiled with --fmad=false: the outline arithmetic follows the reference's f32 operations one by one
// and the trigonometric functions go through f64 (cos_rn): outlines equal the CPU oracle's bit for bit.
#include <mutex>

#include "crm_sim.h"

using crm::blocks_for;

namespace {

constexpr int MAXV = 256;  // outline points per rod (8 vertices at delta_angle = pi/4 give ~40)
constexpr float PI_F = 3.14159265358979323846f;

struct P2 { float x, y; };

// cos / sin / atan2 through f64, rounded to f32: the same f32 as the oracle's (glibc) on every platform, which the
// f32 routines of libm and libdevice are not (decision R0 of oracle/raster_oracle.c).  One thread per rod, a few
// dozen calls: the f64 rate does not matter here.
__device__ __forceinline__ float cos_rn(float x) { return (float)cos((double)x); }
__device__ __forceinline__ float sin_rn(float x) { return (float)sin((double)x); }
__device__ __forceinline__ float atan2_rn(float y, float x) { return (float)atan2((double)y, (double)x); }

// get_n_resolution, src/fitting.rs:403-414
__device__ int n_resolution(float a0, float a1, float delta) {
    const float dangle = fabsf(a0 - a1);
    const float n = fmaxf(dangle / delta, 1.0f);
    const float frac = fmodf(n, 1.0f);
    int out = (fabsf(frac) <= 0.001f) ? (int)(n - 2.0f) : (int)(n - 1.0f);
    if (out < 0) out = 0;  // `as usize` saturates
    return out < 1 ? 1 : out;
}

// writes the points through `put(k, point)`; generate_coordinates_sphere, src/fitting.rs:348-372
template <class F>
__device__ int arc_points(P2 p, float r, float a0, float a1, float delta, F put) {
    if (a0 == a1) return 0;
    const float end = (a0 < a1) ? a1 : a1 + 2.0f * PI_F;
    const int n = n_resolution(a0, end, delta);
    const float dangle = fabsf(end - a0) / ((float)n + 1.0f);
    for (int k = 1; k <= n; ++k) {
        const float ang = a0 + dangle * (float)k;
        put(k - 1, P2{p.x + cos_rn(ang) * r, p.y + sin_rn(ang) * r});
    }
    return n;
}

// generate_coordinates_rectangle, src/fitting.rs:374-396
__device__ void rect_points(P2 p1, P2 p2, float r, P2 c[4]) {
    const float dx = p2.x - p1.x, dy = p2.y - p1.y;
    const float norm = sqrtf(dx * dx + dy * dy);
    const float d0 = dx / norm, d1 = dy / norm;
    c[0] = P2{p1.x - r * d1, p1.y + r * d0};
    c[1] = P2{p2.x - r * d1, p2.y + r * d0};
    c[2] = P2{p2.x + r * d1, p2.y - r * d0};
    c[3] = P2{p1.x + r * d1, p1.y - r * d0};
}

__device__ float angle_x(float dx, float dy) {  // atan2(y, x).rem_euclid(2 pi), src/fitting.rs:398-400
    const float a = atan2_rn(dy, dx);
    float m = fmodf(a, 2.0f * PI_F);
    if (m < 0.0f) m += 2.0f * PI_F;
    return m;
}

// geo::line_intersection restated (single point: proper crossing or touching endpoint)
__device__ bool seg_intersection(P2 a, P2 b, P2 c, P2 d, P2& out) {
    const float rx = b.x - a.x, ry = b.y - a.y, sx = d.x - c.x, sy = d.y - c.y;
    const float den = rx * sy - ry * sx;
    const float qx = c.x - a.x, qy = c.y - a.y;
    if (den == 0.0f) {
        if (qx * ry - qy * rx != 0.0f) return false;
        if (b.x == c.x && b.y == c.y) { out = b; return true; }
        if (a.x == d.x && a.y == d.y) { out = a; return true; }
        return false;
    }
    const float t = (qx * sy - qy * sx) / den;
    const float u = (qx * ry - qy * rx) / den;
    if (t < 0.0f || t > 1.0f || u < 0.0f || u > 1.0f) return false;
    out = P2{a.x + t * rx, a.y + t * ry};
    return true;
}

// plotters RangedCoordf32::map with limit (0, res): floor(res * (v - 0)/(size - 0) + 1e-3), f64
__device__ int to_pixel(float v, float size, int res) {
    const double logic = ((double)v - 0.0) / ((double)size - 0.0);
    return (int)floor((double)res * logic + 1e-3);
}

// One thread per rod: outline (calculate_polygon_hull) into hull[rod][MAXV], count, orientation, float and
// pixel bounding boxes, pixel-snapped vertices.  The forward list grows from the front of the row, the
// backward list from its end (stored reversed: that IS the order the polygon closes in).
__global__ void k_hull(const float* __restrict__ pos, const float* __restrict__ radii, int n, int nv, float delta, float eps,
                       float size_x, float size_y, int res_x, int res_y, float2* __restrict__ hull, int2* __restrict__ hull_px,
                       int* __restrict__ count, float4* __restrict__ bbox, int4* __restrict__ bbox_px, float* __restrict__ orient,
                       int* __restrict__ status) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float2* const H = hull + (size_t)i * MAXV;
    const float r = radii[i];
    auto P = [&](int k) { return P2{pos[((size_t)i * nv + k) * 3], pos[((size_t)i * nv + k) * 3 + 1]}; };
    int nf = 0, nb = 0;
    bool bad = false, full = false;
    auto push_f = [&](P2 q) { if (nf + nb < MAXV) H[nf++] = make_float2(q.x, q.y); else full = true; };
    auto push_b = [&](P2 q) { if (nf + nb < MAXV) { H[MAXV - 1 - nb] = make_float2(q.x, q.y); ++nb; } else full = true; };
    // get_coordinates_at_tip, src/fitting.rs:474-500
    auto tip = [&](P2 p1, P2 p2, P2& c0, P2& c3, float& a0, float& a1) {
        const float dx = p2.x - p1.x, dy = p2.y - p1.y;
        const float norm = sqrtf(dx * dx + dy * dy);
        const float d0 = dx / norm, d1 = dy / norm;
        a0 = angle_x(-d1, d0);
        a1 = angle_x(d1, -d0);
        P2 c[4];
        rect_points(p1, p2, r, c);
        c0 = c[0];
        c3 = c[3];
    };
    {
        P2 c0, c3;
        float a0, a1;
        tip(P(0), P(1), c0, c3, a0, a1);
        arc_points(P(0), r, a0, a1, delta, [&](int, P2 q) { push_b(q); });
        push_b(c3);
        push_f(c0);
    }
    for (int m = 2; m < nv && !bad; ++m) {
        const P2 p1 = P(m - 2), p2 = P(m - 1), p3 = P(m);
        P2 c[4], d[4];
        rect_points(p1, p2, r, c);
        rect_points(p2, p3, r, d);
        // determine_spheroid_coordinates_between_rectangles, src/fitting.rs:416-466
        const float dir1x = p2.x - p1.x, dir1y = p2.y - p1.y, dir2x = p3.x - p2.x, dir2y = p3.y - p2.y;
        const float ang = atan2_rn(dir1x * dir2y - dir1y * dir2x, dir1x * dir2x + dir1y * dir2y);
        const bool right = ang < 0.0f;
        float a0, a1;
        if (right) { a0 = angle_x(d[0].x - p2.x, d[0].y - p2.y); a1 = angle_x(c[1].x - p2.x, c[1].y - p2.y); }
        else { a0 = angle_x(c[2].x - p2.x, c[2].y - p2.y); a1 = angle_x(d[3].x - p2.x, d[3].y - p2.y); }
        P2 x;
        if (right) {
            push_f(c[1]);
            // the arc is appended reversed: reserve its slots, fill them back to front
            const int base = nf;
            const int na = arc_points(p2, r, a0, a1, delta, [&](int, P2) {});
            if (nf + nb + na < MAXV) {
                arc_points(p2, r, a0, a1, delta, [&](int k, P2 q) { H[base + (na - 1 - k)] = make_float2(q.x, q.y); });
                nf += na;
            } else {
                full = true;
            }
            push_f(d[0]);
            if (seg_intersection(c[3], c[2], d[3], d[2], x)) push_b(x);
            else if (fabsf(c[2].x - d[3].x) <= eps && fabsf(c[2].y - d[3].y) <= eps) push_b(c[2]);
            else bad = true;
        } else {
            push_b(c[2]);
            arc_points(p2, r, a0, a1, delta, [&](int, P2 q) { push_b(q); });
            push_b(d[3]);
            if (seg_intersection(c[0], c[1], d[0], d[1], x)) push_f(x);
            else if (fabsf(c[1].x - d[0].x) <= eps && fabsf(c[1].y - d[0].y) <= eps) push_f(c[1]);
            else bad = true;
        }
    }
    if (!bad) {
        P2 c0, c3;
        float a0, a1;
        tip(P(nv - 1), P(nv - 2), c0, c3, a0, a1);
        push_b(c0);
        // forward list += reversed(arc ++ [c3]) = c3, then the arc back to front
        push_f(c3);
        const int base = nf;
        const int na = arc_points(P(nv - 1), r, a0, a1, delta, [&](int, P2) {});
        if (nf + nb + na < MAXV) {
            arc_points(P(nv - 1), r, a0, a1, delta, [&](int k, P2 q) { H[base + (na - 1 - k)] = make_float2(q.x, q.y); });
            nf += na;
        } else {
            full = true;
        }
    }
    if (bad || full) {
        atomicCAS(status, 0, bad ? -(1 + i) : -(0x40000000 + i));
        count[i] = 0;
        return;
    }
    for (int k = 0; k < nb; ++k) H[nf + k] = H[MAXV - nb + k];  // close the polygon: forward ++ reversed(backward)
    const int m = nf + nb;
    count[i] = m;
    float xmin = 3e38f, ymin = 3e38f, xmax = -3e38f, ymax = -3e38f;
    int pxmin = 1 << 30, pymin = 1 << 30, pxmax = -(1 << 30), pymax = -(1 << 30);
    double area2 = 0.0;
    for (int k = 0; k < m; ++k) {
        const float2 a = H[k], b = H[k + 1 == m ? 0 : k + 1];
        area2 += (double)(a.x - H[0].x) * (double)(b.y - H[0].y) - (double)(b.x - H[0].x) * (double)(a.y - H[0].y);
        xmin = fminf(xmin, a.x); xmax = fmaxf(xmax, a.x);
        ymin = fminf(ymin, a.y); ymax = fmaxf(ymax, a.y);
        if (hull_px) {
            const int px = to_pixel(a.x, size_x, res_x), py = to_pixel(a.y, size_y, res_y);
            hull_px[(size_t)i * MAXV + k] = make_int2(px, py);
            pxmin = min(pxmin, px); pxmax = max(pxmax, px);
            pymin = min(pymin, py); pymax = max(pymax, py);
        }
    }
    bbox[i] = make_float4(xmin, ymin, xmax, ymax);
    if (bbox_px) bbox_px[i] = make_int4(pxmin, pymin, pxmax, pymax);
    orient[i] = area2 < 0.0 ? -1.0f : 1.0f;
}

__device__ __forceinline__ long long floor_div(long long a, long long b) {  // b > 0
    long long q = a / b;
    return (a % b != 0 && a < 0) ? q - 1 : q;
}
__device__ __forceinline__ long long ceil_div(long long a, long long b) {  // b > 0
    long long q = a / b;
    return (a % b != 0 && a > 0) ? q + 1 : q;
}

// One block per rod; one thread per scanline of the rod's pixel bounding box: exact rational crossings of
// the half-open edges (ymin <= y < ymax) with the line, sorted, closed even-odd intervals filled (decision R2
// of the oracle).  owner = largest draw index + 1 covering the pixel, cover = number of rods covering it.
__global__ void __launch_bounds__(128) k_raster(const int2* __restrict__ hull_px, const int* __restrict__ count,
                                                const int4* __restrict__ bbox_px, int res_x, int res_y,
                                                unsigned* __restrict__ owner, unsigned* __restrict__ cover) {
    __shared__ int2 V[MAXV];
    const int i = blockIdx.x;
    const int m = count[i];
    if (m == 0) return;
    for (int k = threadIdx.x; k < m; k += blockDim.x) V[k] = hull_px[(size_t)i * MAXV + k];
    __syncthreads();
    const int4 bb = bbox_px[i];
    const int y0 = max(bb.y, 0), y1 = min(bb.w, res_y - 1);
    const bool degenerate = bb.x == bb.z || bb.y == bb.w;
    for (int y = y0 + threadIdx.x; y <= y1; y += blockDim.x) {
        auto fill = [&](long long xa, long long xb) {
            const int a = (int)max(xa, 0ll), b = (int)min(xb, (long long)res_x - 1);
            for (int x = a; x <= b; ++x) {
                const size_t q = (size_t)y * res_x + x;
                atomicMax(&owner[q], (unsigned)i + 1u);
                atomicAdd(&cover[q], 1u);
            }
        };
        if (degenerate) {
            fill(bb.x, bb.z);
            continue;
        }
        // crossings as exact rationals num / den (den > 0); at most 16 per scanline
        long long num[16], den[16];
        int nc = 0;
        for (int k = 0; k < m; ++k) {
            int2 a = V[k], b = V[k + 1 == m ? 0 : k + 1];
            if (a.y == b.y) continue;
            if (a.y > b.y) { const int2 t = a; a = b; b = t; }
            if (y < a.y || y >= b.y) continue;
            if (nc < 16) {
                den[nc] = (long long)(b.y - a.y);
                num[nc] = (long long)a.x * den[nc] + (long long)(y - a.y) * (long long)(b.x - a.x);
                ++nc;
            }
        }
        for (int a = 1; a < nc; ++a) {  // insertion sort by value: n1/d1 < n2/d2 <=> n1 d2 < n2 d1
            const long long n_ = num[a], d_ = den[a];
            int b = a;
            while (b > 0 && num[b - 1] * d_ > n_ * den[b - 1]) { num[b] = num[b - 1]; den[b] = den[b - 1]; --b; }
            num[b] = n_; den[b] = d_;
        }
        // (two intervals of one outline may share their end pixel where it touches itself: covered once)
        long long next = -(1ll << 40);
        for (int k = 0; k + 1 < nc; k += 2) {
            const long long xa = max(ceil_div(num[k], den[k]), next), xb = floor_div(num[k + 1], den[k + 1]);
            fill(xa, xb);
            next = max(next, xb + 1);
        }
    }
}

// per pixel: colour of the owning rod unless two or more rods cover it (decision R3); rows flipped like the
// reference's invert_axis(0)
__global__ void k_resolve(const unsigned* __restrict__ owner, const unsigned* __restrict__ cover, const uint8_t* __restrict__ colors,
                          int res_x, int res_y, uint8_t* __restrict__ mask, uint8_t* __restrict__ over) {
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= (size_t)res_x * res_y) return;
    const int y = (int)(q / res_x), x = (int)(q % res_x);
    const unsigned o = owner[q], c = cover[q];
    uint8_t r = 0, g = 0, b = 0;
    if (o && c < 2) { r = colors[3 * (size_t)(o - 1)]; g = colors[3 * (size_t)(o - 1) + 1]; b = colors[3 * (size_t)(o - 1) + 2]; }
    const size_t w = ((size_t)(res_y - 1 - y) * res_x + x) * 3;
    mask[w] = r; mask[w + 1] = g; mask[w + 2] = b;
    const uint8_t v = c >= 2 ? 255 : 0;
    over[w] = v; over[w + 1] = v; over[w + 2] = v;
}

// ---- overlap(): one warp per rod i, all partners j > i whose boxes meet; Green's theorem over the boundary
// pieces of each outline inside the other (decision R4).  Coordinates relative to the box corner of i.
__device__ bool inside_even_odd(const float2* __restrict__ Q, int nq, float ox, float oy, float x, float y) {
    bool c = false;
    for (int k = 0; k < nq; ++k) {
        const float2 a = Q[k], b = Q[k + 1 == nq ? 0 : k + 1];
        const float ya = a.y - oy, yb = b.y - oy;
        if ((ya > y) != (yb > y)) {
            const float xa = a.x - ox, xb = b.x - ox;
            const float xc = xa + (y - ya) * (xb - xa) / (yb - ya);
            if (x < xc) c = !c;
        }
    }
    return c;
}
__device__ float green_edge(const float2* __restrict__ Pp, int np_, int e, const float2* __restrict__ Q, int nq, float ox, float oy) {
    const float2 A = Pp[e], B = Pp[e + 1 == np_ ? 0 : e + 1];
    const float ax = A.x - ox, ay = A.y - oy, rx = B.x - A.x, ry = B.y - A.y;
    float ts[24];
    int nt = 0;
    ts[nt++] = 0.0f;
    for (int k = 0; k < nq && nt < 22; ++k) {
        const float2 C = Q[k], D = Q[k + 1 == nq ? 0 : k + 1];
        const float cx = C.x - ox, cy = C.y - oy, sx = D.x - C.x, sy = D.y - C.y;
        const float den = rx * sy - ry * sx;
        if (den == 0.0f) continue;
        const float t = ((cx - ax) * sy - (cy - ay) * sx) / den;
        const float u = ((cx - ax) * ry - (cy - ay) * rx) / den;
        if (t > 0.0f && t < 1.0f && u >= 0.0f && u <= 1.0f) ts[nt++] = t;
    }
    ts[nt++] = 1.0f;
    for (int a = 1; a < nt; ++a) {
        const float v = ts[a];
        int b = a;
        while (b > 0 && ts[b - 1] > v) { ts[b] = ts[b - 1]; --b; }
        ts[b] = v;
    }
    float acc = 0.0f;
    for (int a = 0; a + 1 < nt; ++a) {
        const float t0 = ts[a], t1 = ts[a + 1];
        if (!(t1 > t0)) continue;
        const float tm = 0.5f * (t0 + t1);
        if (!inside_even_odd(Q, nq, ox, oy, ax + tm * rx, ay + tm * ry)) continue;
        const float x0 = ax + t0 * rx, y0 = ay + t0 * ry, x1 = ax + t1 * rx, y1 = ay + t1 * ry;
        acc += 0.5f * (x0 * y1 - x1 * y0);
    }
    return acc;
}
__global__ void __launch_bounds__(128) k_overlap(const float2* __restrict__ hull, const int* __restrict__ count,
                                                 const float4* __restrict__ bbox, const float* __restrict__ orient, int n,
                                                 double* __restrict__ total) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x & 31;
    if (warp >= n) return;
    const int i = warp, mi = count[i];
    if (mi == 0) return;
    const float4 bi = bbox[i];
    const float2* const Pi = hull + (size_t)i * MAXV;
    const float si = orient[i];
    double mine = 0.0;
    for (int j = i + 1; j < n; ++j) {  // (warp-uniform loop)
        const float4 bj = bbox[j];
        if (bi.x > bj.z || bj.x > bi.z || bi.y > bj.w || bj.y > bi.w) continue;
        const int mj = count[j];
        if (mj == 0) continue;
        const float2* const Pj = hull + (size_t)j * MAXV;
        const float sj = orient[j];
        float acc = 0.0f;
        for (int e = lane; e < mi; e += 32) acc += si * green_edge(Pi, mi, e, Pj, mj, bi.x, bi.y);
        for (int e = lane; e < mj; e += 32) acc += sj * green_edge(Pj, mj, e, Pi, mi, bi.x, bi.y);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        mine += (double)fabsf(acc);
    }
    if (lane == 0 && mine != 0.0) atomicAdd(total, mine);
}

// color_to_counter, reference src/imaging.rs:42-66
__device__ unsigned color_to_counter_dev(unsigned r, unsigned g, unsigned b) {
    const unsigned long long M = 251ull * 251ull * 251ull;
    unsigned long long v = (unsigned long long)r * 251ull * 251ull + (unsigned long long)g * 251ull + b;
    v = (v * 12590168ull) % M;
    v = (v * 13775961ull) % M;
    v = (v * 12157008ull) % M;
    return (unsigned)v;
}
__global__ void k_parents_diff(const uint8_t* __restrict__ m1, const uint8_t* __restrict__ m2, long long n_pixels,
                               const int* __restrict__ cell_of_counter, unsigned n_counters, const int* __restrict__ parent_of,
                               float penalty, float* __restrict__ out, long long* __restrict__ missing) {
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_pixels) return;
    const uint8_t a0 = m1[3 * q], a1 = m1[3 * q + 1], a2 = m1[3 * q + 2];
    const uint8_t b0 = m2[3 * q], b1 = m2[3 * q + 1], b2 = m2[3 * q + 2];
    float v;
    if (a0 == b0 && a1 == b1 && a2 == b2) {
        v = 0.0f;
    } else if ((uint8_t)(a0 + a1 + a2) == 0 || (uint8_t)(b0 + b1 + b2) == 0) {  // u8 `sum()` as written (wraps)
        v = 1.0f;
    } else {
        const unsigned k1 = color_to_counter_dev(a0, a1, a2), k2 = color_to_counter_dev(b0, b1, b2);
        const int i1 = k1 < n_counters ? cell_of_counter[k1] : -1, i2 = k2 < n_counters ? cell_of_counter[k2] : -1;
        if (i1 < 0 || i2 < 0) {
            atomicMin(missing, q);
            v = 1.0f;
        } else {
            v = (parent_of[i2] == i1 || parent_of[i1] == i2) ? penalty : 1.0f;
        }
    }
    out[q] = v;
}

struct Scratch {  // device buffers of one call, released together
    std::vector<void*> ptrs;
    template <class T> int alloc(T*& p, size_t count) {
        p = nullptr;
        if (cudaMalloc(&p, sizeof(T) * (count ? count : 1)) != cudaSuccess) {
            cudaGetLastError();
            return crm_fail(CRM_ERR_NOMEM, "mask path: out of device memory (%zu bytes)", sizeof(T) * count);
        }
        ptrs.push_back(p);
        return 0;
    }
    ~Scratch() { for (void* p : ptrs) cudaFree(p); }
};

struct Outlines {
    float* pos = nullptr;
    float* radii = nullptr;
    float2* hull = nullptr;
    int2* hull_px = nullptr;
    int* count = nullptr;
    float4* bbox = nullptr;
    int4* bbox_px = nullptr;
    float* orient = nullptr;
    int* status = nullptr;
};

// ev_a / ev_b (optional): recorded around the outline kernel alone (the allocations and the copies in front of it
// stall the host, not the device, and are not kernel time)
int outlines(Scratch& sc, Outlines& o, const float* pos, const float* radii, uint32_t n, uint32_t nv, float delta, float eps,
             float size_x, float size_y, int res_x, int res_y, bool pixels, cudaStream_t st, cudaEvent_t ev_a = nullptr,
             cudaEvent_t ev_b = nullptr) {
    int rc;
    if ((rc = sc.alloc(o.pos, (size_t)n * nv * 3)) || (rc = sc.alloc(o.radii, n)) || (rc = sc.alloc(o.hull, (size_t)n * MAXV)) ||
        (rc = sc.alloc(o.count, n)) || (rc = sc.alloc(o.bbox, n)) || (rc = sc.alloc(o.orient, n)) || (rc = sc.alloc(o.status, 1)))
        return rc;
    if (pixels && ((rc = sc.alloc(o.hull_px, (size_t)n * MAXV)) || (rc = sc.alloc(o.bbox_px, n)))) return rc;
    CK(cudaMemcpyAsync(o.pos, pos, sizeof(float) * (size_t)n * nv * 3, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(o.radii, radii, sizeof(float) * n, cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(o.status, 0, sizeof(int), st));
    if (ev_a) CK(cudaEventRecord(ev_a, st));
    if (n) k_hull<<<blocks_for(n, 64), 64, 0, st>>>(o.pos, o.radii, (int)n, (int)nv, delta, eps, size_x, size_y, res_x, res_y, o.hull,
                                                    o.hull_px, o.count, o.bbox, o.bbox_px, o.orient, o.status);
    if (ev_b) CK(cudaEventRecord(ev_b, st));
    CK(cudaGetLastError());
    int h = 0;
    CK(cudaMemcpyAsync(&h, o.status, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (h <= -0x40000000) return crm_fail(CRM_ERR_NOMEM, "outline of rod %d has more than %d points (delta_angle too small)", -h - 0x40000000, MAXV);
    if (h < 0) return crm_fail(CRM_ERR_NUMERIC, "CalcError: lines should be intersecting (outline of rod %d)", -h - 1);
    return 0;
}

}  // namespace

extern "C" {

// calculate_polygon_hull for n rods (reference src/fitting.rs:514-620): out_xy [n][cap][2], out_count [n].
int crm_polygon_hulls(const float* pos, const float* radii, uint32_t n, uint32_t n_vertices, float delta_angle, float epsilon,
                      float* out_xy, uint32_t* out_count, uint32_t cap) {
    if (!pos || !radii || !out_xy || !out_count) return crm_fail(CRM_ERR_INVALID, "null argument");
    if (n_vertices < 2) return crm_fail(CRM_ERR_INVALID, "an outline needs two vertices");
    if (crm_device_count() <= 0) return crm_fail(CRM_ERR_CUDA, "no CUDA device available (cr_mech_coli_b200 has no CPU fallback)");
    Scratch sc;
    Outlines o;
    int rc = outlines(sc, o, pos, radii, n, n_vertices, delta_angle, epsilon, 1.0f, 1.0f, 1, 1, false, 0);
    if (rc) return rc;
    std::vector<float2> h((size_t)n * MAXV);
    std::vector<int> c(n);
    CK(cudaMemcpy(h.data(), o.hull, sizeof(float2) * h.size(), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(c.data(), o.count, sizeof(int) * n, cudaMemcpyDeviceToHost));
    for (uint32_t i = 0; i < n; ++i) {
        if ((uint32_t)c[i] > cap) return crm_fail(CRM_ERR_NOMEM, "outline of rod %u has %d points, the buffer %u", i, c[i], cap);
        out_count[i] = (uint32_t)c[i];
        for (int k = 0; k < c[i]; ++k) {
            out_xy[((size_t)i * cap + k) * 2] = h[(size_t)i * MAXV + k].x;
            out_xy[((size_t)i * cap + k) * 2 + 1] = h[(size_t)i * MAXV + k].y;
        }
    }
    return 0;
}

// render_mask_2d (reference src/fitting.rs:121-244).  Rods in draw order (the reference iterates its BTreeMap
// over CellIdentifier), colors [n][3]; out_mask / out_overlap: [res_y][res_x][3] u8 host buffers.
// kernel_ms (may be NULL): device time of outline + raster + resolve kernels (CUDA events).
int crm_render_mask_2d(const float* pos, const float* radii, const uint8_t* colors, uint32_t n, uint32_t n_vertices,
                       const float domain_size[2], const uint32_t resolution[2], float delta_angle, float epsilon,
                       uint8_t* out_mask, uint8_t* out_overlap, double* kernel_ms) {
    if (!pos || !radii || !colors || !domain_size || !resolution || !out_mask || !out_overlap)
        return crm_fail(CRM_ERR_INVALID, "null argument");
    if (n_vertices < 2) return crm_fail(CRM_ERR_INVALID, "an outline needs two vertices");
    if (crm_device_count() <= 0) return crm_fail(CRM_ERR_CUDA, "no CUDA device available (cr_mech_coli_b200 has no CPU fallback)");
    const int rx = (int)resolution[0], ry = (int)resolution[1];
    const size_t npx = (size_t)rx * ry;
    if (npx == 0) return 0;
    Scratch sc;
    Outlines o;
    cudaEvent_t e0, e1, e2, e3;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaEventCreate(&e2));
    CK(cudaEventCreate(&e3));
    unsigned *owner, *cover;
    uint8_t *dcol, *dmask, *dover;
    int rc;
    if ((rc = sc.alloc(owner, npx)) || (rc = sc.alloc(cover, npx)) || (rc = sc.alloc(dcol, (size_t)n * 3)) ||
        (rc = sc.alloc(dmask, npx * 3)) || (rc = sc.alloc(dover, npx * 3)))
        return rc;
    CK(cudaMemcpyAsync(dcol, colors, (size_t)n * 3, cudaMemcpyHostToDevice, 0));
    if ((rc = outlines(sc, o, pos, radii, n, n_vertices, delta_angle, epsilon, domain_size[0], domain_size[1], rx, ry, true, 0, e2,
                       e3)))
        return rc;
    CK(cudaEventRecord(e0, 0));
    CK(cudaMemsetAsync(owner, 0, sizeof(unsigned) * npx, 0));
    CK(cudaMemsetAsync(cover, 0, sizeof(unsigned) * npx, 0));
    if (n) k_raster<<<n, 128>>>(o.hull_px, o.count, o.bbox_px, rx, ry, owner, cover);
    k_resolve<<<blocks_for(npx, 256), 256>>>(owner, cover, dcol, rx, ry, dmask, dover);
    CK(cudaEventRecord(e1, 0));
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_mask, dmask, npx * 3, cudaMemcpyDeviceToHost, 0));
    CK(cudaMemcpyAsync(out_overlap, dover, npx * 3, cudaMemcpyDeviceToHost, 0));
    CK(cudaStreamSynchronize(0));
    if (kernel_ms) {  // outline kernel + (clear, raster, resolve): the host's wait for the outline status is not in it
        float ms = 0.0f, ms_hull = 0.0f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        CK(cudaEventElapsedTime(&ms_hull, e2, e3));
        *kernel_ms = (double)ms + (double)ms_hull;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaEventDestroy(e2);
    cudaEventDestroy(e3);
    return 0;
}

// overlap (reference src/fitting.rs:299-345): summed intersection area of all pairs of outlines.
int crm_overlap(const float* pos, const float* radii, uint32_t n, uint32_t n_vertices, float delta_angle, float epsilon,
                float* out_area) {
    if (!pos || !radii || !out_area) return crm_fail(CRM_ERR_INVALID, "null argument");
    if (n_vertices < 2) return crm_fail(CRM_ERR_INVALID, "an outline needs two vertices");
    if (crm_device_count() <= 0) return crm_fail(CRM_ERR_CUDA, "no CUDA device available (cr_mech_coli_b200 has no CPU fallback)");
    Scratch sc;
    Outlines o;
    int rc = outlines(sc, o, pos, radii, n, n_vertices, delta_angle, epsilon, 1.0f, 1.0f, 1, 1, false, 0);
    if (rc) return rc;
    double* total;
    if ((rc = sc.alloc(total, 1))) return rc;
    CK(cudaMemsetAsync(total, 0, sizeof(double), 0));
    if (n) k_overlap<<<blocks_for((uint64_t)n * 32, 128), 128>>>(o.hull, o.count, o.bbox, o.orient, (int)n, total);
    CK(cudaGetLastError());
    double h = 0.0;
    CK(cudaMemcpy(&h, total, sizeof(double), cudaMemcpyDeviceToHost));
    *out_area = (float)h;
    return 0;
}

// parents_diff_mask (reference src/fitting.rs:42-109).  cell_of_counter[k] = index of the cell whose colour is
// counter_to_color(k) or -1; parent_of[cell] = index of its parent or -1; out [n_pixels] f32.
// *missing_pixel = first pixel with an unknown colour (the reference raises KeyError) or -1.
int crm_parents_diff_mask(const uint8_t* mask1, const uint8_t* mask2, uint64_t n_pixels, const int32_t* cell_of_counter,
                          uint32_t n_counters, const int32_t* parent_of, uint32_t n_cells, float parent_penalty, float* out,
                          int64_t* missing_pixel) {
    if (!mask1 || !mask2 || !cell_of_counter || !parent_of || !out) return crm_fail(CRM_ERR_INVALID, "null argument");
    if (crm_device_count() <= 0) return crm_fail(CRM_ERR_CUDA, "no CUDA device available (cr_mech_coli_b200 has no CPU fallback)");
    if (missing_pixel) *missing_pixel = -1;
    if (n_pixels == 0) return 0;
    Scratch sc;
    uint8_t *d1, *d2;
    int *dc, *dp;
    float* dout;
    long long* dmiss;
    int rc;
    if ((rc = sc.alloc(d1, n_pixels * 3)) || (rc = sc.alloc(d2, n_pixels * 3)) || (rc = sc.alloc(dc, n_counters)) ||
        (rc = sc.alloc(dp, n_cells)) || (rc = sc.alloc(dout, n_pixels)) || (rc = sc.alloc(dmiss, 1)))
        return rc;
    const long long none = 0x7fffffffffffffffll;
    CK(cudaMemcpyAsync(d1, mask1, n_pixels * 3, cudaMemcpyHostToDevice, 0));
    CK(cudaMemcpyAsync(d2, mask2, n_pixels * 3, cudaMemcpyHostToDevice, 0));
    CK(cudaMemcpyAsync(dc, cell_of_counter, sizeof(int) * n_counters, cudaMemcpyHostToDevice, 0));
    CK(cudaMemcpyAsync(dp, parent_of, sizeof(int) * n_cells, cudaMemcpyHostToDevice, 0));
    CK(cudaMemcpyAsync(dmiss, &none, sizeof(long long), cudaMemcpyHostToDevice, 0));
    k_parents_diff<<<blocks_for(n_pixels, 256), 256>>>(d1, d2, (long long)n_pixels, dc, n_counters, dp, parent_penalty, dout, dmiss);
    CK(cudaGetLastError());
    long long miss = none;
    CK(cudaMemcpyAsync(out, dout, sizeof(float) * n_pixels, cudaMemcpyDeviceToHost, 0));
    CK(cudaMemcpyAsync(&miss, dmiss, sizeof(long long), cudaMemcpyDeviceToHost, 0));
    CK(cudaStreamSynchronize(0));
    if (missing_pixel) *missing_pixel = miss == none ? -1 : (int64_t)miss;
    return 0;
}

}  // extern "C"

// ---- crm_overlap_noalloc: overlap() for callers that cannot allocate (CUDA graph capture, serving loops) ----
// No counterpart in the reference.  The scratch of the outline and
// area kernels (what crm_overlap takes from cudaMalloc) lives in ONE file-scope device array; every GPU that
// loads the library holds it until exit.  All calls share it and it keeps the last call's contents: the words that
// must start at zero are cleared on the stream at the start of every call, and calls that overlap on the GPU race.
namespace {

constexpr uint32_t NOALLOC_MAX_RODS = CRM_NOALLOC_MAX_RODS;
// parts, each 256-B aligned: hull [rods][MAXV] float2 | bbox [rods] float4 | count [rods] int | orient [rods] f32 |
// total (f64) | status (int)
constexpr size_t NA_HULL = 0;
constexpr size_t NA_BBOX = NA_HULL + sizeof(float2) * MAXV * NOALLOC_MAX_RODS;
constexpr size_t NA_COUNT = NA_BBOX + sizeof(float4) * NOALLOC_MAX_RODS;
constexpr size_t NA_ORIENT = NA_COUNT + sizeof(int) * NOALLOC_MAX_RODS;
constexpr size_t NA_TOTAL = NA_ORIENT + sizeof(float) * NOALLOC_MAX_RODS;
constexpr size_t NA_STATUS = NA_TOTAL + 256;
constexpr size_t NA_BYTES = NA_STATUS + 256;
static_assert(NA_BBOX % 256 == 0 && NA_COUNT % 256 == 0 && NA_ORIENT % 256 == 0 && NA_TOTAL % 256 == 0, "parts stay aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the kernels get is
// looked up once per device, with that device current.
constexpr int NA_MAX_DEVICES = 64;
unsigned char* g_noalloc_ptr[NA_MAX_DEVICES] = {};
std::mutex g_noalloc_mutex;

int noalloc_scratch(unsigned char** out) {
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NA_MAX_DEVICES) return crm_fail(CRM_ERR_INVALID, "crm_overlap_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

// status of the outlines as crm_overlap reports it (0, CRM_ERR_NOMEM, CRM_ERR_NUMERIC) and the area, f64 -> f32
__global__ void k_noalloc_finish(const double* __restrict__ total, const int* __restrict__ status, float* __restrict__ out_area,
                                 int* __restrict__ out_status) {
    const int h = *status;
    *out_area = (float)*total;
    *out_status = h <= -0x40000000 ? CRM_ERR_NOMEM : h < 0 ? CRM_ERR_NUMERIC : 0;
}

}  // namespace

extern "C" int crm_overlap_noalloc(const float* d_pos, const float* d_radii, uint32_t n, uint32_t n_vertices, float delta_angle,
                                   float epsilon, float* d_out_area, int32_t* d_out_status, void* stream) {
    if (!d_pos || !d_radii || !d_out_area || !d_out_status) return crm_fail(CRM_ERR_INVALID, "null argument");
    if (n_vertices < 2) return crm_fail(CRM_ERR_INVALID, "an outline needs two vertices");
    if (n > NOALLOC_MAX_RODS)
        return crm_fail(CRM_ERR_INVALID, "crm_overlap_noalloc: %u rods exceed the fixed scratch of %u rods (use crm_overlap)", n,
                        NOALLOC_MAX_RODS);
    if (crm_device_count() <= 0) return crm_fail(CRM_ERR_CUDA, "no CUDA device available (cr_mech_coli_b200 has no CPU fallback)");
    unsigned char* base = nullptr;
    int rc = noalloc_scratch(&base);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float2* hull = reinterpret_cast<float2*>(base + NA_HULL);
    float4* bbox = reinterpret_cast<float4*>(base + NA_BBOX);
    int* count = reinterpret_cast<int*>(base + NA_COUNT);
    float* orient = reinterpret_cast<float*>(base + NA_ORIENT);
    double* total = reinterpret_cast<double*>(base + NA_TOTAL);
    int* status = reinterpret_cast<int*>(base + NA_STATUS);
    // total | status must start at zero (the array keeps the last call's contents); hull / bbox / count / orient of
    // rods 0..n-1 are written by k_hull before k_overlap reads them
    CK(cudaMemsetAsync(total, 0, 512, st));
    if (n) {
        k_hull<<<blocks_for(n, 64), 64, 0, st>>>(d_pos, d_radii, (int)n, (int)n_vertices, delta_angle, epsilon, 1.0f, 1.0f, 1, 1, hull,
                                                  nullptr, count, bbox, nullptr, orient, status);
        k_overlap<<<blocks_for((uint64_t)n * 32, 128), 128, 0, st>>>(hull, count, bbox, orient, (int)n, total);
    }
    k_noalloc_finish<<<1, 1, 0, st>>>(total, status, d_out_area, d_out_status);
    CK(cudaGetLastError());
    return 0;
}
