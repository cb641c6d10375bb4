/*
 * raster_oracle.c — CPU restatement of cr_mech_coli's 2-D mask path.
 *
 * TEST INFRASTRUCTURE ONLY (checker of csrc/crm_raster.cu; same rules as crm_oracle.c).
 *
 * What it follows, reference file:line
 *   calculate_polygon_hull + helpers      src/fitting.rs:347-620   (capsule-chain outline of a rod)
 *   render_mask_2d                        src/fitting.rs:121-244   (instance-colour bitmap + overlap bitmap)
 *   overlap                               src/fitting.rs:299-345   (summed polygon intersection area)
 *   parents_diff_mask                     src/fitting.rs:42-109    (per-pixel difference with parent penalty)
 *
 * PARITY UNPINNED for the pixels: the reference fills polygons through the `plotters` bitmap backend and
 * intersects them with `geo::BooleanOps`; neither crate is vendored.  Pinned by the reference's own unit
 * tests for the outline (src/fitting.rs:755-893: arc points, joint arcs, 12 hull vertices, edge lengths),
 * tests/test_raster.py.  Restatement decisions:
 *   R1  data -> pixel: plotters' RangedCoordf32::map, pixel = floor(res * (v - 0)/(size - 0) + 1e-3) in f64;
 *   R2  a polygon with integer vertices covers pixel (x, y) iff on the scanline y the even-odd crossing
 *       intervals [c0, c1], [c2, c3], ... (closed, half-open edge rule ymin <= y < ymax, exact rational
 *       crossings) contain x; degenerate polygons (one row / one column) cover their bounding line;
 *   R3  the reference paints every pairwise intersection polygon black into the mask and white into the
 *       overlap mask AFTER all cells: restated per pixel as "covered by two or more cells";
 *   R0  cos / sin / atan2 are evaluated in f64 and rounded to f32 (platform-independent, see cos_rn);
 *   R4  overlap(): area of the intersection of two outlines through Green's theorem over the boundary
 *       pieces of each polygon that lie inside the other (exact for simple polygons in general position).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define RASTER_API __attribute__((visibility("default")))
#define PI_F 3.14159265358979323846f

typedef struct { float x, y; } pt;

/* ---- outline ------------------------------------------------------------------------------- */
/* Trigonometric functions through f64, rounded to f32: Rust's f32::cos / sin / atan2 call the platform libm,
 * whose results are within an ulp but not the same on every platform (nor on the GPU).  The f64 routines of
 * glibc and of the device library round to the same f32 (barring a 1e-9 tie), so the CUDA outline is the
 * oracle's bit for bit; either may differ from a given libm's cosf by one ulp. */
static float cos_rn(float x) { return (float)cos((double)x); }
static float sin_rn(float x) { return (float)sin((double)x); }
static float atan2_rn(float y, float x) { return (float)atan2((double)y, (double)x); }

/* get_n_resolution, src/fitting.rs:403-414 */
static int n_resolution(float a0, float a1, float delta) {
    float dangle = fabsf(a0 - a1);
    float n = fmaxf(dangle / delta, 1.0f);
    float frac = fmodf(n, 1.0f);
    int out = (fabsf(frac - 0.0f) <= 0.001f) ? (int)(n - 2.0f) : (int)(n - 1.0f);
    /* `as usize` saturates negative values to 0 */
    if (out < 0) out = 0;
    return out < 1 ? 1 : out;
}

/* generate_coordinates_sphere, src/fitting.rs:348-372 */
static int arc_points(pt p, float r, float a0, float a1, float delta, pt* out) {
    if (a0 == a1) return 0;
    float end = (a0 < a1) ? a1 : a1 + 2.0f * PI_F;
    int n = n_resolution(a0, end, delta);
    float dangle = fabsf(end - a0) / ((float)n + 1.0f);
    for (int k = 1; k <= n; ++k) {
        float ang = a0 + dangle * (float)k;
        out[k - 1].x = p.x + cos_rn(ang) * r;
        out[k - 1].y = p.y + sin_rn(ang) * r;
    }
    return n;
}

/* generate_coordinates_rectangle, src/fitting.rs:374-396 */
static void rect_points(pt p1, pt p2, float r, pt c[4]) {
    float dx = p2.x - p1.x, dy = p2.y - p1.y;
    float norm = sqrtf(dx * dx + dy * dy);
    float d0 = dx / norm, d1 = dy / norm;
    c[0].x = p1.x - r * d1; c[0].y = p1.y + r * d0;
    c[1].x = p2.x - r * d1; c[1].y = p2.y + r * d0;
    c[2].x = p2.x + r * d1; c[2].y = p2.y - r * d0;
    c[3].x = p1.x + r * d1; c[3].y = p1.y - r * d0;
}

/* angle_to_x_axis: atan2(y, x).rem_euclid(2 pi), src/fitting.rs:398-400 */
static float angle_x(float dx, float dy) {
    float a = atan2_rn(dy, dx);
    float m = fmodf(a, 2.0f * PI_F);
    if (m < 0.0f) m += 2.0f * PI_F;
    return m;
}

/* geo::line_intersection for two segments, restated: a single intersection point (proper crossing or a
 * touching endpoint); collinear overlaps and disjoint segments report none (the caller's epsilon fallback). */
static int seg_intersection(pt a, pt b, pt c, pt d, pt* out) {
    float rx = b.x - a.x, ry = b.y - a.y, sx = d.x - c.x, sy = d.y - c.y;
    float den = rx * sy - ry * sx;
    float qx = c.x - a.x, qy = c.y - a.y;
    if (den == 0.0f) {
        /* parallel: a shared endpoint counts as a single point when the segments do not overlap further */
        if (qx * ry - qy * rx != 0.0f) return 0;
        if (b.x == c.x && b.y == c.y) { *out = b; return 1; }
        if (a.x == d.x && a.y == d.y) { *out = a; return 1; }
        return 0;
    }
    float t = (qx * sy - qy * sx) / den;
    float u = (qx * ry - qy * rx) / den;
    if (t < 0.0f || t > 1.0f || u < 0.0f || u > 1.0f) return 0;
    out->x = a.x + t * rx;
    out->y = a.y + t * ry;
    return 1;
}

/* get_coordinates_at_tip, src/fitting.rs:474-500: arc around p1 + c3; returns c0 */
static int tip_points(pt p1, pt p2, float r, float delta, pt* out, pt* c0_out) {
    float dx = p2.x - p1.x, dy = p2.y - p1.y;
    float norm = sqrtf(dx * dx + dy * dy);
    float d0 = dx / norm, d1 = dy / norm;
    float a0 = angle_x(-d1, d0), a1 = angle_x(d1, -d0);
    pt c[4];
    rect_points(p1, p2, r, c);
    int n = arc_points(p1, r, a0, a1, delta, out);
    out[n++] = c[3];
    *c0_out = c[0];
    return n;
}

/* determine_spheroid_coordinates_between_rectangles, src/fitting.rs:416-466: the arc that closes the outer side
 * of the joint at p2; *right = the polyline bends right there. */
static int joint_points(pt p1, pt p2, pt p3, float r, float delta, pt* out, int* right, pt c[4], pt d[4]) {
    rect_points(p1, p2, r, c);
    rect_points(p2, p3, r, d);
    float dir1x = p2.x - p1.x, dir1y = p2.y - p1.y, dir2x = p3.x - p2.x, dir2y = p3.y - p2.y;
    float ang = atan2_rn(dir1x * dir2y - dir1y * dir2x, dir1x * dir2x + dir1y * dir2y);
    *right = ang < 0.0f;
    float a0, a1;
    if (*right) { a0 = angle_x(d[0].x - p2.x, d[0].y - p2.y); a1 = angle_x(c[1].x - p2.x, c[1].y - p2.y); }
    else { a0 = angle_x(c[2].x - p2.x, c[2].y - p2.y); a1 = angle_x(d[3].x - p2.x, d[3].y - p2.y); }
    return arc_points(p2, r, a0, a1, delta, out);
}
/* the two helpers the reference unit-tests directly (src/fitting.rs:755-852) */
RASTER_API int raster_oracle_arc(float px, float py, float r, float a0, float a1, float delta, float* out_xy) {
    pt tmp[512];
    int n = arc_points((pt){px, py}, r, a0, a1, delta, tmp);
    for (int k = 0; k < n; ++k) { out_xy[2 * k] = tmp[k].x; out_xy[2 * k + 1] = tmp[k].y; }
    return n;
}
RASTER_API int raster_oracle_joint(const float* p123, float r, float delta, float* out_xy, int* right) {
    pt tmp[512], c[4], d[4];
    int n = joint_points((pt){p123[0], p123[1]}, (pt){p123[2], p123[3]}, (pt){p123[4], p123[5]}, r, delta, tmp, right, c, d);
    for (int k = 0; k < n; ++k) { out_xy[2 * k] = tmp[k].x; out_xy[2 * k + 1] = tmp[k].y; }
    return n;
}

/* calculate_polygon_hull, src/fitting.rs:514-620.  p: [nv][2]; out: at most cap points; returns the
 * number of outline points, -1 on "lines should be intersecting", -2 when cap is too small. */
RASTER_API int raster_oracle_hull(const float* p, int nv, float r, float delta, float eps, float* out_xy, int cap) {
    const int W = 4096;
    pt* fw = (pt*)malloc(sizeof(pt) * W * 2);
    pt* bw = fw + W;
    pt tmp[512];
    int nf = 0, nb = 0, rc = 0;
#define P(i) ((pt){p[2 * (i)], p[2 * (i) + 1]})
    {
        pt c0;
        int n = tip_points(P(0), P(1), r, delta, tmp, &c0);
        for (int k = 0; k < n; ++k) bw[nb++] = tmp[k];
        fw[nf++] = c0;
    }
    for (int n = 2; n < nv && rc == 0; ++n) {
        pt p1 = P(n - 2), p2 = P(n - 1), p3 = P(n);
        pt c[4], d[4];
        int right;
        int na = joint_points(p1, p2, p3, r, delta, tmp, &right, c, d);
        pt x;
        if (right) {
            fw[nf++] = c[1];
            for (int k = na - 1; k >= 0; --k) fw[nf++] = tmp[k];
            fw[nf++] = d[0];
            if (seg_intersection(c[3], c[2], d[3], d[2], &x)) bw[nb++] = x;
            else if (fabsf(c[2].x - d[3].x) <= eps && fabsf(c[2].y - d[3].y) <= eps) bw[nb++] = c[2];
            else rc = -1;
        } else {
            bw[nb++] = c[2];
            for (int k = 0; k < na; ++k) bw[nb++] = tmp[k];
            bw[nb++] = d[3];
            if (seg_intersection(c[0], c[1], d[0], d[1], &x)) fw[nf++] = x;
            else if (fabsf(c[1].x - d[0].x) <= eps && fabsf(c[1].y - d[0].y) <= eps) fw[nf++] = c[1];
            else rc = -1;
        }
        if (nf > W - 600 || nb > W - 600) rc = -2;
    }
    if (rc == 0) {
        pt cfin;
        int n = tip_points(P(nv - 1), P(nv - 2), r, delta, tmp, &cfin);
        bw[nb++] = cfin;
        for (int k = n - 1; k >= 0; --k) fw[nf++] = tmp[k];
        if (nf + nb > cap) rc = -2;
        else {
            int o = 0;
            for (int k = 0; k < nf; ++k) { out_xy[2 * o] = fw[k].x; out_xy[2 * o + 1] = fw[k].y; ++o; }
            for (int k = nb - 1; k >= 0; --k) { out_xy[2 * o] = bw[k].x; out_xy[2 * o + 1] = bw[k].y; ++o; }
            rc = o;
        }
    }
#undef P
    free(fw);
    return rc;
}

/* ---- R1: data -> pixel (plotters RangedCoordf32::map with limit (0, res)) ------------------- */
RASTER_API int raster_oracle_to_pixel(float v, float size, int res) {
    double logic = ((double)v - 0.0) / ((double)size - 0.0);
    return (int)floor((double)res * logic + 1e-3);
}

/* ---- R2: does the integer polygon cover pixel (x, y)? ------------------------------------------ */
static int covers(const int* vx, const int* vy, int n, int xmin, int xmax, int ymin, int ymax, int x, int y) {
    if (x < xmin || x > xmax || y < ymin || y > ymax) return 0;
    if (xmin == xmax || ymin == ymax) return 1; /* degenerate: the bounding line */
    int less = 0, leq = 0;
    for (int i = 0; i < n; ++i) {
        int j = (i + 1 == n) ? 0 : i + 1;
        int xa = vx[i], ya = vy[i], xb = vx[j], yb = vy[j];
        if (ya == yb) continue;
        if (ya > yb) { int t = xa; xa = xb; xb = t; t = ya; ya = yb; yb = t; }
        if (y < ya || y >= yb) continue;
        /* crossing xc = xa + (y - ya)(xb - xa)/(yb - ya); compare with x exactly */
        int64_t lhs = (int64_t)(y - ya) * (int64_t)(xb - xa);
        int64_t rhs = (int64_t)(x - xa) * (int64_t)(yb - ya);
        less += lhs < rhs;
        leq += lhs <= rhs;
    }
    return (leq > less) || (less & 1);
}

/* render_mask_2d.  pos: [n][nv][3] (xy used), radii [n], colors [n][3] in draw order (the reference iterates a
 * BTreeMap over CellIdentifier).  out_mask / out_overlap: [res_y][res_x][3], row 0 = largest y (the reference
 * inverts axis 0).  Returns 0, or -1 - i when the outline of rod i fails. */
RASTER_API int raster_oracle_render(const float* pos, const float* radii, const uint8_t* colors, int n, int nv,
                                    float size_x, float size_y, int res_x, int res_y, float delta, float eps,
                                    uint8_t* out_mask, uint8_t* out_overlap) {
    const int CAP = 2048;
    size_t npx = (size_t)res_x * res_y;
    uint32_t* owner = (uint32_t*)calloc(npx, sizeof(uint32_t));
    uint8_t* cnt = (uint8_t*)calloc(npx, 1);
    float* xy = (float*)malloc(sizeof(float) * 2 * CAP);
    float* p2 = (float*)malloc(sizeof(float) * 2 * nv);
    int* vx = (int*)malloc(sizeof(int) * CAP);
    int* vy = (int*)malloc(sizeof(int) * CAP);
    int rc = 0;
    for (int i = 0; i < n && rc == 0; ++i) {
        for (int k = 0; k < nv; ++k) {
            p2[2 * k] = pos[((size_t)i * nv + k) * 3];
            p2[2 * k + 1] = pos[((size_t)i * nv + k) * 3 + 1];
        }
        int m = raster_oracle_hull(p2, nv, radii[i], delta, eps, xy, CAP);
        if (m < 0) { rc = -1 - i; break; }
        int xmin = 1 << 30, xmax = -(1 << 30), ymin = 1 << 30, ymax = -(1 << 30);
        for (int k = 0; k < m; ++k) {
            vx[k] = raster_oracle_to_pixel(xy[2 * k], size_x, res_x);
            vy[k] = raster_oracle_to_pixel(xy[2 * k + 1], size_y, res_y);
            if (vx[k] < xmin) xmin = vx[k];
            if (vx[k] > xmax) xmax = vx[k];
            if (vy[k] < ymin) ymin = vy[k];
            if (vy[k] > ymax) ymax = vy[k];
        }
        int x0 = xmin < 0 ? 0 : xmin, x1 = xmax >= res_x ? res_x - 1 : xmax;
        int y0 = ymin < 0 ? 0 : ymin, y1 = ymax >= res_y ? res_y - 1 : ymax;
        for (int y = y0; y <= y1; ++y)
            for (int x = x0; x <= x1; ++x)
                if (covers(vx, vy, m, xmin, xmax, ymin, ymax, x, y)) {
                    size_t q = (size_t)y * res_x + x;
                    owner[q] = (uint32_t)i + 1u; /* later cells paint over earlier ones */
                    if (cnt[q] < 2) cnt[q]++;
                }
    }
    if (rc == 0)
        for (int y = 0; y < res_y; ++y)
            for (int x = 0; x < res_x; ++x) {
                size_t q = (size_t)y * res_x + x;
                size_t o = ((size_t)(res_y - 1 - y) * res_x + x) * 3;
                uint8_t c[3] = {0, 0, 0};
                if (owner[q] && cnt[q] < 2) memcpy(c, colors + 3 * (size_t)(owner[q] - 1), 3);
                memcpy(out_mask + o, c, 3);
                memset(out_overlap + o, cnt[q] >= 2 ? 255 : 0, 3);
            }
    free(owner); free(cnt); free(xy); free(p2); free(vx); free(vy);
    return rc;
}

/* ---- R4: intersection area of two simple polygons ------------------------------------------------ */
static int inside_even_odd(const double* q, int n, double x, double y) {
    int c = 0;
    for (int i = 0; i < n; ++i) {
        int j = (i + 1 == n) ? 0 : i + 1;
        double xa = q[2 * i], ya = q[2 * i + 1], xb = q[2 * j], yb = q[2 * j + 1];
        if ((ya > y) != (yb > y)) {
            double xc = xa + (y - ya) * (xb - xa) / (yb - ya);
            if (x < xc) c ^= 1;
        }
    }
    return c;
}
static double signed_area(const double* p, int n) {
    double a = 0;
    for (int i = 0; i < n; ++i) {
        int j = (i + 1 == n) ? 0 : i + 1;
        a += p[2 * i] * p[2 * j + 1] - p[2 * j] * p[2 * i + 1];
    }
    return 0.5 * a;
}
/* boundary pieces of p that lie inside q, integrated as 1/2 (x dy - y dx) */
static double green_part(const double* p, int np_, const double* q, int nq) {
    double acc = 0;
    for (int i = 0; i < np_; ++i) {
        int j = (i + 1 == np_) ? 0 : i + 1;
        double ax = p[2 * i], ay = p[2 * i + 1], bx = p[2 * j], by = p[2 * j + 1];
        double rx = bx - ax, ry = by - ay;
        double ts[64];
        int nt = 0;
        ts[nt++] = 0.0;
        for (int k = 0; k < nq && nt < 62; ++k) {
            int l = (k + 1 == nq) ? 0 : k + 1;
            double cx = q[2 * k], cy = q[2 * k + 1], sx = q[2 * l] - cx, sy = q[2 * l + 1] - cy;
            double den = rx * sy - ry * sx;
            if (den == 0.0) continue;
            double t = ((cx - ax) * sy - (cy - ay) * sx) / den;
            double u = ((cx - ax) * ry - (cy - ay) * rx) / den;
            if (t > 0.0 && t < 1.0 && u >= 0.0 && u <= 1.0) ts[nt++] = t;
        }
        ts[nt++] = 1.0;
        for (int a = 1; a < nt; ++a) { /* insertion sort */
            double v = ts[a];
            int b = a;
            while (b > 0 && ts[b - 1] > v) { ts[b] = ts[b - 1]; --b; }
            ts[b] = v;
        }
        for (int a = 0; a + 1 < nt; ++a) {
            double t0 = ts[a], t1 = ts[a + 1];
            if (!(t1 > t0)) continue;
            double tm = 0.5 * (t0 + t1);
            if (!inside_even_odd(q, nq, ax + tm * rx, ay + tm * ry)) continue;
            double x0 = ax + t0 * rx, y0 = ay + t0 * ry, x1 = ax + t1 * rx, y1 = ay + t1 * ry;
            acc += 0.5 * (x0 * y1 - x1 * y0);
        }
    }
    return acc;
}

/* overlap(positions, radii, delta_angle, epsilon), src/fitting.rs:299-345: summed intersection area of all
 * pairs of outlines (the reference's bounding-box filter passes every pair, :262-297: its boxes start at
 * -inf/+inf; a true bounding-box reject gives the same sum). */
RASTER_API int raster_oracle_overlap(const float* pos, const float* radii, int n, int nv, float delta, float eps,
                                     double* out_area) {
    const int CAP = 2048;
    float* xy = (float*)malloc(sizeof(float) * 2 * CAP);
    float* p2 = (float*)malloc(sizeof(float) * 2 * nv);
    double** poly = (double**)calloc(n, sizeof(double*));
    int* cnt = (int*)calloc(n, sizeof(int));
    double* bb = (double*)malloc(sizeof(double) * 4 * (n ? n : 1));
    int rc = 0;
    for (int i = 0; i < n && rc == 0; ++i) {
        for (int k = 0; k < nv; ++k) {
            p2[2 * k] = pos[((size_t)i * nv + k) * 3];
            p2[2 * k + 1] = pos[((size_t)i * nv + k) * 3 + 1];
        }
        int m = raster_oracle_hull(p2, nv, radii[i], delta, eps, xy, CAP);
        if (m < 0) { rc = -1 - i; break; }
        poly[i] = (double*)malloc(sizeof(double) * 2 * m);
        cnt[i] = m;
        bb[4 * i] = bb[4 * i + 1] = 1e300;
        bb[4 * i + 2] = bb[4 * i + 3] = -1e300;
        for (int k = 0; k < m; ++k) {
            poly[i][2 * k] = xy[2 * k];
            poly[i][2 * k + 1] = xy[2 * k + 1];
            if (xy[2 * k] < bb[4 * i]) bb[4 * i] = xy[2 * k];
            if (xy[2 * k + 1] < bb[4 * i + 1]) bb[4 * i + 1] = xy[2 * k + 1];
            if (xy[2 * k] > bb[4 * i + 2]) bb[4 * i + 2] = xy[2 * k];
            if (xy[2 * k + 1] > bb[4 * i + 3]) bb[4 * i + 3] = xy[2 * k + 1];
        }
        if (signed_area(poly[i], m) < 0) /* orient counter-clockwise */
            for (int k = 0; k < m / 2; ++k) {
                double tx = poly[i][2 * k], ty = poly[i][2 * k + 1];
                poly[i][2 * k] = poly[i][2 * (m - 1 - k)];
                poly[i][2 * k + 1] = poly[i][2 * (m - 1 - k) + 1];
                poly[i][2 * (m - 1 - k)] = tx;
                poly[i][2 * (m - 1 - k) + 1] = ty;
            }
    }
    double total = 0;
    if (rc == 0)
        for (int i = 0; i < n; ++i)
            for (int j = i + 1; j < n; ++j) {
                if (bb[4 * i] > bb[4 * j + 2] || bb[4 * j] > bb[4 * i + 2] || bb[4 * i + 1] > bb[4 * j + 3] ||
                    bb[4 * j + 1] > bb[4 * i + 3])
                    continue;
                /* local origin: the cross products cancel badly far from it */
                double ox = bb[4 * i], oy = bb[4 * i + 1];
                int mi = cnt[i], mj = cnt[j];
                double* a = (double*)malloc(sizeof(double) * 2 * (mi + mj));
                double* b = a + 2 * mi;
                for (int k = 0; k < mi; ++k) { a[2 * k] = poly[i][2 * k] - ox; a[2 * k + 1] = poly[i][2 * k + 1] - oy; }
                for (int k = 0; k < mj; ++k) { b[2 * k] = poly[j][2 * k] - ox; b[2 * k + 1] = poly[j][2 * k + 1] - oy; }
                double ar = green_part(a, mi, b, mj) + green_part(b, mj, a, mi);
                total += fabs(ar);
                free(a);
            }
    for (int i = 0; i < n; ++i) free(poly[i]);
    free(poly); free(cnt); free(bb); free(xy); free(p2);
    *out_area = total;
    return rc;
}

/* parents_diff_mask, src/fitting.rs:42-109.  Colours are decoded with color_to_counter (src/imaging.rs:42-66;
 * the bijection the container's cell_to_color is built from, src/cell_container.rs:76-80): cell_of_counter[k] =
 * index of the cell with colour counter k or -1; parent_of[cell] = index of its parent or -1.
 * Returns 0 or -(1 + first pixel whose colour is unknown) (the reference raises KeyError). */
static uint32_t color_to_counter(const uint8_t* c) {
    const uint64_t M = 251ull * 251ull * 251ull;
    uint64_t v = (uint64_t)c[0] * 251ull * 251ull + (uint64_t)c[1] * 251ull + c[2];
    v = (v * 12590168ull) % M;
    v = (v * 13775961ull) % M;
    v = (v * 12157008ull) % M;
    return (uint32_t)v;
}
RASTER_API int64_t raster_oracle_parents_diff(const uint8_t* m1, const uint8_t* m2, int64_t n_pixels,
                                              const int32_t* cell_of_counter, uint32_t n_counters,
                                              const int32_t* parent_of, float penalty, float* out) {
    for (int64_t q = 0; q < n_pixels; ++q) {
        const uint8_t* c1 = m1 + 3 * q;
        const uint8_t* c2 = m2 + 3 * q;
        int same = c1[0] == c2[0] && c1[1] == c2[1] && c1[2] == c2[2];
        /* `c.sum() != 0` on u8 arrays wraps modulo 256 in ndarray's release builds; restated as written */
        uint8_t s1 = (uint8_t)(c1[0] + c1[1] + c1[2]), s2 = (uint8_t)(c2[0] + c2[1] + c2[2]);
        if (same) { out[q] = 0.0f; continue; }
        if (s1 == 0 || s2 == 0) { out[q] = 1.0f; continue; }
        uint32_t k1 = color_to_counter(c1), k2 = color_to_counter(c2);
        if (k1 >= n_counters || k2 >= n_counters || cell_of_counter[k1] < 0 || cell_of_counter[k2] < 0) return -(1 + q);
        int32_t i1 = cell_of_counter[k1], i2 = cell_of_counter[k2];
        out[q] = (parent_of[i2] == i1 || parent_of[i1] == i2) ? penalty : 1.0f;
    }
    return 0;
}
