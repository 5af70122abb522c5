// View-dependent conservative film-space bounds of one triangle slot ("film rect"), shared by the per-frame CUDA
// pre-pass (render.cu: k_view_bounds) and the host (capi.cpp: atray_debug_film_rects, used by the CPU soundness tests).
//
// Purpose. Every camera ray of a frame starts at the eye and is identified by its film coordinates (fx, fy)
// (renderer.cpp:317,329,340-341,350-351: pixel_pos = frame_center + camera_x*fx + camera_y*fy). The reference accepts a
// triangle for a ray only if its float Möller–Trumbore test (model.h:75-103) passes. This file computes, per triangle
// and per frame, a rectangle R in film space such that
//
//        reference accepts the triangle for the ray with film coordinates (fx, fy)   ==>   (fx, fy) lies in R.
//
// Rays outside R can skip the triangle without changing any result bit. The implication is proved for the float
// arithmetic of the reference (not for exact geometry) by the error analysis below; DESIGN.md §4b restates it.
//
// Error analysis (u = 2^-24, all quantities real unless hatted; a, ab, ac = the float record values, N = ac x ab,
// S = |ab||ac|, T = |eye - a|, tau = 1e-4f). With the standard model fl(x∘y) = (x∘y)(1+δ), |δ| <= u:
//   |pvec^ - pvec| <= 2.42u|d||ac|;  |det^ - det| <= 5.43u S|d|;  |Nu^ - Nu| <= 6.43u T|d||ac|;  |Nv^ - Nv| <= 6.43u T|d||ab|
//   u^ = Nu^/det^ (1+2δ), v^ = Nv^/det^ (1+2δ), accepted  =>  det^ >= tau, 0 <= u^, 0 <= v^, u^+v^ <= 1+u.
// All constants are rounded up to 8u. Let c = 8u(1+1e-6)/max(tau, det_lo). Then the barycentric errors satisfy
//   Eu <= (c|ac|(T+|ab|) + 3u)/(1 - cS - 3u),  Ev <= (c|ab|(T+|ac|) + 3u)/(1 - cS - 3u)
// and the true line/plane intersection P* = a + u* ab + v* ac lies within
//   m = [cS(2T + |ab| + |ac|) + 3u(|ab| + |ac|)]/(1 - cS - 3u) + u max(|ab|,|ac|)
// of the triangle. Hence an accepted ray passes within m of the triangle. det_lo is a lower bound of det^ over all
// rays from the eye that pass within m of the triangle: det = d.N = ((a-eye).N + δ.N)/s with |δ| <= m and
// s <= (Tmax + m)(1+2u), so det^ >= ((a-eye).N - m|N|)/((Tmax+m)(1+1e-6)) - 8uS. m is first computed with det_lo = tau
// (valid for every accepted ray), then refined twice (each step is valid given the previous one). A triangle with
// (a-eye).N + m|N| <= 0 has det^ <= 8uS < tau for every such ray and is never accepted (back-facing for this eye).
// The float ray generation (camera_ray in render.cu = renderer.cpp:350-351 + SetRay) deviates from the ideal line
// through the film point by an angle <= ang (ViewParams.ang, bounded on the host); this adds ang*(Tmax+m) to m.
// The inflated triangle (vertex balls of radius m) is projected to film space with interval arithmetic in double and
// rounded outward to float. Triangles that cannot be bounded (non-finite input, huge coordinates, a vertex ball that
// reaches the eye plane, cS >= 0.5) get the infinite rectangle.
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define ATRAY_HD __host__ __device__ __forceinline__
#else
#define ATRAY_HD inline
#endif

namespace atray {

struct ViewParams
{
    double eye[3];
    double minv[3][3];   // (alpha, beta, gamma) = minv * (x - eye); film point of x = (alpha/gamma, beta/gamma)
    double row[3];       // Euclidean norms of the rows of minv, rounded up
    double ang;          // angular deviation bound of the float ray generation from the ideal line (radians)
    double kx;           // h_fov * aspect: film_x(x) = (-1 + 2x/W) * kx   (renderer.cpp:329)
    double px_slack_x, px_slack_y;  // pixel slack of film_rect_px: float rounding of film_x / film_y plus the jitter of an AA sample
    double inv_kx;                  // 1 / kx
    double fx_max, fy_max;          // bounds of |film_x|, |film_y| over every ray of the frame (jitter included)
    int32_t W, H;
    int32_t tiles_x, tiles_y;      // 8x4-pixel tiles
    int32_t valid;                 // 0: camera not usable for culling (singular basis / non-finite) → unculled kernel
    uint32_t frame_id;             // stamp value of this frame
};

struct FilmRect { float xlo, ylo, xhi, yhi; };
// (Groundwork, verified on the CPU by tests/test_view_bounds.py; the kernels do not read it yet — measured this round:
// testing (slot, ray) pairs against it cuts the triangle tests 4.6x but adds a dependent load in front of every test and
// the latency-bound render kernel got slower, 269 -> 292 us.)
// Three half-planes a[k]*fx + b[k]*fy <= c[k] (evaluated in float, any association, with or without FMA) that hold for the
// film point of every ray the reference accepts for the triangle. The film point lies in the projection of the inflated
// triangle = convex hull of the three projected vertex balls (a perspective map with gamma > 0 sends segments to
// segments), each ball lies in its axis-parallel film box, so any half-plane containing the three boxes is sound. One
// half-plane per edge, oriented away from the third vertex, offset = maximum over the 12 box corners plus a slack for
// every float rounding on the way (see tri_film_rect). (0, 0, 0) is the trivial half-plane.
struct FilmHull { float a[3], b[3], c[3]; };
ATRAY_HD FilmHull film_hull_all()
{
    FilmHull h;
    for (int k = 0; k < 3; ++k) { h.a[k] = 0.f; h.b[k] = 0.f; h.c[k] = 0.f; }
    return h;
}

ATRAY_HD FilmRect film_rect_empty()
{
    FilmRect r; r.xlo = INFINITY; r.ylo = INFINITY; r.xhi = -INFINITY; r.yhi = -INFINITY; return r;
}
ATRAY_HD FilmRect film_rect_all()
{
    FilmRect r; r.xlo = -INFINITY; r.ylo = -INFINITY; r.xhi = INFINITY; r.yhi = INFINITY; return r;
}
ATRAY_HD bool film_rect_is_empty(const FilmRect& r) { return r.xlo > r.xhi || r.ylo > r.yhi; }
ATRAY_HD FilmRect film_rect_union(const FilmRect& a, const FilmRect& b)
{
    FilmRect r;
    r.xlo = a.xlo < b.xlo ? a.xlo : b.xlo; r.ylo = a.ylo < b.ylo ? a.ylo : b.ylo;
    r.xhi = a.xhi > b.xhi ? a.xhi : b.xhi; r.yhi = a.yhi > b.yhi ? a.yhi : b.yhi;
    return r;
}

ATRAY_HD float round_down_f(double x)
{
#if defined(__CUDA_ARCH__)
    return __double2float_rd(x);
#else
    float f = (float)x;
    if ((double)f > x) f = nextafterf(f, -INFINITY);
    return f;
#endif
}
ATRAY_HD float round_up_f(double x)
{
#if defined(__CUDA_ARCH__)
    return __double2float_ru(x);
#else
    float f = (float)x;
    if ((double)f < x) f = nextafterf(f, INFINITY);
    return f;
#endif
}

// Refinements of the lower bound of det among accepted rays (each costs three FP64 divides per triangle and frame): the
// first replaces the cull threshold 1e-4 by the triangle's own bound and shrinks the margin ~100x; a third changes the
// rectangle by less than 1e-3 pixel on the 4K frame. ATRAY_MARGIN_ITERS can be defined at build time for experiments.
#ifndef ATRAY_MARGIN_ITERS
#define ATRAY_MARGIN_ITERS 2
#endif
static constexpr int kMarginIters = ATRAY_MARGIN_ITERS;

// Cheap single-precision pre-test: true ONLY IF tri_film_rect would return the empty rectangle because the triangle faces
// away from the eye (its first rule: g + m|N| <= 0 with g = (a - eye).N, N = ac x ab). About half of a closed mesh passes, and
// those slots then skip the double-precision path altogether (k_view_bounds compacts the rest).
// With u = 2^-24 and S >= |ab||ac|, T >= |a - eye|, L >= max(|ab|, |ac|) (all rounded up):
//   * |N| <= S, and for cS <= 1/4 (c = 8u(1+1e-6)/tau) the margin obeys m <= 1.34 [cS(2T + 2L) + 6uL] + uL + 1e-25
//     <= 2.7 cS(T + L) + 10uL, hence m|N| <= S (2.7 cS(T + L) + 10uL);
//   * g evaluated in float (w = a - eye, the cross product, the dot product: every operation rounded once) differs from the
//     real g by at most 12u T S (each of the three terms carries <= 3 roundings of relative size u on magnitudes <= T S, plus
//     the summation).
// So g_float <= -(1 + 1e-3)[S (2.7 cS(T + L) + 10uL) + 12u T S] - 1e-30 implies g + m|N| <= 0. The norms are float square roots
// of float sums of squares scaled by (1 + 1e-5); anything non-finite or out of range (overflow, S > 52, tiny values below the
// absolute floor) simply fails the test and takes the exact path. The eye of ViewParams is a float value held in double.
ATRAY_HD bool tri_backfacing_f32(const ViewParams& vp, const float* a, const float* ab, const float* ac)
{
    const float U = 5.9604644775390625e-8f;
    const float wx = a[0] - (float)vp.eye[0], wy = a[1] - (float)vp.eye[1], wz = a[2] - (float)vp.eye[2];
    const float nx = ac[1] * ab[2] - ac[2] * ab[1], ny = ac[2] * ab[0] - ac[0] * ab[2], nz = ac[0] * ab[1] - ac[1] * ab[0];
    const float g = wx * nx + wy * ny + wz * nz;
    const float k = 1.00001f;
    const float lab = sqrtf(ab[0] * ab[0] + ab[1] * ab[1] + ab[2] * ab[2]) * k, lac = sqrtf(ac[0] * ac[0] + ac[1] * ac[1] + ac[2] * ac[2]) * k;
    const float T = sqrtf(wx * wx + wy * wy + wz * wz) * k;
    const float S = lab * lac * k, L = (lab > lac ? lab : lac);
    const float c = 8.0f * U * 1.0001f / 9.9999e-5f;      // > 8u(1+1e-6)/tau
    const float cS = c * S;
    // (false for NaN too; coordinates beyond 2^30 take the exact path, which gives them the unbounded rectangle)
    if (!(cS <= 0.25f) || !(T < 1e9f) || !(L < 1e9f) || !(fabsf(a[0]) < 1e9f && fabsf(a[1]) < 1e9f && fabsf(a[2]) < 1e9f)) return false;
    const float bound = (S * (2.7f * cS * (T + L) + 10.0f * U * L) + 12.0f * U * T * S) * 1.001f + 1e-30f;
    return g <= -bound;
}

// Film rectangle of one triangle record {a, ab, ac} (float values widened to double) for the view `vp`.
ATRAY_HD FilmRect tri_film_rect(const ViewParams& vp, const float* a_f, const float* ab_f, const float* ac_f, FilmHull* hull = nullptr)
{
    if (hull) *hull = film_hull_all();
    if (tri_backfacing_f32(vp, a_f, ab_f, ac_f)) return film_rect_empty();
    const double U = 5.9604644775390625e-8;   // 2^-24
    const double TAU = 9.9999e-5;             // just below (double)1e-4f = 9.99999974738e-05
    const double BIG = 1073741824.0;          // 2^30: beyond this the products of the test may overflow → no bound
    double a[3], ab[3], ac[3], w[3][3];
    bool ok = true;
    for (int i = 0; i < 3; ++i)
    {
        a[i] = (double)a_f[i]; ab[i] = (double)ab_f[i]; ac[i] = (double)ac_f[i];
        ok = ok && fabs(a[i]) < BIG && fabs(ab[i]) < BIG && fabs(ac[i]) < BIG;   // false for NaN/Inf too
    }
    if (!ok) return film_rect_all();
    for (int i = 0; i < 3; ++i)
    {
        w[0][i] = a[i] - vp.eye[i];            // vertex a, b' = a + ab, c' = a + ac relative to the eye
        w[1][i] = w[0][i] + ab[i];
        w[2][i] = w[0][i] + ac[i];
    }
    const double lab = sqrt(ab[0] * ab[0] + ab[1] * ab[1] + ab[2] * ab[2]);
    const double lac = sqrt(ac[0] * ac[0] + ac[1] * ac[1] + ac[2] * ac[2]);
    const double S = lab * lac * (1.0 + 1e-12);
    const double L = lab > lac ? lab : lac;
    // det = ab.(d x ac) = d.(ac x ab)  (model.h:81-82)  →  N = ac x ab
    const double N[3] = {ac[1] * ab[2] - ac[2] * ab[1], ac[2] * ab[0] - ac[0] * ab[2], ac[0] * ab[1] - ac[1] * ab[0]};
    const double lN = sqrt(N[0] * N[0] + N[1] * N[1] + N[2] * N[2]) * (1.0 + 1e-12);
    const double g = w[0][0] * N[0] + w[0][1] * N[1] + w[0][2] * N[2];   // (a - eye).N
    double T[3], Tmax = 0.0;
    for (int k = 0; k < 3; ++k)
    {
        T[k] = sqrt(w[k][0] * w[k][0] + w[k][1] * w[k][1] + w[k][2] * w[k][2]) * (1.0 + 1e-12);
        if (T[k] > Tmax) Tmax = T[k];
    }
    if (!(Tmax < 4.0 * BIG)) return film_rect_all();
    const double err_det = 8.0 * U * S * (1.0 + 1e-6);
    // margin of the float test as a function of the lower bound `den` of det^ among accepted rays
    double m = 0.0, den = TAU;
    for (int it = 0; it < kMarginIters; ++it)
    {
        const double c = 8.0 * U * (1.0 + 1e-6) / den;
        const double cS = c * S;
        if (!(cS < 0.5)) return film_rect_all();
        m = (cS * (2.0 * T[0] + lab + lac) + 3.0 * U * (lab + lac)) / (1.0 - cS - 3.0 * U) + U * L + 1e-25;
        m *= (1.0 + 1e-9);
        if (it == 0)
        {
            // never accepted from this eye: det^ <= det + 8uS with det <= ((a-eye).N + m|N|)/s
            if (err_det < TAU)
            {
                const double top = g + m * lN;
                if (top <= 0.0) return film_rect_empty();
                const double smin = (fabs(g) / lN - m) * (1.0 - 1e-6);   // distance eye → plane, minus the margin
                if (smin > 0.0 && top / smin + err_det < TAU) return film_rect_empty();
            }
        }
        const double lo = (g - m * lN) / ((Tmax + m) * (1.0 + 1e-6)) - err_det;
        if (!(lo > den)) break;   // no improvement (also NaN-safe)
        den = lo;
    }
    // total radius of the vertex balls: float-test margin + deviation of the float ray from the ideal line
    const double rad = (m + vp.ang * (Tmax + m)) * (1.0 + 1e-9);
    double xlo = INFINITY, xhi = -INFINITY, ylo = INFINITY, yhi = -INFINITY;
    double bx0[3], bx1[3], by0[3], by1[3];  // film box of each inflated vertex
    for (int k = 0; k < 3; ++k)
    {
        const double al = vp.minv[0][0] * w[k][0] + vp.minv[0][1] * w[k][1] + vp.minv[0][2] * w[k][2];
        const double be = vp.minv[1][0] * w[k][0] + vp.minv[1][1] * w[k][1] + vp.minv[1][2] * w[k][2];
        const double ga = vp.minv[2][0] * w[k][0] + vp.minv[2][1] * w[k][1] + vp.minv[2][2] * w[k][2];
        // rounding of the three dot products above (double): 4 ulp of the largest term
        const double e = 1e-15 * (fabs(w[k][0]) + fabs(w[k][1]) + fabs(w[k][2])) * (vp.row[0] + vp.row[1] + vp.row[2]);
        const double ra = vp.row[0] * rad + e, rb = vp.row[1] * rad + e, rg = vp.row[2] * rad + e;
        const double g0 = ga - rg, g1 = ga + rg;
        if (!(g0 > 1e-9 * (1.0 + fabs(al) + fabs(be)))) return film_rect_all();   // ball reaches the eye plane
        const double a0 = al - ra, a1 = al + ra, b0 = be - rb, b1 = be + rb;
        // two reciprocals instead of four quotients (FP64 divides are what the pre-pass spends its time on): the product
        // is within 1.5 ulp of the quotient, far inside the 1e-12 relative slack applied to the bounds below
        const double i0 = 1.0 / g0, i1 = 1.0 / g1;
        const double x0 = a0 * (a0 < 0.0 ? i0 : i1), x1 = a1 * (a1 < 0.0 ? i1 : i0);
        const double y0 = b0 * (b0 < 0.0 ? i0 : i1), y1 = b1 * (b1 < 0.0 ? i1 : i0);
        if (x0 < xlo) xlo = x0;
        if (x1 > xhi) xhi = x1;
        if (y0 < ylo) ylo = y0;
        if (y1 > yhi) yhi = y1;
        bx0[k] = x0 - 1e-12 * (1.0 + fabs(x0)); bx1[k] = x1 + 1e-12 * (1.0 + fabs(x1));
        by0[k] = y0 - 1e-12 * (1.0 + fabs(y0)); by1[k] = y1 + 1e-12 * (1.0 + fabs(y1));
    }
    if (!(xlo <= xhi && ylo <= yhi)) return film_rect_all();   // NaN somewhere
    if (hull)
    {
        // In float (the FP64 pipe is what bounds the pre-pass): the boxes are rounded outward to float, the edge normal and
        // the offset are float expressions, and every rounding — of the offset's own evaluation, of its sum with the slack,
        // and of the kernel's a*fx + b*fy — is below 3u(|a| Mx + |b| My) with Mx, My bounding every coordinate involved
        // (u = 2^-24); 24u is added.
        float fx0[3], fx1[3], fy0[3], fy1[3];
        float Mx = round_up_f(vp.fx_max), My = round_up_f(vp.fy_max);
        for (int k = 0; k < 3; ++k)
        {
            fx0[k] = round_down_f(bx0[k]); fx1[k] = round_up_f(bx1[k]); fy0[k] = round_down_f(by0[k]); fy1[k] = round_up_f(by1[k]);
            Mx = fmaxf(Mx, fmaxf(fabsf(fx0[k]), fabsf(fx1[k])));
            My = fmaxf(My, fmaxf(fabsf(fy0[k]), fabsf(fy1[k])));
        }
        for (int e = 0; e < 3; ++e)
        {
            const int i = e, j = (e + 1) % 3, k = (e + 2) % 3;
            const float cix = 0.5f * (fx0[i] + fx1[i]), ciy = 0.5f * (fy0[i] + fy1[i]);
            float nx = 0.5f * (fy0[j] + fy1[j]) - ciy, ny = -(0.5f * (fx0[j] + fx1[j]) - cix);
            if (!(fabsf(nx) + fabsf(ny) > 0.0f)) continue;       // coincident centres or NaN: trivial half-plane
            if (nx * (0.5f * (fx0[k] + fx1[k]) - cix) + ny * (0.5f * (fy0[k] + fy1[k]) - ciy) > 0.0f) { nx = -nx; ny = -ny; }
            float c = -INFINITY;
            for (int v = 0; v < 3; ++v)
            {
                const float px = nx > 0.0f ? fx1[v] : fx0[v], py = ny > 0.0f ? fy1[v] : fy0[v];   // corner maximising nx x + ny y
                c = fmaxf(c, nx * px + ny * py);
            }
            c += 24.0f * 5.9604644775390625e-8f * (fabsf(nx) * Mx + fabsf(ny) * My) + 1e-30f;
            if (!(c == c) || !(fabsf(c) < 1e30f)) continue;
            hull->a[e] = nx; hull->b[e] = ny; hull->c[e] = c;
        }
    }
    FilmRect r;
    r.xlo = round_down_f(xlo - 1e-12 * (1.0 + fabs(xlo))); r.ylo = round_down_f(ylo - 1e-12 * (1.0 + fabs(ylo)));
    r.xhi = round_up_f(xhi + 1e-12 * (1.0 + fabs(xhi))); r.yhi = round_up_f(yhi + 1e-12 * (1.0 + fabs(yhi)));
    return r;
}

// Pixels that can hold a ray whose film point lies in `r` (inclusive range; false when there is none).
//   film_x(x) = fl(fl(fc(x)*h_fov)*aspect), fc(i) = fl(-1 + 2*fl(i/n))   (renderer.cpp:317,329; render.cu: film_coord)
// deviates from the real F(x) = (-1 + 2x/W)*kx, kx = h_fov*aspect (exact product of the two floats), by at most 5.1u|kx|
// (u = 2^-24: one rounding of i/n, one of the sum — both on values <= 1 — and one per product), i.e. by 2.6u*W = 1.6e-7*W
// pixels; film_y by 1.5u*H pixels. A jittered sample (renderer.cpp:340-341: rand_bi*half_pixel_width + film_x, |rand_bi| <= 1)
// moves the film point by at most |half_pixel_width| (1 + u) + u|film_x|. ViewParams::px_slack_x/y = 2e-6*W (+ the jitter
// in pixels) covers both with a factor 10 to spare, so
//        film point of a ray of pixel (x, y) lies in r   ==>   x0 <= x <= x1 and y0 <= y <= y1.
// The inversion runs in double (relative error 1e-15). An unbounded rectangle gives the whole image.
struct PxRect { int x0, x1, y0, y1; };
ATRAY_HD bool film_rect_px(const ViewParams& vp, const FilmRect& r, PxRect& p)
{
    p.x0 = 0; p.x1 = -1; p.y0 = 0; p.y1 = -1;
    if (film_rect_is_empty(r)) return false;
    const double W = (double)vp.W, H = (double)vp.H;
    double xa = ((double)r.xlo * vp.inv_kx + 1.0) * 0.5 * W, xb = ((double)r.xhi * vp.inv_kx + 1.0) * 0.5 * W;
    if (xa > xb) { const double t = xa; xa = xb; xb = t; }
    double ya = ((double)r.ylo + 1.0) * 0.5 * H, yb = ((double)r.yhi + 1.0) * 0.5 * H;
    if (!(xa == xa && xb == xb && ya == ya && yb == yb)) { p.x1 = vp.W - 1; p.y1 = vp.H - 1; return true; }   // NaN (inf*0 …)
    xa -= vp.px_slack_x; xb += vp.px_slack_x; ya -= vp.px_slack_y; yb += vp.px_slack_y;
    if (xb < 0.0 || xa > W - 1.0 || yb < 0.0 || ya > H - 1.0) return false;
    p.x0 = xa <= 0.0 ? 0 : (int)ceil(xa); p.x1 = xb >= W - 1.0 ? vp.W - 1 : (int)floor(xb);
    p.y0 = ya <= 0.0 ? 0 : (int)ceil(ya); p.y1 = yb >= H - 1.0 ? vp.H - 1 : (int)floor(yb);
    return p.x0 <= p.x1 && p.y0 <= p.y1;
}

}  // namespace atray
