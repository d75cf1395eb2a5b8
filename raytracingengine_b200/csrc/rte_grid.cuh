// rte_grid.cuh — per-origin direction grids: the acceleration structure of the FP32 production path and of its exact fix-up.
//
// Every ray of a frame starts at the camera (primary rays) or ends at a light (shadow rays, RE/Scene.h:86-104; "RE/" =
// /root/reference/RaytracingEngine/); a shadow ray is looked up REVERSED, from the light towards the shaded point.  So all
// rays share 1 + nl origins O_j, and the brute-force loop of Scene::IntersectClosest (RE/Scene.h:218-257) over all spheres
// can be replaced, per origin, by a table over ray DIRECTIONS: a cube map around O_j with G x G cells per face; cell (face,
// iu, iv) lists every sphere that a ray from O_j through that cell can touch.  A ray then tests the handful of spheres of ONE
// cell instead of all of them (C2: ~3 instead of 1024) — no shared-memory staging, no per-tile culling, O(1) per ray, and no
// capacity limit (the lists live in global memory and are read through L1 / L2; neighbouring pixels read the same cell).
//
// Conservative by construction (grid_rect below): seen from O_j a sphere is a cone of directions; its bounding rectangle in a
// face's gnomonic coordinates (u, v) = (d_a1, d_a2) / |d_a| is exact — the extreme u is attained on a plane through the
// face's v-axis tangent to the sphere —, is widened by the angular slack of an FP32 direction, and for a light's grid the
// radius is inflated by `bias` because the true shadow ray starts at P + N*bias (RE/Scene.h:104) and passes the light at a
// distance of at most `bias`.  Spheres that contain the origin or subtend more than 30 degrees go to a per-origin "near list"
// that every ray of that origin tests.  A cell may list spheres the ray misses (cost only), never the reverse; the caller
// still decides every candidate itself (rte_render32.cu: FP32 intervals; rte_exact_grid.cuh: the exact FP64 tests).
//
// Layout (CSR): cells of view j start at j * cells_per_view; cell = face * G*G + iv * G + iu; the view's near list is its
// last cell (index 6 G*G).  cell_start[c] .. cell_start[c + 1] index `entries` (unordered primitive indices: spheres first,
// then the triangles — standalone ones and every model's — through their bounding spheres).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rte {

struct DevGrid {
    const uint32_t* cell_start;  // n_views * cells_per_view + 1 offsets
    const uint32_t* entries;     // primitive indices: spheres [0, ns), triangles ns + (flattened triangle index)
    int32_t G;                   // cells per edge of a cube face
    int32_t cells_per_view;      // 6 G^2 + 1
};

// Cell of a unit direction (FP32) in one view's cube map.  |d| = 1 +- 1e-5 is enough (only ratios are used).
__device__ __forceinline__ uint32_t grid_cell(const DevGrid& g, int view, float dx, float dy, float dz) {
    const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    int axis;
    float w, p, q;  // major component, the two others in cyclic order
    if (ax >= ay && ax >= az) { axis = 0; w = dx; p = dy; q = dz; }
    else if (ay >= az)        { axis = 1; w = dy; p = dz; q = dx; }
    else                      { axis = 2; w = dz; p = dx; q = dy; }
    const int face = 2 * axis + (w < 0.0f ? 1 : 0);
    float inv;  // (approximate reciprocal: the rectangles of the build carry a margin of 1e-5 in u, v)
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(fabsf(w)));
    const float half = 0.5f * (float)g.G;
    // (u + 1) * G / 2 with u = p / |w| in [-1, 1]
    int iu = (int)__fmaf_rn(p * inv, half, half), iv = (int)__fmaf_rn(q * inv, half, half);
    iu = min(max(iu, 0), g.G - 1); iv = min(max(iv, 0), g.G - 1);
    return (uint32_t)view * (uint32_t)g.cells_per_view + (uint32_t)(face * g.G * g.G + iv * g.G + iu);
}
__device__ __forceinline__ uint32_t grid_near_cell(const DevGrid& g, int view) {
    return (uint32_t)view * (uint32_t)g.cells_per_view + (uint32_t)(6 * g.G * g.G);
}

// ------------------------------------------------------------------------------ cheap FP64 (no DDIV / DSQRT sequences)
// 1/sqrt(x) to ~1e-15 relative for x within FP32's exponent range: MUFU.RSQ seed (1e-7) + two Newton steps in FP64.
// (Outside that range the result is 0, inf or NaN: callers either cannot get there or treat NaN conservatively.)
__device__ __forceinline__ double rsqrt64_fast(double x) {
    double y = (double)rsqrtf((float)x);
    const double hx = 0.5 * x;
    y = y * fma(-hx * y, y, 1.5);
    y = y * fma(-hx * y, y, 1.5);
    return y;
}
// 1/x to ~1e-15 relative: MUFU.RCP seed + two Newton steps.
__device__ __forceinline__ double rcp64_fast(double x) {
    double y = (double)__frcp_rn((float)x);
    y = y * fma(-x, y, 2.0);
    y = y * fma(-x, y, 2.0);
    return y;
}
__device__ __forceinline__ double sqrt64_fast(double x) { return x > 0.0 ? x * rsqrt64_fast(x) : 0.0; }

// Cells of one face that the sphere {f = C - O (view frame), r} can be seen through: [iu0, iu1] x [iv0, iv1], or false.
// Spheres of the view's near list (grid_sphere_is_near) never get here.  The arithmetic is the cheap FP64 above (a (view,
// primitive) pair used to cost ~700 warp instructions, most of them IEEE division / square-root sequences): its 1e-15 errors
// disappear in the margin of 1e-5; anything that comes out as NaN (coordinates beyond FP32's range) takes the whole face.
struct GridRect { int iu0, iu1, iv0, iv1; };
__device__ __forceinline__ bool grid_sphere_is_near(double fx, double fy, double fz, double r) {
    return !(fx * fx + fy * fy + fz * fz > 4.0 * r * r);  // origin inside, or the cone's half-angle reaches 30 degrees (also NaN)
}
// Straight-line on purpose (selects, no early exits): the six lanes of a warp that work out the six faces of one primitive
// must not diverge — divergent exits cost the warp one pass per face.
__device__ __forceinline__ bool grid_rect(int G, int face, double fx, double fy, double fz, double r, GridRect& out) {
    const int axis = face >> 1;
    const double sgn = (face & 1) ? -1.0 : 1.0;
    // depth along the face normal, the two others in cyclic order (same convention as grid_cell)
    const double w = sgn * (axis == 0 ? fx : axis == 1 ? fy : fz);
    const double x = axis == 0 ? fy : axis == 1 ? fz : fx, y = axis == 0 ? fz : axis == 1 ? fx : fy;
    const double A = w * w - r * r;
    // in front of the plane w = 0 (and not within 1e-6 of touching it: the differences below cancel like w / (w - r), and
    // 1e-15 * 1e6 stays far below the margin): exact bounds of the projected conic,  u = (x w +- r sqrt(x^2 + A)) / A
    const bool front = w > r && A > 1e-6 * w * w;
    const double As = front ? A : 1.0, iA = rcp64_fast(As);
    const double Du = r * sqrt64_fast(x * x + As), Dv = r * sqrt64_fast(y * y + As);
    // ... unless it lies outside the face's pyramid |x| <= w, |y| <= w, beyond one of its four planes through the origin
    // (distance (x - w) / sqrt 2)
    const double k = r * 1.4142136;
    const bool outside = x - w > k || -x - w > k || y - w > k || -y - w > k;
    // behind, straddling or nearly touching w = 0: the cone (half-angle < 30 deg) reaches the face's directions (within 54.74
    // deg of its normal) only if angle(axis, normal) - theta < 54.74 deg; then it may touch any cell of the face
    const double iF = rsqrt64_fast(fx * fx + fy * fy + fz * fz), ca = w * iF, st = r * iF;
    const double ct = sqrt64_fast(1.0 - st * st), sa = sqrt64_fast(1.0 - ca * ca);
    const bool reach = !(ca * ct + sa * st <= 0.5773);  // (NaN: reaches)
    const bool any = front ? !outside : (reach && !(w < -r));
    double u0 = front ? (x * w - Du) * iA : -1.0, u1 = front ? (x * w + Du) * iA : 1.0;
    double v0 = front ? (y * w - Dv) * iA : -1.0, v1 = front ? (y * w + Dv) * iA : 1.0;
    const double m = 1e-5;  // FP32 direction of the ray + rounding of its cell arithmetic, in units of u
    u0 -= m * (1.0 + fabs(u0)); u1 += m * (1.0 + fabs(u1)); v0 -= m * (1.0 + fabs(v0)); v1 += m * (1.0 + fabs(v1));
    const bool nan = !(u0 == u0 && u1 == u1 && v0 == v0 && v1 == v1);  // coordinates beyond FP32's range: the whole face
    u0 = nan ? -1.0 : u0; u1 = nan ? 1.0 : u1; v0 = nan ? -1.0 : v0; v1 = nan ? 1.0 : v1;
    const bool seen = any && !(u1 < -1.0 || u0 > 1.0 || v1 < -1.0 || v0 > 1.0);
    const double half = 0.5 * (double)G;
    out.iu0 = max(0, min(G - 1, (int)floor((fmax(u0, -1.0) + 1.0) * half - 1e-3)));
    out.iu1 = max(0, min(G - 1, (int)floor((fmin(u1, 1.0) + 1.0) * half + 1e-3)));
    out.iv0 = max(0, min(G - 1, (int)floor((fmax(v0, -1.0) + 1.0) * half - 1e-3)));
    out.iv1 = max(0, min(G - 1, (int)floor((fmin(v1, 1.0) + 1.0) * half + 1e-3)));
    return seen;
}

}  // namespace rte
