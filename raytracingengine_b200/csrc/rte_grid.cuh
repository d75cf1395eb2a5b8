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

// Cells of one face that the sphere {f = C - O (view frame), r} can be seen through: [iu0, iu1] x [iv0, iv1], or false.
// `near` is set when the sphere belongs to the view's near list instead (then every face returns false).
struct GridRect { int iu0, iu1, iv0, iv1; };
__device__ __forceinline__ bool grid_sphere_is_near(double fx, double fy, double fz, double r) {
    return !(fx * fx + fy * fy + fz * fz > 4.0 * r * r);  // origin inside, or the cone's half-angle reaches 30 degrees (also NaN)
}
__device__ __forceinline__ bool grid_rect(int G, int face, double fx, double fy, double fz, double r, GridRect& out) {
    const int axis = face >> 1;
    const double sgn = (face & 1) ? -1.0 : 1.0;
    double w, x, y;  // depth along the face normal, the two others in cyclic order (same convention as grid_cell)
    if (axis == 0) { w = sgn * fx; x = fy; y = fz; }
    else if (axis == 1) { w = sgn * fy; x = fz; y = fx; }
    else { w = sgn * fz; x = fx; y = fy; }
    double u0 = -1.0, u1 = 1.0, v0 = -1.0, v1 = 1.0;
    if (w > r) {  // entirely in front of the plane w = 0: exact bounds of the projected conic
        const double A = w * w - r * r;
        const double Du = r * sqrt(fmax(x * x + A, 0.0)), Dv = r * sqrt(fmax(y * y + A, 0.0));
        u0 = (x * w - Du) / A; u1 = (x * w + Du) / A;
        v0 = (y * w - Dv) / A; v1 = (y * w + Dv) / A;
    } else {
        // straddles w = 0: the cone (half-angle < 30 deg) reaches the face's directions (within 54.74 deg of its normal) only
        // if angle(axis, normal) - theta < 54.74 deg; then it may touch any cell of the face
        const double F = sqrt(fx * fx + fy * fy + fz * fz), ca = w / F, st = r / F;
        const double ct = sqrt(fmax(1.0 - st * st, 0.0)), sa = sqrt(fmax(1.0 - ca * ca, 0.0));
        if (!(ca * ct + sa * st > 0.5773)) return false;
    }
    const double m = 1e-5;  // FP32 direction of the ray + rounding of its cell arithmetic, in units of u
    u0 -= m * (1.0 + fabs(u0)); u1 += m * (1.0 + fabs(u1)); v0 -= m * (1.0 + fabs(v0)); v1 += m * (1.0 + fabs(v1));
    if (!(u1 >= -1.0 && u0 <= 1.0 && v1 >= -1.0 && v0 <= 1.0)) return false;
    const double half = 0.5 * (double)G;
    out.iu0 = max(0, min(G - 1, (int)floor((fmax(u0, -1.0) + 1.0) * half - 1e-3)));
    out.iu1 = max(0, min(G - 1, (int)floor((fmin(u1, 1.0) + 1.0) * half + 1e-3)));
    out.iv0 = max(0, min(G - 1, (int)floor((fmax(v0, -1.0) + 1.0) * half - 1e-3)));
    out.iv1 = max(0, min(G - 1, (int)floor((fmin(v1, 1.0) + 1.0) * half + 1e-3)));
    return true;
}

}  // namespace rte
