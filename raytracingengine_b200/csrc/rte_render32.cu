// rte_render32.cu — the FP32 production path of the render call (sm_100a): diffuse scenes (spheres, planes, triangles, models) over per-origin
// direction grids (rte_grid.cuh; built here), outputs HDR float4 / RGB8 (+ the integer AOVs prim_id, light_mask).
//
// Replaces Scene::RenderImage / TraceRay / directLightning + tonemap()/toColor for those scenes (RE/Scene.h:311-328, 131-198,
// 79-129; RE/RaytracingEngine.cpp:113-121,165-174; "RE/" = /root/reference/RaytracingEngine/) to the bar BASELINE.json states:
// nearest-hit index and shadow visibility bit-exact, depth / normals 1e-5, RGB8 +-1 LSB.  (The FP64 path of rte_render.cu,
// bit-identical to the reference, stays behind RTE_FLAG_EXACT_FP64 and every FP64 output.)
//
//  * Values (hit point, normal, N.L, 1/d^2, the light sum, the operator) are FP32; the few places where FP32 would cancel
//    (hit point, C - origin of a candidate) take a handful of FP64 adds / FMAs — never an FP64 division or square root
//    (those are ~25 / ~40 SASS instructions each): reciprocal square roots come from MUFU + Newton steps.
//  * DECISIONS (which primitive is nearest, is a shadow ray cast, is it occluded) are taken from FP32 INTERVALS with explicit
//    error bounds.  A pixel with any inconclusive decision is queued by its warp and rendered again with the reference's
//    exact FP64 operations (fix_batch, rte_exact_grid.cuh).  So every decision that reaches an output is either certain
//    or exact, and those pixels (~1e-3 of a frame) are the reference's own bits.
//  * The operator + toColor run in the kernel (tm32, rte_tonemap.cuh) and a tile row of 8 pixels leaves as six aligned
//    32-bit words: no FP64 HDR scratch, no second pass.
#include <algorithm>

#include "rte_exact_grid.cuh"
#include "rte_grid.cuh"
#include "rte_interval.cuh"
#include "rte_kernels.h"
#include "rte_tonemap.cuh"
#include "rte_trace.cuh"

namespace rte {

constexpr int kR32Threads = 1024;  // one CTA per SM at 64 registers (768 / 512 threads with more registers measured slower)
constexpr int kR32Planes = 8, kR32Lights = 16;  // scenes with more take the exact path

// (cheap FP64 reciprocals / square roots without DDIV / DSQRT sequences: rsqrt64_fast, rcp64_fast in rte_grid.cuh)

// ------------------------------------------------------------------------------ scene constants in shared memory
struct Scene32 {
    Plane32 pl[kR32Planes];
    float li_hi[kR32Lights][3], li_lo[kR32Lights][3];  // light position = hi + lo
    float li_E[kR32Lights][3];
    float plA[kR32Lights][kR32Planes];  // n_i . Lp_j - d_i: the part of (n_i . L) * dist that does not depend on the pixel
    double cam[3];
    float cam_hi[3], cam_lo[3];
    int np, nl;
};

// ------------------------------------------------------------------------------ building the direction grids (rte_grid.cuh)
// Per frame (camera, lights or bias changed): count the (sphere, cell) incidences, exclusive scan, fill.  One thread per
// (view, sphere); the count pass leaves its counters at zero again after the fill pass has consumed them.
struct GridBuildParams {
    DevScene sc;
    double cam[3];
    double bias;
    int32_t G, cells_per_view, n_views;
    uint32_t* count;       // n_cells counters (zero before the count pass and after the fill pass)
    uint32_t* cell_start;  // n_cells + 1
    uint32_t* entries;
    uint32_t capacity;     // of `entries`
};
// Primitive i of the grid: spheres [0, ns), then the triangles [ns, ns + nt) through their bounding spheres (centroid, farthest
// vertex: a ray that hits the triangle hits that sphere).
__device__ __forceinline__ void grid_view_prim(const GridBuildParams& B, int j, int i, double& fx, double& fy, double& fz, double& r) {
    double cx, cy, cz, rad;
    if (i < B.sc.ns) {
        const double4 s = B.sc.sph64[i];
        cx = s.x; cy = s.y; cz = s.z; rad = fabs(s.w);
    } else {
        const DevTri64 T = B.sc.tri64[i - B.sc.ns];
        const double ax = T.a0[0], ay = T.a0[1], az = T.a0[2];
        const double bx = ax + T.e1[0], by = ay + T.e1[1], bz = az + T.e1[2], qx = ax + T.e2[0], qy = ay + T.e2[1], qz = az + T.e2[2];
        cx = (ax + bx + qx) * (1.0 / 3.0); cy = (ay + by + qy) * (1.0 / 3.0); cz = (az + bz + qz) * (1.0 / 3.0);
        const double da = (ax - cx) * (ax - cx) + (ay - cy) * (ay - cy) + (az - cz) * (az - cz);
        const double db = (bx - cx) * (bx - cx) + (by - cy) * (by - cy) + (bz - cz) * (bz - cz);
        const double dc = (qx - cx) * (qx - cx) + (qy - cy) * (qy - cy) + (qz - cz) * (qz - cz);
        rad = sqrt64_fast(fmax(da, fmax(db, dc))) * (1.0 + 1e-9) + 1e-300;
    }
    const double ox = j == 0 ? B.cam[0] : B.sc.lights[j - 1].pos[0], oy = j == 0 ? B.cam[1] : B.sc.lights[j - 1].pos[1],
                 oz = j == 0 ? B.cam[2] : B.sc.lights[j - 1].pos[2];
    fx = cx - ox; fy = cy - oy; fz = cz - oz;
    const double F = sqrt64_fast(fx * fx + fy * fy + fz * fz);  // (only scales the slack below)
    // radius (+ bias for a light's grid: the shadow line passes the light within `bias`), + the lateral reach of an FP32 direction
    r = (rad + (j > 0 ? B.bias : 0.0)) * (1.0 + 1e-9) + 4e-6 * F + 1e-12;
}
// One WARP per (view, primitive).  Lane f < 6 works out face f's rectangle (the FP64 square roots and divisions once per face);
// the warp walks small rectangles together, rectangles of more than kGridBigRect cells (a sphere close to a light covers
// thousands) are handed to the whole block, and every thread keeps several atomics in flight: the fill pass waits for each
// atomic's return value, and one warp walking such a rectangle alone was a 0.4 ms tail on C5.
constexpr int kGridBuildThreads = 256, kGridBuildWarps = kGridBuildThreads / 32, kGridBigRect = 128;
template <bool kFill>
__global__ void __launch_bounds__(kGridBuildThreads) grid_build_kernel(const GridBuildParams B) {
    __shared__ int s_big[kGridBuildWarps * 6][6];  // face base cell, iu0, iv0, width, cells, primitive
    __shared__ int s_nbig;
    const int n_prims = B.sc.ns + B.sc.nt, total = B.n_views * n_prims, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int GG = B.G * B.G;
    // cells [c0, c0 + step * kUnroll) of a rectangle, `step` apart: the atomics are independent, the stores follow
    auto add_cells = [&](uint32_t face_base, int iu0, int iv0, int w, int n, int prim, int c0, int step) {
        constexpr int kUnroll = 4;
        const float inv_w = 1.0f / (float)w;
        for (int c = c0; c < n; c += step * kUnroll) {
            uint32_t cell[kUnroll], pos[kUnroll];
#pragma unroll
            for (int k = 0; k < kUnroll; ++k) {
                const int cc = c + k * step;
                int row = (int)((float)cc * inv_w), col = cc - row * w;  // cc / w without the integer-division sequence (n <= 512^2)
                if (col < 0) { --row; col += w; } else if (col >= w) { ++row; col -= w; }
                cell[k] = cc < n ? face_base + (uint32_t)((iv0 + row) * B.G + iu0 + col) : 0xFFFFFFFFu;
            }
            if (!kFill) {
#pragma unroll
                for (int k = 0; k < kUnroll; ++k) if (cell[k] != 0xFFFFFFFFu) atomicAdd(B.count + cell[k], 1u);
            } else {
#pragma unroll
                for (int k = 0; k < kUnroll; ++k) if (cell[k] != 0xFFFFFFFFu) pos[k] = B.cell_start[cell[k]] + (atomicSub(B.count + cell[k], 1u) - 1u);
#pragma unroll
                for (int k = 0; k < kUnroll; ++k) if (cell[k] != 0xFFFFFFFFu && pos[k] < B.capacity) B.entries[pos[k]] = (uint32_t)prim;
            }
        }
    };
    for (int t0 = blockIdx.x * kGridBuildWarps; t0 < total; t0 += gridDim.x * kGridBuildWarps) {  // (block-uniform)
        if (threadIdx.x == 0) s_nbig = 0;
        __syncthreads();
        const int t = t0 + warp;
        if (t < total) {
            const int j = t / n_prims, i = t - j * n_prims;
            double fx, fy, fz, r;
            grid_view_prim(B, j, i, fx, fy, fz, r);
            const uint32_t base = (uint32_t)j * (uint32_t)B.cells_per_view;
            if (grid_sphere_is_near(fx, fy, fz, r)) {
                if (lane == 0) add_cells(base + (uint32_t)(6 * GG), 0, 0, 1, 1, i, 0, 1);
            } else {
                GridRect mine;
                mine.iu0 = mine.iu1 = mine.iv0 = mine.iv1 = 0;
                const bool have = lane < 6 && grid_rect(B.G, lane, fx, fy, fz, r, mine);
                const unsigned faces = __ballot_sync(0xFFFFFFFFu, have);
                for (int face = 0; face < 6; ++face) {
                    if (!(faces >> face & 1u)) continue;
                    const int iu0 = __shfl_sync(0xFFFFFFFFu, mine.iu0, face), iu1 = __shfl_sync(0xFFFFFFFFu, mine.iu1, face);
                    const int iv0 = __shfl_sync(0xFFFFFFFFu, mine.iv0, face), iv1 = __shfl_sync(0xFFFFFFFFu, mine.iv1, face);
                    const int w = iu1 - iu0 + 1, n = w * (iv1 - iv0 + 1);
                    const uint32_t face_base = base + (uint32_t)(face * GG);
                    if (n <= kGridBigRect) add_cells(face_base, iu0, iv0, w, n, i, lane, 32);
                    else if (lane == 0) {
                        int* e = s_big[atomicAdd(&s_nbig, 1)];
                        e[0] = (int)face_base; e[1] = iu0; e[2] = iv0; e[3] = w; e[4] = n; e[5] = i;
                    }
                }
            }
        }
        __syncthreads();
        const int nbig = s_nbig;
        for (int k = 0; k < nbig; ++k)
            add_cells((uint32_t)s_big[k][0], s_big[k][1], s_big[k][2], s_big[k][3], s_big[k][4], s_big[k][5], threadIdx.x, kGridBuildThreads);
        __syncthreads();  // (s_big / s_nbig are rewritten by the next round)
    }
}

// Exclusive scan of `count` into `cell_start`, cell_start[n] = total.  Block b scans cells [b, b + 1) * kScanChunk locally,
// publishes its total tagged with the build's epoch (no reset between builds), adds the totals of the blocks before it
// (they all run at the same time: the grid is far smaller than one wave) and writes its offsets.
constexpr int kScanThreads = 256, kScanPer = 16, kScanChunk = kScanThreads * kScanPer;
__global__ void __launch_bounds__(kScanThreads) grid_scan_kernel(const uint32_t* __restrict__ count, uint32_t* __restrict__ cell_start, int n,
                                                                 unsigned long long* aggregates, uint32_t epoch, uint32_t* total_host) {
    __shared__ uint32_t s_warp[kScanThreads / 32];
    __shared__ uint32_t s_prefix;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i0 = blockIdx.x * kScanChunk + threadIdx.x * kScanPer;
    uint32_t v[kScanPer], sum = 0;
#pragma unroll
    for (int k = 0; k < kScanPer; k += 4) {
        uint4 q = make_uint4(0u, 0u, 0u, 0u);
        if (i0 + k + 3 < n) q = *reinterpret_cast<const uint4*>(count + i0 + k);
        else { if (i0 + k < n) q.x = count[i0 + k]; if (i0 + k + 1 < n) q.y = count[i0 + k + 1]; if (i0 + k + 2 < n) q.z = count[i0 + k + 2]; }
        v[k] = q.x; v[k + 1] = q.y; v[k + 2] = q.z; v[k + 3] = q.w;
        sum += q.x + q.y + q.z + q.w;
    }
    uint32_t inc = sum;  // inclusive scan of the per-thread sums: warp shuffles, then the warp totals
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, btotal = 0;
#pragma unroll
    for (int k = 0; k < kScanThreads / 32; ++k) { const uint32_t t = s_warp[k]; if (k < warp) woff += t; btotal += t; }
    if (threadIdx.x == 0) {
        const unsigned long long tagged = ((unsigned long long)epoch << 32) | btotal;
        asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(aggregates + blockIdx.x), "l"(tagged) : "memory");
        s_prefix = 0u;
    }
    __syncthreads();
    uint32_t part = 0;  // totals of the blocks before this one
    for (int k = threadIdx.x; k < (int)blockIdx.x; k += kScanThreads) {
        unsigned long long a;
        do {  // (volatile: a plain load may legally be read once and the loop assumed to end)
            asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(a) : "l"(aggregates + k) : "memory");
        } while ((uint32_t)(a >> 32) != epoch);
        part += (uint32_t)a;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
    if (lane == 0 && part) atomicAdd(&s_prefix, part);
    __syncthreads();
    uint32_t run = s_prefix + woff + inc - sum;
#pragma unroll
    for (int k = 0; k < kScanPer; ++k) { if (i0 + k < n) cell_start[i0 + k] = run; run += v[k]; }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kScanThreads - 1) {
        cell_start[n] = s_prefix + btotal;
        if (total_host) *total_host = s_prefix + btotal;  // pinned host word (unified addressing): no copy node on the stream
    }
}

static GridBuildParams grid_params(const DevScene& sc, const DevCamera& cam, double bias, int G, uint32_t* count, uint32_t* cell_start, uint32_t* entries,
                                   uint32_t capacity) {
    GridBuildParams B{};
    B.sc = sc; B.cam[0] = cam.pos[0]; B.cam[1] = cam.pos[1]; B.cam[2] = cam.pos[2]; B.bias = bias;
    B.G = G; B.cells_per_view = 6 * G * G + 1; B.n_views = 1 + sc.nl;
    B.count = count; B.cell_start = cell_start; B.entries = entries; B.capacity = capacity;
    return B;
}
int grid_scan_blocks(int n_cells) { return (n_cells + kScanChunk - 1) / kScanChunk; }
cudaError_t launch_grid_count_scan(const DevScene& sc, const DevCamera& cam, double bias, int G, uint32_t* count, uint32_t* cell_start,
                                   unsigned long long* aggregates, uint32_t epoch, uint32_t* total_host, cudaStream_t stream) {
    const GridBuildParams B = grid_params(sc, cam, bias, G, count, cell_start, nullptr, 0);
    const long long total = (long long)B.n_views * (sc.ns + sc.nt);
    const int n_cells = B.n_views * B.cells_per_view;
    if (total > 0) grid_build_kernel<false><<<(unsigned)std::min<long long>((total + 7) / 8, 1 << 20), 256, 0, stream>>>(B);
    grid_scan_kernel<<<grid_scan_blocks(n_cells), kScanThreads, 0, stream>>>(count, cell_start, n_cells, aggregates, epoch, total_host);
    return cudaGetLastError();
}
cudaError_t launch_grid_fill(const DevScene& sc, const DevCamera& cam, double bias, int G, uint32_t* count, uint32_t* cell_start, uint32_t* entries,
                             uint32_t capacity, cudaStream_t stream) {
    const GridBuildParams B = grid_params(sc, cam, bias, G, count, cell_start, entries, capacity);
    const long long total = (long long)B.n_views * (sc.ns + sc.nt);
    if (total > 0) grid_build_kernel<true><<<(unsigned)std::min<long long>((total + 7) / 8, 1 << 20), 256, 0, stream>>>(B);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------ the candidates of one ray
// The spheres of the ray's cell, then the view's near list.  `on_cand(idx)` may clear `go` to stop early.
template <class Fn>
__device__ __forceinline__ void grid_candidates(const DevGrid& g, int view, float dx, float dy, float dz, bool& go, Fn&& on_cand) {
    if (!go) return;
    uint32_t cell = grid_cell(g, view, dx, dy, dz);
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const uint32_t e0 = __ldg(g.cell_start + cell), e1 = __ldg(g.cell_start + cell + 1);
#pragma unroll 1
        for (uint32_t e = e0; e < e1 && go; ++e) on_cand((int)__ldg(g.entries + e));
        cell = grid_near_cell(g, view);
    }
}

// ------------------------------------------------------------------------------ operator + toColor for one pixel
__device__ __forceinline__ uint32_t tonemap_rgb24(int op, float x, float y, float z, const tm32::Consts& K) {
    const tm32::Shared s = tm32::prepare<0u>(tm32::mk3(x, y, z), 1u << op);
    tm32::f3 c;
    switch (op) {
        case RTE_TM_SIMPLE: c = tm32::apply_op<RTE_TM_SIMPLE>(s, K); break;
        case RTE_TM_REINHARD_SIMPLE: c = tm32::apply_op<RTE_TM_REINHARD_SIMPLE>(s, K); break;
        case RTE_TM_REINHARD_EXTENDED: c = tm32::apply_op<RTE_TM_REINHARD_EXTENDED>(s, K); break;
        case RTE_TM_REINHARD_EXTENDED_LUMINANCE: c = tm32::apply_op<RTE_TM_REINHARD_EXTENDED_LUMINANCE>(s, K); break;
        case RTE_TM_REINHARD_JODIE: c = tm32::apply_op<RTE_TM_REINHARD_JODIE>(s, K); break;
        case RTE_TM_UNCHARTED2: c = tm32::apply_op<RTE_TM_UNCHARTED2>(s, K); break;
        default: c = tm32::apply_op<RTE_TM_ACES>(s, K); break;
    }
    return (tm32::qbyte(c.x) & 255u) | ((tm32::qbyte(c.y) & 255u) << 8) | ((tm32::qbyte(c.z) & 255u) << 16);
}

// One sample of one pixel per lane, all 32 lanes of the warp in lock step.  Returns false when some decision of the pixel
// was inconclusive (the pixel then goes to the warp's exact queue, fix_batch, and nothing else of the result is meaningful).
struct Px32 {
    float r, g, b;       // linear HDR of the sample
    uint32_t prim, lmask;
    uint32_t c_shadow, c_sph, c_pl, c_tri;  // statistics of this sample (SURVEY §8(d) algorithmic counts)
};
template <bool kTris>  // scenes with triangles: a second instance, so that sphere scenes keep their 64 registers without spills (C2 0.85 -> 0.98 ms otherwise)
__device__ __forceinline__ bool trace32(const RenderParams& P, const Scene32& S, float* __restrict__ pl_num, int px, int py, bool alive, int sample, Px32& out,
                                        uint32_t& executed) {
    const DevScene& sc = P.sc;
    const DevGrid& grid = P.grid;
    const int ns = sc.ns, nt = kTris ? sc.nt : 0, np = S.np, nl = S.nl;
    const float bias = (float)P.bias;
    bool ok = true;
    out.r = out.g = out.b = 0.0f;
    out.prim = kMiss; out.lmask = 0u; out.c_shadow = 0u; out.c_sph = (uint32_t)ns; out.c_pl = (uint32_t)np; out.c_tri = (uint32_t)nt;
    // ---- Camera::getRay, RE/Math.h:99-121 (the reference's FP64 operations up to the normalisation)
    double dx, dy, dz;
    {
        double sx = (double)px - P.cam.half_w, sy = P.cam.half_h - (double)py;
        if (sample > 0) {
            const int64_t pixel = (int64_t)py * P.cam.width + px;
            sx += jitter01(P.cam.jitter_seed, pixel, sample, 0);
            sy += jitter01(P.cam.jitter_seed, pixel, sample, 1);
        }
        const double vx = sx - S.cam[0], vy = sy - S.cam[1], vz = (S.cam[2] + P.cam.focal) - S.cam[2];
        const double len2 = vx * vx + vy * vy + vz * vz;
        if (!(len2 > 1e-20) || !(len2 < 1e30)) ok = false;  // (normalize() returns 0 below 1e-12: leave it to the exact code)
        const double inv = rsqrt64_fast(len2);
        dx = vx * inv; dy = vy * inv; dz = vz * inv;
    }
    struct { float ux, uy, uz; } r;  // fl32 of the unit direction
    r.ux = (float)dx; r.uy = (float)dy; r.uz = (float)dz;
    // ---- Scene::IntersectClosest, RE/Scene.h:218-257, from intervals
    Near32 nr;
    nr.kind = RTE_HIT_NONE; nr.idx = 0; nr.lo = 0.0f; nr.hi = 0.0f; nr.amb_lo = INFINITY;
    const float cam32x = S.cam_hi[0], cam32y = S.cam_hi[1], cam32z = S.cam_hi[2];
    const float cam1 = fabsf(cam32x) + fabsf(cam32y) + fabsf(cam32z);
    bool go0 = alive;
    grid_candidates(grid, 0, r.ux, r.uy, r.uz, go0, [&](int idx) {
        executed += 1u;
        if (kTris && idx >= ns) {  // a triangle (standalone or a model's): the cell lists hold it through its bounding sphere
            float lo, hi;
            const int cls = tri_interval(sc.tri_iv[idx - ns], cam32x, cam32y, cam32z, S.cam_lo[0], S.cam_lo[1], S.cam_lo[2], r.ux, r.uy, r.uz, 0.0f, 4.0f * kEps32, lo, hi);
            if (cls == kCandHit) near_hit(nr, RTE_HIT_TRIANGLE, idx - ns, lo, hi);
            else if (cls == kCandUnsure) near_unsure(nr, lo);
            return;
        }
        const float4 shi = __ldg(sc.sph_hl + 2 * idx);
        if (far_miss(shi, cam32x, cam32y, cam32z, cam1, r.ux, r.uy, r.uz)) return;
        const Cand32 c = classify_sphere(shi, __ldg(sc.sph_hl + 2 * idx + 1), cam32x, cam32y, cam32z, S.cam_lo[0], S.cam_lo[1], S.cam_lo[2], r.ux, r.uy, r.uz);
        float lo, hi;
        const int cls = reported_t(c, lo, hi);
        if (cls == kCandHit) near_hit(nr, RTE_HIT_SPHERE, idx, lo, hi);
        else if (cls == kCandUnsure) near_unsure(nr, lo);
    });
#pragma unroll 1
    for (int i = 0; i < np; ++i) {  // planes after the spheres: a plane wins only when it is certainly closer
        float lo, hi;
        const int cls = plane_interval(S.pl[i], cam32x, cam32y, cam32z, cam1, r.ux, r.uy, r.uz, lo, hi);
        if (cls == kCandHit) near_hit(nr, RTE_HIT_PLANE, i, lo, hi);
        else if (cls == kCandUnsure) near_unsure(nr, 0.0f);
    }
    if (near_ambiguous(nr)) ok = false;
    const bool hitv = alive && ok && nr.kind != RTE_HIT_NONE;
    if (!alive) { out.c_sph = 0u; out.c_pl = 0u; out.c_tri = 0u; }
    if (alive && nr.kind == RTE_HIT_NONE) {  // Scene::backgroundColor, RE/Scene.h:30-33
        const float t = 0.5f * (r.uy + 1.0f);
        out.r = __fmaf_rn(0.5f, t, 1.0f - t); out.g = __fmaf_rn(0.7f, t, 1.0f - t); out.b = 1.0f;
    }
    // ---- the hit: distance in cheap FP64 (no DDIV / DSQRT), point, viewer-facing unit normal
    double Px = 0.0, Py = 0.0, Pz = 0.0;
    float Nx = 0.0f, Ny = 0.0f, Nz = 1.0f;
    int mat = sc.nm, self = -1;
    if (hitv && nr.kind == RTE_HIT_SPHERE) {
        const double4 s = sc.sph64[nr.idx];
        const double ocx = S.cam[0] - s.x, ocy = S.cam[1] - s.y, ocz = S.cam[2] - s.z;
        const double bq = ocx * dx + ocy * dy + ocz * dz;            // oc.d, |d| = 1
        const double cq = fma(ocx, ocx, fma(ocy, ocy, ocz * ocz)) - s.w * s.w;
        const double disc = fma(bq, bq, -cq);
        const bool far_root = cq < 0.0;                              // origin inside the sphere: the far root was reported
        const double sq = disc > 0.0 ? disc * rsqrt64_fast(disc) : 0.0;
        const double t = far_root ? -bq + sq : -bq - sq;
        Px = fma(dx, t, S.cam[0]); Py = fma(dy, t, S.cam[1]); Pz = fma(dz, t, S.cam[2]);
        const float ir = __frcp_rn((float)fabs(s.w));
        float nx = (float)(Px - s.x) * ir, ny = (float)(Py - s.y) * ir, nz = (float)(Pz - s.z) * ir;  // outward
        if (far_root) { nx = -nx; ny = -ny; nz = -nz; } else if (bias >= 1e-5f) self = nr.idx;       // frontFace = n.d < 0
        Nx = nx; Ny = ny; Nz = nz;
        mat = sc.sph_mat[nr.idx];
        out.prim = ((uint32_t)RTE_HIT_SPHERE << 30) | (uint32_t)nr.idx;
    } else if (kTris && hitv && nr.kind == RTE_HIT_TRIANGLE) {
        const DevTri64& T = sc.tri64[nr.idx];  // Moeller-Trumbore's t (RE/Shape.h:208-218) in cheap FP64
        const double hx = dy * T.e2[2] - dz * T.e2[1], hy = dz * T.e2[0] - dx * T.e2[2], hz = dx * T.e2[1] - dy * T.e2[0];
        const double a = T.e1[0] * hx + T.e1[1] * hy + T.e1[2] * hz;
        const double sx = S.cam[0] - T.a0[0], sy = S.cam[1] - T.a0[1], sz = S.cam[2] - T.a0[2];
        const double qx = sy * T.e1[2] - sz * T.e1[1], qy = sz * T.e1[0] - sx * T.e1[2], qz = sx * T.e1[1] - sy * T.e1[0];
        const double t = (T.e2[0] * qx + T.e2[1] * qy + T.e2[2] * qz) * rcp64_fast(a);
        Px = fma(dx, t, S.cam[0]); Py = fma(dy, t, S.cam[1]); Pz = fma(dz, t, S.cam[2]);
        const double nd = T.n[0] * dx + T.n[1] * dy + T.n[2] * dz;
        if (fabs(nd) < 1e-9) ok = false;            // frontFace = n.d < 0 must be certain
        const float sg = nd < 0.0 ? 1.0f : -1.0f;
        Nx = sg * (float)T.n[0]; Ny = sg * (float)T.n[1]; Nz = sg * (float)T.n[2];
        mat = sc.tri_info[nr.idx].x;
        out.prim = ((uint32_t)RTE_HIT_TRIANGLE << 30) | (uint32_t)sc.tri_info[nr.idx].y;
    } else if (hitv) {
        const DevPlane& pl = sc.planes[nr.idx];
        const double num = (pl.p0[0] - S.cam[0]) * pl.n[0] + (pl.p0[1] - S.cam[1]) * pl.n[1] + (pl.p0[2] - S.cam[2]) * pl.n[2];
        const double den = pl.n[0] * dx + pl.n[1] * dy + pl.n[2] * dz;
        const double t = num * rcp64_fast(den);
        Px = fma(dx, t, S.cam[0]); Py = fma(dy, t, S.cam[1]); Pz = fma(dz, t, S.cam[2]);
        const float sg = den < 0.0 ? 1.0f : -1.0f;  // frontFace = n.d < 0 (|n.d| > 1e-6 is certain here)
        Nx = sg * S.pl[nr.idx].nx; Ny = sg * S.pl[nr.idx].ny; Nz = sg * S.pl[nr.idx].nz;
        mat = S.pl[nr.idx].mat;
        out.prim = ((uint32_t)RTE_HIT_PLANE << 30) | (uint32_t)nr.idx;
    }
    mat = (mat >= 0 && mat < sc.nm) ? mat : sc.nm;
    // hit point as hi + lo floats; shadow origin so = P + N * bias (RE/Scene.h:104) as hi + lo floats through an error-free FP32 sum
    const float Phx = (float)Px, Phy = (float)Py, Phz = (float)Pz;
    const float Plx = (float)(Px - (double)Phx), Ply = (float)(Py - (double)Phy), Plz = (float)(Pz - (double)Phz);
    float so32x, so32y, so32z, solx, soly, solz;
    {
        const float bx = Nx * bias, by = Ny * bias, bz = Nz * bias;
        so32x = Phx + bx; so32y = Phy + by; so32z = Phz + bz;
        // TwoSum: s = a + b exactly equals fl(s) + err
        const float tx = so32x - Phx, ty = so32y - Phy, tz = so32z - Phz;
        solx = Plx + ((Phx - (so32x - tx)) + (bx - tx)); soly = Ply + ((Phy - (so32y - ty)) + (by - ty)); solz = Plz + ((Phz - (so32z - tz)) + (bz - tz));
    }
    const float so1 = fabsf(so32x) + fabsf(so32y) + fabsf(so32z);
    // num_i = d_i - n_i . so of every plane, once per pixel (this thread's slots of pl_num: [plane][thread])
#pragma unroll 1
    for (int i = 0; i < np; ++i)
        pl_num[i * kR32Threads] = S.pl[i].d - __fmaf_rn(S.pl[i].nz, so32z, __fmaf_rn(S.pl[i].ny, so32y, S.pl[i].nx * so32x));
    float ar = 0.0f, ag = 0.0f, ab = 0.0f;
    // ---- Scene::directLightning, RE/Scene.h:79-129 (diffuse term)
#pragma unroll 1
    for (int j = 0; j < nl; ++j) {
        const float vx = (S.li_hi[j][0] - Phx) + (S.li_lo[j][0] - Plx), vy = (S.li_hi[j][1] - Phy) + (S.li_lo[j][1] - Ply), vz = (S.li_hi[j][2] - Phz) + (S.li_lo[j][2] - Plz);
        const float d2 = __fmaf_rn(vz, vz, __fmaf_rn(vy, vy, vx * vx));
        float inv = rsqrtf(d2);
        inv = inv * __fmaf_rn(-0.5f * d2 * inv, inv, 1.5f);  // one Newton step: <= 1.5 ulp
        const float dist = d2 * inv;
        const float Lx = vx * inv, Ly = vy * inv, Lz = vz * inv;
        const float ndl = __fmaf_rn(Nz, Lz, __fmaf_rn(Ny, Ly, Nx * Lx));
        bool cast = hitv && ok && d2 > 0.0f;
        if (cast) {
            // guards of RE/Scene.h:89-101: ndl > 0, dist > bias — certain, or the pixel is flagged
            if (fabsf(ndl) < 32.0f * kEps32 || fabsf(dist - bias) < 1e-5f * bias + 1e-30f) { ok = false; cast = false; }
            else cast = ndl > 0.0f && dist > bias;
        }
        if (!__any_sync(kFullMask, cast)) continue;
        const float maxDist = dist - bias;
        const float dt = 24.0f * kEps32 * (dist + so1) + 1e-7f;  // slack of the range ends (maxDist, bias) against the candidate bounds
        const float tmax_lo = maxDist - dt, tmax_hi = maxDist + dt, bias_hi = bias + dt;
        bool occl = false, unsure = false, occl_t = false;
        uint32_t first = (uint32_t)ns, first_t = (uint32_t)nt;
        bool go = cast;
        // the light's grid is looked up along the reversed ray (from the light towards the shaded point); the candidates
        // are tested against the ray itself: origin so, direction L
        grid_candidates(grid, 1 + j, -Lx, -Ly, -Lz, go, [&](int idx) {
            if (idx == self) return;  // so lies outside the sphere the point is on, the ray leaves it: both roots < 0
            executed += 1u;
            if (kTris && idx >= ns) {  // a triangle
                float lo, hi2;
                const int cls = tri_interval(sc.tri_iv[idx - ns], so32x, so32y, so32z, solx, soly, solz, Lx, Ly, Lz, 0.0f, 8.0f * kEps32, lo, hi2);
                if (cls == kCandMiss || lo > tmax_hi) return;
                if (cls == kCandHit && lo > bias_hi && hi2 < tmax_lo) { first_t = min(first_t, (uint32_t)(idx - ns)); occl_t = true; return; }
                unsure = true; go = false;
                return;
            }
            const float4 hi = __ldg(sc.sph_hl + 2 * idx);
            if (far_miss(hi, so32x, so32y, so32z, so1, Lx, Ly, Lz)) return;
            const Cand32 c = classify_sphere(hi, __ldg(sc.sph_hl + 2 * idx + 1), so32x, so32y, so32z, solx, soly, solz, Lx, Ly, Lz);
            float lo, hi2;
            const int cls = reported_t(c, lo, hi2);
            if (cls == kCandMiss) return;
            if (lo > tmax_hi) return;  // behind the light: not an occluder, not `near`
            if (cls == kCandHit && lo > bias_hi && hi2 < tmax_lo) {  // certainly bias < t < maxDist
                first = min(first, (uint32_t)idx);  // (the lists are unordered: the first occluder in canonical order is the smallest index)
                occl = true;
                return;
            }
            unsure = true; go = false;
        });
        if (unsure) ok = false;
        uint32_t firstp = (uint32_t)np;
        if (cast && !occl && !unsure) {  // planes, any-hit in canonical order
#pragma unroll 1
            for (int i = 0; i < np; ++i) {
                const float en = __fmaf_rn(4.0f * kEps32, fabsf(S.pl[i].d) + so1, 1e-9f);
                const int cls = plane_shadow(pl_num[i * kR32Threads], en, S.plA[j][i], dist, bias_hi, tmax_lo, tmax_hi);
                if (cls == kCandMiss) continue;
                if (cls == kCandHit) { occl = true; firstp = (uint32_t)i; break; }
                ok = false; break;
            }
        }
        if (cast) {
            // canonical order of the reference's scan: spheres, planes, triangles — tests up to and including the first occluder
            out.c_shadow += 1u;
            out.c_sph += first < (uint32_t)ns ? first + 1u : (uint32_t)ns;
            out.c_pl += first < (uint32_t)ns ? 0u : (firstp < (uint32_t)np ? firstp + 1u : (uint32_t)np);
            out.c_tri += (first < (uint32_t)ns || firstp < (uint32_t)np) ? 0u : (first_t < (uint32_t)nt ? first_t + 1u : (uint32_t)nt);
            occl = occl || occl_t;
            if (j < 16) out.lmask |= 1u << (16 + j);
            if (!occl) {
                if (j < 16) out.lmask |= 1u << j;
                const float w = inv * inv * ndl;  // (1 / d^2) * N.L
                ar = __fmaf_rn(S.li_E[j][0], w, ar); ag = __fmaf_rn(S.li_E[j][1], w, ag); ab = __fmaf_rn(S.li_E[j][2], w, ab);
            }
        }
    }
    if (hitv) {
        const rte_material& M = sc.materials[mat];
        out.r = (float)M.color[0] * ar; out.g = (float)M.color[1] * ag; out.b = (float)M.color[2] * ab;
    }
    return ok || !alive;
}

// ------------------------------------------------------------------------------ deferred exact pixels
// A warp queues the (band-local) indices of its pixels with an inconclusive FP32 decision and renders them in batches with the
// reference's exact FP64 operations (rte_exact_grid.cuh).  Lanes = (pixel, light): every lane of a pixel's group evaluates the
// primary hit, lane j the shadow ray of light j, and the contributions are added in the lights' order (RE/Scene.h:86,113) —
// all 32 lanes busy, and the dependent FP64 chain of a pixel is 1 + 1 rays long instead of 1 + nl.
constexpr int kFixCap = 64;  // per-warp queue
static __device__ __noinline__ void fix_batch(const RenderParams& P, const uint32_t* __restrict__ q, int n, unsigned long long* __restrict__ stats_row) {
    const int lane = threadIdx.x & 31;
    const DevScene& sc = P.sc;
    const int nl = sc.nl, W = P.cam.width, sr = P.stripe_rows, sp = P.stripe_period, aa = P.cam.aa;
    int lpp = 1;
    while (lpp < nl) lpp <<= 1;  // lanes per pixel (nl <= 16)
    const int per = 32 / lpp, slot = lane / lpp, j = lane % lpp, lead = lane & ~(lpp - 1);
    uint32_t st[kStatCount] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 1
    for (int base = 0; base < n; base += per) {
        const bool act = base + slot < n;
        const int64_t pix = act ? (int64_t)q[base + slot] : 0;
        const int band_y = (int)(pix / W), px = (int)(pix - (int64_t)band_y * W);
        const int py = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
        d3 v = mk(0.0, 0.0, 0.0);
        uint32_t prim0 = kMiss, lmask0 = 0u;
#pragma unroll 1
        for (int s = 0; s < aa; ++s) {  // GeneratePixelAt, RE/Scene.h:283-304
            ExactHit H;
            H.hit = false; H.sky = mk(0.0, 0.0, 0.0); H.prim = kMiss; H.mat = sc.nm; H.Pn = mk(0.0, 0.0, 0.0); H.normal = mk(0.0, 0.0, 1.0);
            H.ngeo = mk(0.0, 0.0, 0.0); H.t = 0.0;
            if (act) exact_primary(P, sc, px, py, s, H);
            PixelCounters c = {0, 0, 0, 0, 0, 0, 0, 0};
            d3 contrib = mk(0.0, 0.0, 0.0);
            uint32_t bits = 0u;
            if (act && j < nl) contrib = exact_light(P, sc, H, j, bits, c);
            d3 diffuse = mk(0.0, 0.0, 0.0);
            uint32_t lm = 0u;
            for (int k = 0; k < nl; ++k) {  // (warp-uniform) diffuseAccumulation += contribution * transmittance, in the lights' order
                const double cx = __shfl_sync(kFullMask, contrib.x, lead + k), cy = __shfl_sync(kFullMask, contrib.y, lead + k),
                             cz = __shfl_sync(kFullMask, contrib.z, lead + k);
                const uint32_t b = __shfl_sync(kFullMask, bits, lead + k);
                if (b & 1u) diffuse = diffuse + mk(cx, cy, cz);
                if (k < 16) lm |= (((b >> 16) & 1u) << (16 + k)) | ((b & 1u) << k);
            }
            const d3 e = exact_finish(sc, H, diffuse);
            if (aa > 1) v = v + e; else v = e;
            if (s == 0) { prim0 = H.prim; lmask0 = lm; }
            st[kStatShadow] += c.shadow; st[kStatSphTests] += c.sph; st[kStatPlTests] += c.pl; st[kStatTriTests] += c.tri; st[kStatMarches] += c.marches;
            if (act && j == 0) { st[kStatPrimary] += 1u; st[kStatSphTests] += (uint32_t)sc.ns; st[kStatPlTests] += (uint32_t)sc.np; st[kStatTriTests] += (uint32_t)sc.nt; st[kStatConfirms] += 1u; }
        }
        if (aa > 1) v = v / (double)aa;
        if (act && j == 0) {
            if (P.out.hdr) reinterpret_cast<float4*>(P.out.hdr)[pix] = make_float4((float)v.x, (float)v.y, (float)v.z, 0.f);
            if (P.out.prim_id) P.out.prim_id[pix] = prim0;
            if (P.out.light_mask) P.out.light_mask[pix] = lmask0;
            if (P.out.rgb8) {
                const uchar3 u = tm64::to_color(tm64::apply(P.tonemap_op, v, P.max_white, P.jodie_a));  // the reference's own bytes
                uint8_t* d = P.out.rgb8 + 3 * pix;
                d[0] = u.x; d[1] = u.y; d[2] = u.z;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < kStatCount; ++k) {
        const uint32_t t = __reduce_add_sync(kFullMask, st[k]);
        if (lane == 0 && t) stats_row[k] += t;
    }
    __syncwarp();
}

// ------------------------------------------------------------------------------ the kernel
// Persistent CTAs, one 8 x 4 pixel tile per warp at a time from a global counter (like the exact kernels), one pixel per lane.
#ifndef RTE_TW
#define RTE_TW 8
#define RTE_TH 4
#endif
constexpr int kTile32W = RTE_TW, kTile32H = RTE_TH;

template <bool kAov, bool kTris>
__global__ void __launch_bounds__(kR32Threads, 1) render32_kernel(const __grid_constant__ RenderParams P) {
    __shared__ Scene32 s32;
    __shared__ unsigned long long s_stats[kR32Threads / 32][kStatCount];
    __shared__ uint32_t s_fixq[kR32Threads / 32][kFixCap];
    __shared__ float s_plnum[kR32Planes][kR32Threads];  // per pixel: d_i - n_i . so of every plane (trace32)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int fix_n = 0;  // (warp-uniform) pixels waiting in s_fixq[warp]
    int fix_full = 32;  // pixels of one full batch of fix_batch: 32 lanes / lanes per pixel (the lights, rounded up to a power of two)
    { int lpp = 1; while (lpp < P.sc.nl) lpp <<= 1; fix_full = 32 / (lpp > 32 ? 32 : lpp); if (fix_full < 1) fix_full = 1; }
    {   // scene constants: one thread per plane / light / (light, plane) pair (one thread doing all of it was ~15 us of serial loads)
        const DevScene& g = P.sc;
        const int t = (int)threadIdx.x;
        if (t == 0) { s32.np = g.np; s32.nl = g.nl; }
        if (t < 3) { s32.cam[t] = P.cam.pos[t]; s32.cam_hi[t] = (float)P.cam.pos[t]; s32.cam_lo[t] = (float)(P.cam.pos[t] - (double)s32.cam_hi[t]); }
        if (t >= 32 && t < 32 + g.np) {
            const int i = t - 32;
            const DevPlane p = g.planes[i];
            s32.pl[i].nx = (float)p.n[0]; s32.pl[i].ny = (float)p.n[1]; s32.pl[i].nz = (float)p.n[2];
            s32.pl[i].d = (float)(p.n[0] * p.p0[0] + p.n[1] * p.p0[1] + p.n[2] * p.p0[2]);
            s32.pl[i].mat = p.mat;
        }
        if (t >= 64 && t < 64 + g.nl) {
            const int j = t - 64;
            const DevLight l = g.lights[j];
            for (int c = 0; c < 3; ++c) {
                s32.li_hi[j][c] = (float)l.pos[c]; s32.li_lo[j][c] = (float)(l.pos[c] - (double)s32.li_hi[j][c]);
                s32.li_E[j][c] = (float)l.E[c];
            }
        }
        if (t >= 128 && t < 128 + g.nl * g.np) {
            const int j = (t - 128) / g.np, i = (t - 128) - j * g.np;
            const DevPlane p = g.planes[i];
            const DevLight l = g.lights[j];
            const double d = p.n[0] * p.p0[0] + p.n[1] * p.p0[1] + p.n[2] * p.p0[2];
            s32.plA[j][i] = (float)((p.n[0] * l.pos[0] + p.n[1] * l.pos[1] + p.n[2] * l.pos[2]) - d);
        }
    }
    if (lane < kStatCount) s_stats[warp][lane] = 0ull;
    __syncthreads();
    tm32::Consts K;
    {
        const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f, x = 11.2f;  // 1 / uncharted2_tonemap_partial(11.2)
        K.white_scale = 1.0f / (((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F);
        K.inv_wsq = (float)(1.0 / (P.max_white * P.max_white));
        K.jodie_a = (float)P.jodie_a;
    }
    const int W = P.cam.width, rows = P.n_rows;
    const int sr = P.stripe_rows, sp = P.stripe_period;
    const int tiles_x = (W + kTile32W - 1) / kTile32W, tiles_y = (rows + kTile32H - 1) / kTile32H;
    const unsigned int n_tiles = (unsigned int)(tiles_x * tiles_y);
    const float inv_tiles_x = 1.0f / (float)tiles_x;
    const int lx = lane & (kTile32W - 1), ly = lane / kTile32W;
    const int aa = P.cam.aa;
    const bool words = (W & 3) == 0;  // rows of RGB8 are word aligned: a tile row of 8 pixels is six aligned words

    // first tile by position, the following ones from the global counter (every warp of the grid fetching its first tile from
    // one address was ~9 us of serialised atomics in front of the first ray: a quarter of a 1/8-frame band's overhead)
    const unsigned int n_static = gridDim.x * (kR32Threads / 32);
    unsigned int tile = blockIdx.x * (kR32Threads / 32) + warp;
    while (tile < n_tiles) {
        unsigned int next = 0;
        if (lane == 0) next = n_static + atomicAdd(P.tile_counter, 1u);  // the next tile id arrives behind the work
#ifdef RTE_TILE_CLOCK
        unsigned long long tc0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tc0));
#endif
        int ty = (int)((float)tile * inv_tiles_x);  // tile / tiles_x without the integer-division sequence (tile < 2^24 for every frame <= 16K x 16K)
        if ((unsigned)(ty * tiles_x) > tile) --ty; else if ((unsigned)((ty + 1) * tiles_x) <= tile) ++ty;
        const int tx = (int)tile - ty * tiles_x;
        const int band_y = ty * kTile32H + ly, px = tx * kTile32W + lx;
        const int py = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
        const bool inb = px < W && band_y < rows;
        const int64_t pix = (int64_t)band_y * W + px;
        // GeneratePixelAt, RE/Scene.h:283-304: mean of the samples
        float cr = 0.0f, cg = 0.0f, cb = 0.0f;
        uint32_t prim = kMiss, lmask = 0u, n_shadow = 0u, n_sph = 0u, n_pl = 0u, n_tri = 0u, executed = 0u;
        bool ok = true;
#pragma unroll 1
        for (int s = 0; s < aa; ++s) {
            Px32 o;
            ok = trace32<kTris>(P, s32, &s_plnum[0][threadIdx.x], px, py, inb, s, o, executed) && ok;
            cr += o.r; cg += o.g; cb += o.b;
            n_shadow += o.c_shadow; n_sph += o.c_sph; n_pl += o.c_pl; n_tri += o.c_tri;
            if (s == 0) { prim = o.prim; lmask = o.lmask; }
        }
        if (aa > 1) { const float ia = 1.0f / (float)aa; cr *= ia; cg *= ia; cb *= ia; }
        const bool flagged = inb && !ok, mine = inb && ok;  // flagged pixels wait in the warp's queue for an exact batch (fix_batch)
        // ---- outputs
        if (mine) {
            if (P.out.hdr) __stcs(reinterpret_cast<float4*>(P.out.hdr) + pix, make_float4(cr, cg, cb, 0.f));
            if (kAov) {
                if (P.out.prim_id) P.out.prim_id[pix] = prim;
                if (P.out.light_mask) P.out.light_mask[pix] = lmask;
            }
        }
        if (P.out.rgb8) {
            const uint32_t rgb = tonemap_rgb24(P.tonemap_op, cr, cg, cb, K);  // (a flagged pixel's bytes are overwritten by fix_batch)
            const bool row_full = (tx + 1) * kTile32W <= W && band_y < rows;  // (uniform over the 8 lanes of a tile row)
            if (words) {
                // bytes 4 lx .. 4 lx + 3 of the row's 24: pixel q = 4 lx / 3 from its byte o = 4 lx % 3 on, then pixel q + 1
                const int q = lx < 6 ? (4 * lx) / 3 : 6, o = lx < 6 ? (4 * lx) % 3 : 0;
                const uint32_t a = __shfl_sync(kFullMask, rgb, (lane & ~7) + q), b = __shfl_sync(kFullMask, rgb, (lane & ~7) + q + 1);
                const uint32_t word = o == 0 ? (a | (b << 24)) : o == 1 ? ((a >> 8) | (b << 16)) : ((a >> 16) | (b << 8));
                if (row_full && lx < 6) __stcs(reinterpret_cast<uint32_t*>(P.out.rgb8 + 3 * ((int64_t)band_y * W + tx * kTile32W)) + lx, word);
            }
            if (mine && !(words && row_full)) {
                uint8_t* d = P.out.rgb8 + 3 * pix;
                d[0] = (uint8_t)rgb; d[1] = (uint8_t)(rgb >> 8); d[2] = (uint8_t)(rgb >> 16);
            }
        }
        // ---- statistics of the pixels shaded here (fix_batch counts its own; fp64_confirms = pixel samples rendered again in FP64)
        const uint32_t v[kStatCount] = {mine ? (uint32_t)aa : 0u, mine ? n_shadow : 0u, 0u, mine ? n_sph : 0u, mine ? n_pl : 0u, mine ? n_tri : 0u, 0u, 0u, executed};
        uint32_t sum_mine = 0;
#pragma unroll
        for (int k = 0; k < kStatCount; ++k) {
            const uint32_t t = __reduce_add_sync(kFullMask, v[k]);
            if (lane == k) sum_mine = t;
        }
        if (lane < kStatCount) s_stats[warp][lane] += sum_mine;
        // ---- queue the flagged pixels; a full batch (every lane a (pixel, light) pair) is rendered exactly right away
        const uint32_t fm = __ballot_sync(kFullMask, flagged);
        if (fm) {
            if (flagged) s_fixq[warp][fix_n + __popc(fm & ((1u << lane) - 1u))] = (uint32_t)pix;
            fix_n += __popc(fm);
            __syncwarp();
            // a batch as soon as every lane of it has a (pixel, light) pair: waiting for a fuller queue left up to eight batches
            // (C5: 8 lights, 4 pixels per batch) to the end of the launch, where nothing hides them — ~0.1 ms per launch
            if (fix_n >= fix_full || fix_n > kFixCap - 32) {
#ifdef RTE_TILE_CLOCK
                unsigned long long fc0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(fc0));
#endif
                fix_batch(P, s_fixq[warp], fix_n, s_stats[warp]);
#ifdef RTE_TILE_CLOCK
                unsigned long long fc1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(fc1));
                if (lane == 0) atomicMax(&P.stats[kStatTriTests], (fc1 - fc0) * 1000ull + (unsigned long long)fix_n);  // slowest batch: ns * 1000 + pixels
#endif
                fix_n = 0;
            }
        }
#ifdef RTE_TILE_CLOCK
        {   // developer probe: slowest tile of the launch in ns (reported as secondary_rays), tiles over 100 us (as exact_marches)
            unsigned long long tc1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tc1));
            if (lane == 0) { atomicMax(&P.stats[kStatSecondary], tc1 - tc0); }
        }
#endif
        tile = __shfl_sync(kFullMask, next, 0);
    }
    if (fix_n > 0) fix_batch(P, s_fixq[warp], fix_n, s_stats[warp]);  // what is left of the warp's queue
    __syncwarp();
    if (lane < kStatCount && s_stats[warp][lane]) atomicAdd(&P.stats[lane], s_stats[warp][lane]);
}

// Can this render call take the FP32 path?  (rte_api.cu decides with it; everything else takes the exact kernels.)
bool fp32_path_eligible(const RenderParams& p) {
    const rte_frame& f = p.out;
    return !p.sc.reflective && !p.sc.transparent && p.sc.ns + p.sc.nt > 0 && p.sc.np <= kR32Planes && p.sc.nl <= kR32Lights &&
           !p.exact_shadows && !p.no_filter && !p.force_general && !p.no_cull && p.pass_mode == 0 && !p.exact_fp64 &&
           !f.hdr64 && !f.depth && !f.depth64 && !f.depth_norm && !f.normal && !f.albedo && p.bias >= 1e-6 && p.bias <= 1e-1;
}

cudaError_t launch_render32(const RenderParams& p, int sm_count, cudaStream_t stream) {
    const bool aov = p.out.prim_id || p.out.light_mask;
    const bool tris = p.sc.nt > 0;
    const void* fn = tris ? (aov ? (const void*)render32_kernel<true, true> : (const void*)render32_kernel<false, true>)
                          : (aov ? (const void*)render32_kernel<true, false> : (const void*)render32_kernel<false, false>);
    const long long n_tiles = (long long)((p.cam.width + kTile32W - 1) / kTile32W) * ((p.n_rows + kTile32H - 1) / kTile32H);
    long long blocks = sm_count;  // persistent: one CTA of 32 warps per SM (64 registers per thread fill the register file)
    const long long needed = (n_tiles + kR32Threads / 32 - 1) / (kR32Threads / 32);
    if (blocks > needed) blocks = needed > 0 ? needed : 1;
    void* args[] = {(void*)&p};
    return cudaLaunchKernel(fn, dim3((unsigned)blocks), dim3(kR32Threads), args, 0, stream);
}

}  // namespace rte
