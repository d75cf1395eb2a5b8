// rte_views.cuh — origin-specialised sphere filters ("views") with packed FP32x2 FMAs.
//
// Every ray of the frame starts at the camera (primary rays) or ends at a light (shadow rays, RE/Scene.h:
// 86-104; "RE/" = /root/reference/RaytracingEngine/).  A shadow ray is filtered *reversed*, from the light
// towards the shaded point, so both families share their origin O_j across all pixels.  Everything in the
// ray-sphere test that depends only on (sphere i, origin j) is hoisted into per-frame "view" arrays.
//
// HOT array (shared memory, scanned for every ray): seen from O_j, sphere i is a cone of directions
//     c_ij = (C_i - O_j) / F,  F = |C_i - O_j|            (FP64, rounded to FP32)
//     w_ij = kConeMargin - cos(theta_ij),  cos(theta_ij) = sqrt(1 - (r_i'/F)^2)   (FP64, rounded UP to FP32)
// and a unit direction u can hit it only if  h = u.c + w >= 0  — 3 FMA per test, two spheres per instruction with
// fma.rn.f32x2 (SASS FFMA2, sm_100+).  The sign bits of h are shifted into a 32-bit mask (1 SHF per test) and only
// the rare clear bits are looked at again.  r_i' = r_i for the camera and r_i + bias for a light: the true shadow
// line starts at P + N*bias (RE/Scene.h:104) and passes the light at a distance of at most `bias`.  An origin inside
// (or on) the sphere makes it a candidate for every direction (w = 4).
//
// COLD array (global memory, read once per candidate): {f_ij = C_i - O_j, ng_ij = -(|f|^2 - r^2) + M_ij}, from which
// the resolution computes the chord of the ray's line through the sphere in FP32 interval arithmetic (b = f.u,
// h' = b^2 + ng) to prune / classify the candidate before the FP64 confirmation in the reference's own operation
// order (rte_device.cuh).
//
// The filter is conservative: with u = fl32(d), |d| = 1 (FP64-normalised by the caller) and
// eps = 2^-24, the rounding of u and c perturbs u.c by at most 2 eps, the three FMAs by at most 2 eps each (partial
// sums <= 2), w is rounded up: the computed h is at least the exact u.c - cos(theta) + kConeMargin - 8 eps, so with
// kConeMargin = 16 eps a sphere whose exact test can succeed always has h >= 0.  It may accept extra spheres (cost only).
#pragma once
#include "rte_trace.cuh"

namespace rte {

constexpr int kViewBlock = 32;  // spheres per candidate mask (ns_pad is a multiple of this)
constexpr double kConeMargin = 16.0 * 5.9604644775390625e-08;  // 16 * 2^-24

// ------------------------------------------------------------------------------ packed f32x2 helpers
typedef unsigned long long f32x2;  // {lo, hi} = two FP32 values in one 64-bit register pair

__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2_bits(f32x2 v, uint32_t& lo, uint32_t& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// ------------------------------------------------------------------------------ view layout
// One view = ns_pad/2 sphere pairs of 32 bytes:  hot {cx0,cx1, cy0,cy1} {cz0,cz1, w0,w1},  cold {fx0,fx1, fy0,fy1} {fz0,fz1, ng0,ng1}.
__host__ __device__ __forceinline__ size_t view_bytes(int ns_pad) { return (size_t)ns_pad * 16; }

struct ViewEntry { float fx, fy, fz, ng; };
__device__ __forceinline__ ViewEntry view_entry(const float* __restrict__ view, int idx) {
    const float* p = view + (size_t)(idx >> 1) * 8 + (idx & 1);
    ViewEntry e;
    e.fx = p[0]; e.fy = p[2]; e.fz = p[4]; e.ng = p[6];
    return e;
}

// Builds view j (0 = camera, 1.. = lights) for every sphere — or, kTri, for the bounding sphere of every triangle
// (centroid, radius = farthest vertex: a ray that hits the triangle hits that sphere).  One thread per (view, primitive).
template <bool kTri>
__global__ void __launch_bounds__(256) build_views_kernel(const DevScene sc, const DevCamera cam, double bias, float* __restrict__ views,
                                                          float* __restrict__ cold) {
    const int n = kTri ? sc.nt : sc.ns, n_pad = kTri ? sc.nt_pad32 : sc.ns_pad, nv = 1 + sc.nl;
    const int total = nv * n_pad;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
        const int j = t / n_pad, i = t - j * n_pad;
        float fx = 0.f, fy = 0.f, fz = 0.f, ng = -INFINITY;  // cold padding: h' = -inf
        float cx = 0.f, cy = 0.f, cz = 0.f, w = -INFINITY;   // hot padding: never a candidate
        if (i < n) {
            const d3 O = j == 0 ? ld3(cam.pos) : ld3(sc.lights[j - 1].pos);
            d3 C;
            double r;
            if (kTri) {
                const DevTri64 T = sc.tri64[i];
                const d3 a = ld3(T.a0), b = a + ld3(T.e1), c = a + ld3(T.e2);
                C = (a + b + c) * (1.0 / 3.0);
                const d3 da = a - C, db = b - C, dc = c - C;
                r = sqrt(fmax(dot(da, da), fmax(dot(db, db), dot(dc, dc)))) * (1.0 + 1e-9) + 1e-300;
            } else {
                const double4 s = sc.sph64[i];
                C = mk(s.x, s.y, s.z);
                r = fabs(s.w);
            }
            const d3 f = C - O;
            const double F2 = dot(f, f);
            double M = 16.0 * (double)kEps32 * F2 + 1e-30;
            if (j > 0) M += 2.0 * r * bias + bias * bias;
            fx = (float)f.x; fy = (float)f.y; fz = (float)f.z;
            ng = __double2float_ru((r * r - F2) + M);
            if (!(ng == ng)) ng = INFINITY;  // non-finite input: always a candidate (the FP64 test decides)
            // the cone of directions from O that can touch the sphere (inflated by `bias` for a light)
            const double ri = (j > 0 ? r + bias : r) * (1.0 + 1e-12), F = sqrt(F2);
            w = 4.0f;  // origin inside / on the sphere, or non-finite input: every direction is a candidate
            if (F > ri && F < 1e150) {
                const double q = ri / F;
                const d3 c = f / F;
                cx = (float)c.x; cy = (float)c.y; cz = (float)c.z;
                w = __double2float_ru(kConeMargin - sqrt(fmax(0.0, 1.0 - q * q)));
                if (!(w == w) || !(cx == cx) || !(cy == cy) || !(cz == cz)) { cx = cy = cz = 0.f; w = 4.0f; }
            }
        }
        const size_t at = ((size_t)j * n_pad + (size_t)(i & ~1)) * 4 + (i & 1);
        float* p = views + at;
        p[0] = cx; p[2] = cy; p[4] = cz; p[6] = w;
        float* q = cold + at;
        q[0] = fx; q[2] = fy; q[4] = fz; q[6] = ng;
    }
}

// ------------------------------------------------------------------------------ the scan
// K rays per thread share every LDS.128 of sphere data (a broadcast LDS.128 costs two shared-memory
// wavefronts: with one ray per thread the loop is shared-memory bound, not FMA bound).
struct ViewRay {
    float ux, uy, uz;  // fl32 of the unit direction the view is scanned with
};

__device__ __forceinline__ ViewRay make_view_ray(d3 u) {
    ViewRay r;
    r.ux = (float)u.x; r.uy = (float)u.y; r.uz = (float)u.z;
    return r;
}

// Filters spheres [base, base+32) of one (hot) view against K unit directions; cand[k] bit (31-i) <-> sphere base+i.
// The 16 sphere pairs are walked in chunks of kPairs unrolled pairs: the loop body (kPairs * (2 LDS + 5K) SASS
// instructions) has to stay well inside the ~6 KB L0 instruction cache, or the warps of an SMSP starve on fetch.
template <int K>
__device__ __forceinline__ void view_block_masks(const ulonglong2* __restrict__ v, const ViewRay (&r)[K], uint32_t (&cand)[K]) {
    constexpr int kPairs = K == 1 ? 16 : (K == 2 ? 8 : 4);
    uint32_t m[K];
    f32x2 ux2[K], uy2[K], uz2[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        m[k] = 0;
        ux2[k] = pack2(r[k].ux, r[k].ux); uy2[k] = pack2(r[k].uy, r[k].uy); uz2[k] = pack2(r[k].uz, r[k].uz);  // SASS: scalar-broadcast operand
    }
#pragma unroll 1
    for (int c = 0; c < kViewBlock / 2; c += kPairs) {
#pragma unroll
        for (int p = 0; p < kPairs; ++p) {
            const ulonglong2 A = v[2 * (c + p)], B = v[2 * (c + p) + 1];
#pragma unroll
            for (int k = 0; k < K; ++k) {
                f32x2 h = fma2(B.x, uz2[k], B.y);  // h = u.c + w
                h = fma2(A.y, uy2[k], h);
                h = fma2(A.x, ux2[k], h);
                uint32_t h0, h1;
                unpack2_bits(h, h0, h1);
                m[k] = __funnelshift_l(h0, m[k], 1);  // m = (m << 1) | sign(h0)
                m[k] = __funnelshift_l(h1, m[k], 1);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) cand[k] = ~m[k];  // sign clear (h >= 0) -> candidate
}

// The FP64 confirmation as ONE out-of-line copy: it is the rare path of every scan site, and inlining it K times
// per site bloats the kernel far beyond the instruction caches.
static __device__ __noinline__ bool confirm_sphere(const double4* __restrict__ sph64, int idx, d3 o, d3 d, double& t) {
    return sphere_exact(sph64[idx], o, d, t);
}
static __device__ __noinline__ bool confirm_triangle(const DevTri64* __restrict__ tri64, int idx, d3 o, d3 d, double& t) {
    return tri_exact(tri64[idx], o, d, t);
}
// what a view holds: the spheres of the scene, or (kTri) the bounding spheres of its triangles
template <bool kTri> __device__ __forceinline__ int view_count(const DevScene& sc) { return kTri ? sc.nt : sc.ns; }
template <bool kTri> __device__ __forceinline__ int view_padded(const DevScene& sc) { return kTri ? sc.nt_pad32 : sc.ns_pad; }
template <bool kTri> __device__ __forceinline__ float view_extent(const DevScene& sc) { return kTri ? sc.tri_extent + sc.tri_edge : sc.sph_extent; }
template <bool kTri> __device__ __forceinline__ bool confirm_prim(const DevScene& sc, int idx, d3 o, d3 d, double& t) {
    return kTri ? confirm_triangle(sc.tri64, idx, o, d, t) : confirm_sphere(sc.sph64, idx, o, d, t);
}

// Per-view constants of the FP32 interval arithmetic on candidates.
struct ViewBounds {
    float slack;   // absolute error bound on distances along the ray
    float m2;      // 2 * (upper bound of every M_ij of the view): h~ - m2 <= true half-chord^2 <= h~
};
__device__ __forceinline__ ViewBounds view_bounds(float extent, d3 origin, double bias, bool light_view) {  // extent >= |C_i| + r_i
    const float S = sqrtf((float)dot(origin, origin)) * 1.0000002f + extent;
    ViewBounds vb;
    vb.slack = 24.0f * kEps32 * S + 1e-6f;
    float m = 24.0f * kEps32 * S * S + 1e-30f;
    if (light_view) m += (float)(2.0 * bias) * extent + (float)(bias * bias);
    vb.m2 = 2.0f * m * 1.0001f;
    return vb;
}

// Centre parameter and half-chord bounds of candidate `idx` along the scan line.
struct ViewChord { float sc, hc_lo, hc_hi; bool miss; };  // miss: the line certainly passes outside the sphere
__device__ __forceinline__ ViewChord view_chord(const float* __restrict__ view, int idx, const ViewRay& r, const ViewBounds& vb) {
    const ViewEntry e = view_entry(view, idx);
    const float b = __fmaf_rn(e.fz, r.uz, __fmaf_rn(e.fy, r.uy, e.fx * r.ux));
    const float h = __fmaf_rn(b, b, e.ng);
    ViewChord c;
    c.sc = b;
    c.miss = h < 0.0f;
    c.hc_hi = sqrtf(fmaxf(h, 0.0f)) * 1.000001f;
    c.hc_lo = sqrtf(fmaxf(h - vb.m2, 0.0f)) * 0.999999f;
    return c;
}

// ------------------------------------------------------------------------------ warp-level culling
// The K*32 rays a warp sends through one view (its 8 x 4K pixel tile from the camera, or the shadow rays of those pixels
// from one light) are coherent: they lie in a narrow cone around their mean direction.  Before the per-ray scan the warp
// tests every sphere of the view ONCE against that tile cone — lanes take different spheres, coalesced LDS.128 — and
// packs the survivors (typically a few percent) into a private mini-view of the same layout, which the per-ray scan then
// walks instead of the whole view.  Cone-cone test on the unit sphere: sphere cone (axis c, half-angle ts) and tile cone
// (axis a, half-angle tt) can share a direction only if  angle(c, a) <= ts + tt,  i.e.  c.a >= cos ts cos tt - sin ts sin tt.
// Conservative like the per-ray test (bounds rounded outwards); the survivors keep their ascending index order, so the
// resolution still sees candidates in canonical order.  Tiles whose rays diverge (silhouettes: cos tt < kCullMinCos) or
// with more than kCullCap survivors scan the full view as before.
#ifndef RTE_CULL_UNROLL
#define RTE_CULL_UNROLL 1  // measured: 1 -> 3.68 ms, 2 -> 4.07, 4 -> 4.29 on C2 (the kernel waits on instruction fetch)
#endif
constexpr int kCullUnroll = RTE_CULL_UNROLL;
#ifndef RTE_CULL_CAP
#define RTE_CULL_CAP 128
#endif
constexpr int kCullCap = RTE_CULL_CAP;  // survivors per (tile, view); a multiple of 32
constexpr int kCullMin = 128;         // views with fewer entries are scanned directly
constexpr float kCullMinCos = 0.5f;   // tile cones wider than 60 degrees are not worth testing
constexpr int kCullBytesPerWarp = kCullCap * 16 + kCullCap * 2;
struct CullScratch {
    ulonglong2* mini;  // kCullCap entries in pair layout (shared memory, private to the warp)
    uint16_t* map;     // position in the mini-view -> index in the view
};

// Tests the n_pad entries of `view` against the tile cone (ax,ay,az | ct = lower bound of cos tt, st2 = upper bound of
// sin^2 tt) and packs the survivors.  All 32 lanes; returns the survivor count (> kCullCap: overflow, scratch unusable).
// No square root: with x = c.a - cos ts cos tt (+ slack), the entry survives iff x >= 0 or x^2 <= sin^2 ts sin^2 tt.
// Both spheres of a pair at once (packed FP32x2, SASS FFMA2): with cs = -w (a lower bound of cos ts),
//     x   = c.a + 4e-6 + w ct            >= the exact c.a - cos ts cos tt
//     rhs = st2' (1 - w^2) + 1e-12       >= sin^2 ts sin^2 tt          (st2' = st2 * 1.00001)
//     y   = x^2 - rhs
// and the entry survives iff x >= 0 or y < 0, i.e. iff sign(x) is clear or sign(y) is set.  (y == 0 may be rejected: rhs is
// inflated by 1e-5 relative, far more than the rounding of x^2.)  The special entries need no test of their own: the origin
// inside a sphere (w = 4) gives x >= 4 ct - 1 > 0 because ct >= kCullMinCos = 0.5, padding (w = -inf) gives x = -inf, y = +inf.
struct CullCone {
    f32x2 ax, ay, az, ct, slack, st2p, nrc;  // nrc = -(st2' + 1e-12)
};
__device__ __forceinline__ CullCone make_cull_cone(float ax, float ay, float az, float ct, float st2) {
    const float st2p = st2 * 1.00001f, nrc = -(st2p + 1e-12f);
    CullCone c;
    c.ax = pack2(ax, ax); c.ay = pack2(ay, ay); c.az = pack2(az, az); c.ct = pack2(ct, ct);
    c.slack = pack2(4e-6f, 4e-6f); c.st2p = pack2(st2p, st2p); c.nrc = pack2(nrc, nrc);
    return c;
}
// A = {cx0,cx1 | cy0,cy1}, B = {cz0,cz1 | w0,w1}; the sign bit of k0 / k1 says "keep"
__device__ __forceinline__ void cull_keep2(const ulonglong2& A, const ulonglong2& B, const CullCone& c, uint32_t& k0, uint32_t& k1) {
    f32x2 x = fma2(A.x, c.ax, c.slack);
    x = fma2(A.y, c.ay, x);
    x = fma2(B.x, c.az, x);
    x = fma2(B.y, c.ct, x);
    const f32x2 nrhs = fma2(mul2(B.y, B.y), c.st2p, c.nrc);  // -(st2' (1 - w^2) + 1e-12)
    const f32x2 y = fma2(x, x, nrhs);
    uint32_t x0, x1, y0, y1;
    unpack2_bits(x, x0, x1);
    unpack2_bits(y, y0, y1);
    k0 = ~x0 | y0;
    k1 = ~x1 | y1;
}
__device__ __forceinline__ ulonglong2 lds_u2(uint32_t addr) {
    ulonglong2 v;
    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void cull_store(float* __restrict__ mf, uint16_t* __restrict__ map, int pos, float cx, float cy, float cz, float w, int idx) {
    float* e = mf + (size_t)(pos >> 1) * 8 + (pos & 1);
    e[0] = cx; e[2] = cy; e[4] = cz; e[6] = w;
    map[pos] = (uint16_t)idx;
}
// `md`: this lane's smallest u.a over its active rays (1 if it has none).  Returns -1 when the tile cone is too wide to be worth testing.
// `src`: the view is itself a culled one (CTA level): position -> index in the full view, or NULL.
static __device__ __noinline__ int cull_pack(const ulonglong2* __restrict__ view, int n_pad, float ax, float ay, float az, float md,
                                      ulonglong2* __restrict__ mini, uint16_t* __restrict__ map, const uint16_t* __restrict__ src) {
    // min over the warp with one REDUX: non-negative floats order like their bit patterns (a negative dot clamps to 0: too wide anyway)
    md = __uint_as_float(__reduce_min_sync(kFullMask, __float_as_uint(fmaxf(md, 0.0f))));
    const float ct = md * 0.999998f - 2e-6f;  // lower bound of cos(tile half-angle): |a|, |u| and the dots are off by < 1e-6
    if (!(ct >= kCullMinCos)) return -1;
    const float st2 = fmaxf(1.0f - ct * ct, 0.0f) * 1.00001f + 1e-12f;  // upper bound of sin^2(tile half-angle)
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const int n_pairs = n_pad >> 1;  // a multiple of 16
    float* mf = reinterpret_cast<float*>(mini);
    const uint32_t v0 = (uint32_t)__cvta_generic_to_shared(view);  // the views live in shared memory
    const CullCone cone = make_cull_cone(ax, ay, az, ct, st2);
    int count = 0;
    uint32_t va = v0 + 32u * (uint32_t)lane;
    // survivors of one group of 32 pairs (pair p of this lane), packed behind `count` in ascending index
    auto emit = [&](const ulonglong2& A, const ulonglong2& B, uint32_t kw0, uint32_t kw1, int p) {
        const bool k0 = (int)kw0 < 0, k1 = (int)kw1 < 0;
        const uint32_t m0 = __ballot_sync(kFullMask, k0), m1 = __ballot_sync(kFullMask, k1);
        if ((m0 | m1) == 0u) return;
        uint32_t a0, a1, a2, a3, b0, b1, b2, b3;  // A = {cx0,cx1,cy0,cy1}, B = {cz0,cz1,w0,w1}
        unpack2_bits(A.x, a0, a1); unpack2_bits(A.y, a2, a3); unpack2_bits(B.x, b0, b1); unpack2_bits(B.y, b2, b3);
        const int pos = count + __popc(m0 & lt) + __popc(m1 & lt);
        if (k0 && pos < kCullCap) cull_store(mf, map, pos, __uint_as_float(a0), __uint_as_float(a2), __uint_as_float(b0), __uint_as_float(b2), src ? (int)src[2 * p] : 2 * p);
        const int pos1 = pos + (k0 ? 1 : 0);
        if (k1 && pos1 < kCullCap) cull_store(mf, map, pos1, __uint_as_float(a1), __uint_as_float(a3), __uint_as_float(b1), __uint_as_float(b3), src ? (int)src[2 * p + 1] : 2 * p + 1);
        count += __popc(m0) + __popc(m1);
    };
    int p0 = 0;
    // (two groups per iteration — two independent FFMA2 chains, one vote — measured slower: 2.70 -> 2.78 ms on C2)
#pragma unroll kCullUnroll
    for (; p0 < n_pairs; p0 += 32, va += 1024u) {  // n_pairs is a multiple of 32 (views of >= kCullMin entries are padded to 64)
        const ulonglong2 A = lds_u2(va), B = lds_u2(va + 16u);
        uint32_t kw0, kw1;
        cull_keep2(A, B, cone, kw0, kw1);
        if (__any_sync(kFullMask, (int)(kw0 | kw1) < 0)) {  // warp-uniform; ~10 % of the groups of 64 spheres have a survivor
            emit(A, B, kw0, kw1, p0 + lane);
            if (count > kCullCap) return count;
        }
    }
    // pad to a multiple of 32 with never-candidate entries
    const int padded = (count + 31) & ~31;
    const int q = count + lane;
    if (q < padded) {
        float* e = mf + (size_t)(q >> 1) * 8 + (q & 1);
        e[0] = 0.f; e[2] = 0.f; e[4] = 0.f; e[6] = -INFINITY;
    }
    __syncwarp();
    return count;
}

// Axis of the tile cone: the normalised sum of the active rays' directions (any unit vector would do: the half-angle is
// measured against it afterwards).  The three sums go through integer REDUX (fixed point, 2^-22: exact enough for an axis)
// instead of 15 shuffles.  One out-of-line copy; returns w = 0 when there is no active ray or the directions cancel.
static __device__ __noinline__ float4 cull_axis(float sx, float sy, float sz) {
    const float kScale = 4194304.0f, kInv = 1.0f / 4194304.0f;  // |sum over the warp| <= 32 K <= 2^7: fits 2^29
    const float fx = (float)__reduce_add_sync(kFullMask, __float2int_rn(sx * kScale)) * kInv;
    const float fy = (float)__reduce_add_sync(kFullMask, __float2int_rn(sy * kScale)) * kInv;
    const float fz = (float)__reduce_add_sync(kFullMask, __float2int_rn(sz * kScale)) * kInv;
    const float l2 = __fmaf_rn(fz, fz, __fmaf_rn(fy, fy, fx * fx));
    if (!(l2 > 1e-6f)) return make_float4(0.f, 0.f, 0.f, 0.f);
    const float inv = rsqrtf(l2);
    return make_float4(fx * inv, fy * inv, fz * inv, 1.0f);  // |a| = 1 +- 5e-7
}

// Tile cone of the active rays of the warp, then cull_pack.  Returns true when `cs` holds a usable mini-view of `count`
// entries.  All 32 lanes call it.
template <int K>
__device__ __forceinline__ bool cull_view(const ulonglong2* __restrict__ view, int n_pad, const ViewRay (&r)[K], const bool (&active)[K],
                                          const CullScratch& cs, int& count, const uint16_t* __restrict__ src = nullptr) {
    float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (active[k]) { sx += r[k].ux; sy += r[k].uy; sz += r[k].uz; }
    }
    const float4 a = cull_axis(sx, sy, sz);
    if (a.w == 0.0f) return false;
    float md = 1.0f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (active[k]) md = fminf(md, __fmaf_rn(r[k].uz, a.z, __fmaf_rn(r[k].uy, a.y, r[k].ux * a.x)));
    }
    count = cull_pack(view, n_pad, a.x, a.y, a.z, md, cs.mini, cs.map, src);  // -1: tile cone too wide
    return count >= 0 && count <= kCullCap;
}

// ------------------------------------------------------------------------------ CTA-level culling (streamed views)
// A scene whose views are streamed (C5: 4096 spheres, 9 views of 64 KB) pays one per-warp cull pass over the WHOLE view
// per (tile, view): 64 groups of 64 spheres.  In that kernel every warp of the CTA walks the same view at the same time
// and the CTA's tiles lie side by side (one batch = 4 x 4 tiles, or one strip of tiles when the rows are striped), so the
// CTA first culls the view ONCE against the bounding cone of all its rays — every warp tests its share of the view,
// two passes (count, then pack behind the lower warps' counts: ascending index) — and the per-warp culls then walk the
// CTA's survivors (a few hundred) instead of the view.  Same test, same conservativeness as the warp level.
#ifndef RTE_CTA_CAP
#define RTE_CTA_CAP 1024
#endif
constexpr int kCtaCap = RTE_CTA_CAP;  // survivors of the CTA-level cull; more: the warps cull the whole view as before
constexpr int kCtaCullBytes = kCtaCap * 16 + kCtaCap * 2 + 32 * 4 * 4;
struct CtaCull {
    ulonglong2* mini;  // kCtaCap entries in pair layout (shared memory, one per CTA)
    uint16_t* map;     // position -> index in the view
    int* part;         // [32 warps][4]: exchange of per-warp partial results
};

// Axis of the CTA's cone.  Called by every thread of the CTA; sx,sy,sz: this lane's sum of active directions.
static __device__ __noinline__ float4 cta_cull_axis(float sx, float sy, float sz, int* __restrict__ part) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const float kScale = 262144.0f, kInv = 1.0f / 262144.0f;  // |sum over the CTA| <= 1024 K <= 2^11: fits 2^29
    const int ix = __reduce_add_sync(kFullMask, __float2int_rn(sx * kScale));
    const int iy = __reduce_add_sync(kFullMask, __float2int_rn(sy * kScale));
    const int iz = __reduce_add_sync(kFullMask, __float2int_rn(sz * kScale));
    if (lane == 0) { part[warp * 4] = ix; part[warp * 4 + 1] = iy; part[warp * 4 + 2] = iz; }
    __syncthreads();
    const bool in = lane < nw;
    const float fx = (float)__reduce_add_sync(kFullMask, in ? part[lane * 4] : 0) * kInv;
    const float fy = (float)__reduce_add_sync(kFullMask, in ? part[lane * 4 + 1] : 0) * kInv;
    const float fz = (float)__reduce_add_sync(kFullMask, in ? part[lane * 4 + 2] : 0) * kInv;
    const float l2 = __fmaf_rn(fz, fz, __fmaf_rn(fy, fy, fx * fx));
    if (!(l2 > 1e-4f)) return make_float4(0.f, 0.f, 0.f, 0.f);  // (CTA-uniform)
    const float inv = rsqrtf(l2);
    return make_float4(fx * inv, fy * inv, fz * inv, 1.0f);
}

// Called by every thread of the CTA.  Returns the survivor count (CTA-uniform): -1 cone too wide, > kCtaCap overflow.
static __device__ __noinline__ int cta_cull_pack(const ulonglong2* __restrict__ view, int n_pad, float ax, float ay, float az, float md,
                                          ulonglong2* __restrict__ mini, uint16_t* __restrict__ map, int* __restrict__ part) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const uint32_t mdw = __reduce_min_sync(kFullMask, __float_as_uint(fmaxf(md, 0.0f)));
    if (lane == 0) part[warp * 4 + 3] = (int)mdw;
    __syncthreads();
    md = __uint_as_float(__reduce_min_sync(kFullMask, lane < nw ? (uint32_t)part[lane * 4 + 3] : 0x3f800000u));
    const float ct = md * 0.999998f - 2e-6f;
    if (!(ct >= kCullMinCos)) return -1;
    const float st2 = fmaxf(1.0f - ct * ct, 0.0f) * 1.00001f + 1e-12f;
    const CullCone cone = make_cull_cone(ax, ay, az, ct, st2);
    const uint32_t lt = (1u << lane) - 1u;
    const int groups = n_pad >> 6;  // of 32 pairs (n_pad is a multiple of 64)
    const int per = (groups + nw - 1) / nw, g0 = min(warp * per, groups), g1 = min(g0 + per, groups);
    const uint32_t v0 = (uint32_t)__cvta_generic_to_shared(view) + 32u * (uint32_t)lane;
    int cnt = 0;
#pragma unroll 1
    for (int g = g0; g < g1; ++g) {  // pass 1: count
        const uint32_t va = v0 + 1024u * (uint32_t)g;
        const ulonglong2 A = lds_u2(va), B = lds_u2(va + 16u);
        uint32_t kw0, kw1;
        cull_keep2(A, B, cone, kw0, kw1);
        cnt += __popc(__ballot_sync(kFullMask, (int)kw0 < 0)) + __popc(__ballot_sync(kFullMask, (int)kw1 < 0));
    }
    if (lane == 0) part[warp * 4] = cnt;
    __syncthreads();
    const int c = lane < nw ? part[lane * 4] : 0;
    const int total = __reduce_add_sync(kFullMask, c);
    if (total > kCtaCap) return total;
    int count = __reduce_add_sync(kFullMask, lane < warp ? c : 0);  // behind the survivors of the lower warps
    float* mf = reinterpret_cast<float*>(mini);
#pragma unroll 1
    for (int g = g0; g < g1 && cnt > 0; ++g) {  // pass 2: pack
        const uint32_t va = v0 + 1024u * (uint32_t)g;
        const ulonglong2 A = lds_u2(va), B = lds_u2(va + 16u);
        uint32_t kw0, kw1;
        cull_keep2(A, B, cone, kw0, kw1);
        const bool k0 = (int)kw0 < 0, k1 = (int)kw1 < 0;
        const uint32_t m0 = __ballot_sync(kFullMask, k0), m1 = __ballot_sync(kFullMask, k1);
        if ((m0 | m1) == 0u) continue;
        uint32_t a0, a1, a2, a3, b0, b1, b2, b3;
        unpack2_bits(A.x, a0, a1); unpack2_bits(A.y, a2, a3); unpack2_bits(B.x, b0, b1); unpack2_bits(B.y, b2, b3);
        const int p = g * 32 + lane;
        const int pos = count + __popc(m0 & lt) + __popc(m1 & lt);
        if (k0) cull_store(mf, map, pos, __uint_as_float(a0), __uint_as_float(a2), __uint_as_float(b0), __uint_as_float(b2), 2 * p);
        if (k1) cull_store(mf, map, pos + (k0 ? 1 : 0), __uint_as_float(a1), __uint_as_float(a3), __uint_as_float(b1), __uint_as_float(b3), 2 * p + 1);
        count += __popc(m0) + __popc(m1);
    }
    if (warp == 0) {  // pad to a multiple of 64 with never-candidate entries
        const int padded = (total + 63) & ~63;
        for (int q = total + lane; q < padded; q += 32) {
            float* e = mf + (size_t)(q >> 1) * 8 + (q & 1);
            e[0] = 0.f; e[2] = 0.f; e[4] = 0.f; e[6] = -INFINITY;
        }
    }
    __syncthreads();
    return total;
}

// CTA cone of the active rays, then cta_cull_pack.  Every thread of the CTA calls it (warps without work pass no active ray).
template <int K>
__device__ __forceinline__ bool cta_cull_view(const ulonglong2* __restrict__ view, int n_pad, const ViewRay (&r)[K], const bool (&active)[K],
                                              const CtaCull& cc, int& count) {
    float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (active[k]) { sx += r[k].ux; sy += r[k].uy; sz += r[k].uz; }
    }
    const float4 a = cta_cull_axis(sx, sy, sz, cc.part);
    if (a.w == 0.0f) return false;
    float md = 1.0f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (active[k]) md = fminf(md, __fmaf_rn(r[k].uz, a.z, __fmaf_rn(r[k].uy, a.y, r[k].ux * a.x)));
    }
    count = cta_cull_pack(view, n_pad, a.x, a.y, a.z, md, cc.mini, cc.map, cc.part);
    return count >= 0 && count <= kCtaCap;
}

// ------------------------------------------------------------------------------ candidate queue
// With ~2 line candidates per ray and 32 lanes, almost every 32-sphere block has a candidate in SOME lane:
// resolving candidates where they are found runs the divergent path ~30 times per scan with 1-2 lanes active.
// Instead every lane appends its candidates, (ray k << 8) | (sphere index - window start) in 16 bits, to a small private queue during
// the scan, and the warp resolves the queues lane-parallel in one compact loop per window: trip count =
// the longest queue in the warp, one copy of the resolution code for all K rays.
constexpr int kWindow = 256;  // spheres between two resolutions (+ warp vote for any-hit rays): bounds the queue
template <int K>
struct CandQueue {
    uint16_t q[kWindow * K];  // (ray k << 8) | (sphere index - window start); local memory, only written entries are touched
    // `n` is passed in and out so that it stays in a register (the dynamically indexed array lives in local memory)
    // rel = base - window start, left = ns - base; the candidate's cold entry is prefetched into L1 now, the resolution at
    // the end of the window reads it some hundred cycles later
    // (`map`: the scan walks a culled mini-view, positions translate to view indices)
    __device__ __forceinline__ int push_block(uint32_t c, int rel, int k, int left, int n, int wbase, const float* __restrict__ cold,
                                              const uint16_t* __restrict__ map) {
        while (c) {  // short, divergent
            const int i = __clz(c);
            c &= ~(0x80000000u >> i);
            if (i < left) {
                q[n++] = (uint16_t)((k << 8) | (rel + i));
                const int idx = map ? (int)map[wbase + rel + i] : wbase + rel + i;
                asm volatile("prefetch.global.L1 [%0];" ::"l"(cold + (size_t)(idx >> 1) * 8));
            }
        }
        return n;
    }
};

template <int K, class T>
__device__ __forceinline__ T pick(const T (&a)[K], int k) {
    T v = a[0];
#pragma unroll
    for (int i = 1; i < K; ++i) if (k == i) v = a[i];
    return v;
}

// More than two pixels per thread: the K rays of a thread share the tile cull (one pass over the view for 32 K rays) but are
// scanned one after the other by the single-ray code (one copy, in a loop): no K-fold registers, no K-fold code.
// Measured on B200 (C2 at 4K, -DRTE_VIEW_PIXELS=4 / 8): bit-identical, but 4.72 / 5.29 ms against 2.95 ms with one pixel per
// thread — the per-pixel FP64 state of K pixels lives in local memory (2.6 / 4.1 KB stack) and that costs more than three
// quarters of the cull passes save.  Kept as an option, not the default.
template <int K> struct SeqScan { static constexpr bool value = K > 2; };

// The window loop of closest_spheres_view: scan (FP32, K rays share every LDS), queue, lane-parallel resolution.
template <int K, bool kTri>
__device__ __forceinline__ void closest_scan(const DevScene& sc, const ulonglong2* __restrict__ view, const uint16_t* __restrict__ map, int ns, int ns_pad,
                                             const float* __restrict__ vf, const ViewBounds& vb, d3 o, const d3 (&d)[K], const ViewRay (&r)[K],
                                             const bool (&alive)[K], bool no_filter, Hit (&best)[K], float (&best_f)[K], bool (&found)[K], uint32_t& confirms) {
    CandQueue<K> Q;
    int qn = 0;
    int base = 0;
#pragma unroll 1
    while (base < ns_pad) {
        const int wbase = base, stop = min(base + kWindow, ns_pad);
#pragma unroll 1
        do {
            uint32_t cand[K];
            view_block_masks<K>(view + base, r, cand);
            uint32_t any = 0;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                cand[k] = no_filter ? 0xFFFFFFFFu : cand[k];
                if (!alive[k]) cand[k] = 0u;
                any |= cand[k];
            }
            if (any) {  // ~5 % of the blocks
#pragma unroll
                for (int k = 0; k < K; ++k) qn = Q.push_block(cand[k], base - wbase, k, ns - base, qn, wbase, vf, map);
            }
            base += kViewBlock;
        } while (base < stop);
        const int nmax = __reduce_max_sync(kFullMask, qn);
#pragma unroll 1
        for (int e = 0; e < nmax; ++e) {  // lane-parallel resolution, ascending index per ray
            if (e >= qn) continue;
            const uint32_t ent = Q.q[e];
            const int k = (int)(ent >> 8), pos = wbase + (int)(ent & 0xFFu);
            const int idx = map ? (int)map[pos] : pos;
            const d3 dk = d[k];  // FP64 ray data is indexed dynamically: it lives in local memory, not in K selects
            if (!no_filter) {
                const ViewChord ch = view_chord(vf, idx, pick<K>(r, k), vb);
                // the line misses it, it lies behind the camera, or its near root cannot beat the best so far
                if (ch.miss || ch.sc + ch.hc_hi + vb.slack < 0.0f || ch.sc - ch.hc_hi - vb.slack > pick<K>(best_f, k)) continue;
            }
            double t;
            ++confirms;
            if (!confirm_prim<kTri>(sc, idx, o, dk, t)) continue;
#pragma unroll
            for (int i = 0; i < K; ++i) {
                if (i == k && (!found[i] || t < best[i].t)) {  // strict '<', RE/Scene.h:223
                    found[i] = true; best[i].t = t; best[i].kind = kTri ? RTE_HIT_TRIANGLE : RTE_HIT_SPHERE; best[i].idx = (uint32_t)idx;
                    best_f[i] = __double2float_ru(t);
                }
            }
        }
        qn = 0;
    }
}

// Scene::IntersectClosest over the spheres (kTri: the triangles, through their bounding spheres) for K PRIMARY rays of one
// thread (view 0, common origin o).  best/found carry the result of the classes scanned before (canonical order, strict '<').
template <int K, bool kTri = false>
__device__ __forceinline__ void closest_spheres_view(const DevScene& sc, const ulonglong2* __restrict__ view, const float* __restrict__ vf, d3 o,
                                                     const d3 (&d)[K], const bool (&alive)[K], bool no_filter, Hit (&best)[K], bool (&found)[K],
                                                     uint32_t& confirms, const CullScratch* cull = nullptr, uint32_t* executed = nullptr,
                                                     const CtaCull* cc = nullptr) {
    // view: hot array (shared), vf: cold array (global); executed: per-lane count of FP32 tests really run (optional)
    // cc: CTA-level cull first (streamed views; then EVERY thread of the CTA is here)
    const ViewBounds vb = view_bounds(view_extent<kTri>(sc), o, 0.0, false);
    ViewRay r[K];
    float best_f[K];
#pragma unroll
    for (int k = 0; k < K; ++k) { r[k] = make_view_ray(d[k]); best_f[k] = found[k] ? __double2float_ru(best[k].t) : INFINITY; }
    int ns = view_count<kTri>(sc), ns_pad = view_padded<kTri>(sc);
    const uint16_t* map = nullptr;  // culled: positions of the mini-view -> indices of the view
    if (cc && !no_filter && (ns_pad & 63) == 0) {
        int cnt = 0;
        if (executed) *executed += (uint32_t)(ns_pad >> 3) / (blockDim.x >> 5);  // two passes over this warp's share of the view
        if (cta_cull_view<K>(view, ns_pad, r, alive, *cc, cnt)) { view = cc->mini; map = cc->map; ns = cnt; ns_pad = (cnt + 63) & ~63; }
    }
    if (cull && !no_filter && ns_pad >= kCullMin && (ns_pad & 63) == 0) {
        int cnt = 0;
        if (executed) *executed += (uint32_t)(ns_pad >> 4);  // the warp's cull pass: ns_pad/32 cone-cone tests per lane, ~2x a ray test
        if (cull_view<K>(view, ns_pad, r, alive, *cull, cnt, map)) { view = cull->mini; map = cull->map; ns = cnt; ns_pad = (cnt + 31) & ~31; }
    }
    uint32_t n_alive = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) n_alive += alive[k] ? 1u : 0u;
    if (executed) *executed += n_alive * (uint32_t)ns_pad;
    if (SeqScan<K>::value) {
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            const bool a1[1] = {alive[k]};
            if (!__any_sync(kFullMask, a1[0])) continue;
            const d3 d1[1] = {d[k]};
            const ViewRay r1[1] = {r[k]};
            Hit b1[1] = {best[k]};
            bool f1[1] = {found[k]};
            float bf1[1] = {best_f[k]};
            closest_scan<1, kTri>(sc, view, map, ns, ns_pad, vf, vb, o, d1, r1, a1, no_filter, b1, bf1, f1, confirms);
            best[k] = b1[0]; found[k] = f1[0];
        }
    } else {
        closest_scan<K, kTri>(sc, view, map, ns, ns_pad, vf, vb, o, d, r, alive, no_filter, best, best_f, found, confirms);
    }
}

// Scene::IntersectClosest for K primary rays: spheres through view 0, planes, then triangles (through their view 0 when
// `tri_view`, else with the generic FP32 filter).  The view pointers are always shared-memory addresses (never selected
// against NULL: the scans must compile to LDS).
template <int K>
__device__ __forceinline__ void closest_hit_view(const DevScene& sc, const ulonglong2* __restrict__ view0, const ulonglong2* __restrict__ tview0,
                                                 const float* cold0, const float* tcold0, bool tri_view, const float4* __restrict__ s_tri, d3 o, const d3 (&d)[K], const bool (&alive)[K],
                                                 bool no_filter, Hit (&best)[K], uint32_t& confirms) {
    bool found[K];
#pragma unroll
    for (int k = 0; k < K; ++k) { best[k].t = 0.0; best[k].kind = RTE_HIT_NONE; best[k].idx = 0; found[k] = false; }
    if (sc.ns > 0) closest_spheres_view<K>(sc, view0, cold0, o, d, alive, no_filter, best, found, confirms);
    if (sc.np > 0 || (sc.nt > 0 && !tri_view)) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const Ray32 r = make_ray32(sc, o, d[k], no_filter);
            closest_planes_triangles(sc, s_tri, r, o, d[k], alive[k], best[k], found[k], confirms, !tri_view);
        }
    }
    if (sc.nt > 0 && tri_view) closest_spheres_view<K, true>(sc, tview0, tcold0, o, d, alive, no_filter, best, found, confirms);
}

// One forward shadow ray of the reference (RE/Scene.h:86-104) plus what the reversed scan needs.
struct ShadowRay {
    d3 o, d;          // shadow origin so = P + N*bias and unit direction L: every FP64 confirmation uses them untouched
    double dist;      // |Lp - P|
    double maxDist;   // dist - bias
    double ndl;       // N.L (> 0 when cast)
    int self;         // sphere the shading point lies on, hit from outside (-1: none): its exact test cannot succeed
    bool cast;
};

// Any-hit shadow query (see occluded() in rte_trace.cuh for the equivalence argument) for K rays of one thread
// towards the light `Lp`, sphere phase scanned through that light's view: the line is walked from the light,
// direction -L, parameter s; the forward distance from the shadow origin is  t = (dist - bias*ndl) - s.
//
// Candidates are resolved in canonical order, each one either
//   * skipped: it is the front-face sphere the shading point lies on (both roots <= -bias: the exact test misses),
//   * decided in FP32 when the interval arithmetic is conclusive: a certain occluder with t in (t_safe, maxDist),
//     or certainly outside (behind the light / beyond the shaded point),
//   * or confirmed in FP64 with the reference's own operation order.
constexpr int kListCap = 24;  // in-range sphere candidates remembered per shadow ray (for march_list)
template <int K>
struct RayLists {
    uint16_t idx[K][kListCap];
    int n[K];          // > kListCap: overflow, the list is unusable
};

// Scene::computeTransmittance, RE/Scene.h:35-77, literally — but every closest-hit query of the march only looks at the
// spheres of `list` (all spheres whose chord overlaps the ray's range: the march moves its origin along the same line, so
// no other sphere can ever be hit) and at the planes.  Scenes of the fast path have no triangles and no transparency.
static __device__ __noinline__ double march_list(const DevScene& sc, const uint16_t* list, int n, d3 o, d3 d, double maxDist, double bias) {
    double T = 1.0, traveled = 0.0;
    int safety = 64;
    while (safety-- > 0 && T > 1e-4 && traveled < maxDist) {
        bool found = false;
        double bt = 0.0;
        int mat = sc.nm;
#pragma unroll 1
        for (int i = 0; i < n; ++i) {  // ascending sphere index, strict '<' (RE/Scene.h:218-257)
            double t;
            const int k = list[i];
            if (sphere_exact(sc.sph64[k], o, d, t) && (!found || t < bt)) { found = true; bt = t; mat = sc.sph_mat[k]; }
        }
#pragma unroll 1
        for (int i = 0; i < sc.np; ++i) {
            double t;
            if (plane_exact(sc.planes[i], o, d, t) && (!found || t < bt)) { found = true; bt = t; mat = sc.planes[i].mat; }
        }
        if (!found) break;
        const double t = bt;
        if (t <= 0.0) { o = o + d * bias; traveled += bias; continue; }
        if (t <= bias) { o = (o + d * t) + d * bias; traveled += t + bias; continue; }
        if (traveled + t >= maxDist) break;
        const int m = (mat >= 0 && mat < sc.nm) ? mat : sc.nm;
        T *= std_clamp(sc.materials[m].transparency, 0.0, 1.0);
        const d3 newOrigin = (o + d * t) + d * bias;
        traveled += t + bias;
        o = newOrigin;
    }
    return std_clamp(T, 0.0, 1.0);
}

// Per-ray FP32 constants of a reversed shadow scan.
struct OccRay32 { float tref, tmax_lo, tmax_hi; };

// The window loop of occluded_spheres_view: scan, queue, lane-parallel resolution in canonical order per ray, warp vote per window.
template <int K, bool kTri>
__device__ __forceinline__ void occluded_scan(const DevScene& sc, const ulonglong2* __restrict__ view, const uint16_t* __restrict__ map, int ns, int ns_pad,
                                              const float* __restrict__ vf, const ViewBounds& vb, const ShadowRay (&ray)[K], const ViewRay (&r)[K],
                                              const OccRay32 (&q32)[K], float tsafe_hi, bool skip_self, bool no_filter, OccState (&st)[K], uint32_t (&first)[K],
                                              RayLists<K>& lists, uint32_t& confirms, uint32_t& executed) {
    CandQueue<K> Q;
    int qn = 0;
    int base = 0;
#pragma unroll 1
    while (base < ns_pad) {
        const int wbase = base, stop = min(base + kWindow, ns_pad);
#pragma unroll 1
        do {
            uint32_t cand[K];
            view_block_masks<K>(view + base, r, cand);
            if (!kTri) executed += kViewBlock;
            uint32_t any = 0;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                cand[k] = no_filter ? 0xFFFFFFFFu : cand[k];
                if (st[k].done) cand[k] = 0u;
                any |= cand[k];
            }
            if (any) {  // ~5 % of the blocks
#pragma unroll
                for (int k = 0; k < K; ++k) qn = Q.push_block(cand[k], base - wbase, k, ns - base, qn, wbase, vf, map);
            }
            base += kViewBlock;
        } while (base < stop);
        const int nmax = __reduce_max_sync(kFullMask, qn);
#pragma unroll 1
        for (int e = 0; e < nmax; ++e) {  // lane-parallel resolution, canonical order per ray
            if (e >= qn) continue;
            const uint32_t ent = Q.q[e];
            const int k = (int)(ent >> 8), pos = wbase + (int)(ent & 0xFFu);
            const int idx = map ? (int)map[pos] : pos;
            bool dn = false;
#pragma unroll
            for (int i = 0; i < K; ++i) if (i == k) dn = st[i].done;
            if (dn) continue;  // decided by an earlier entry of this window
            bool certain = false;
            if (!no_filter) {
                if (skip_self && idx == ray[k].self) continue;
                const ViewChord ch = view_chord(vf, idx, pick<K>(r, k), vb);
                if (ch.miss) continue;
                const float tc = q32[k].tref - ch.sc;  // forward distance of the chord centre
                const float t0_lo = tc - ch.hc_hi - vb.slack, t0_hi = tc - ch.hc_lo + vb.slack, t1_hi = tc + ch.hc_hi + vb.slack;
                if (t1_hi < 0.0f || t0_lo > q32[k].tmax_hi) continue;  // certainly no hit in (1e-6, maxDist), none <= bias
                certain = !kTri && ch.hc_lo > 0.0f && t0_lo > tsafe_hi && t0_hi < q32[k].tmax_lo;  // certainly t in (t_safe, maxDist)
            }
            if (!kTri) {  // remember every in-range candidate of the ray (what an exact march over this ray could ever hit)
                const int ln = lists.n[k];
                if (ln < kListCap && idx < 65536) lists.idx[k][ln] = (uint16_t)idx;
                lists.n[k] = (ln < kListCap && idx < 65536) ? ln + 1 : kListCap + 1;
            }
            double t = 0.0;
            bool hit = false;
            if (!certain) {
                ++confirms;
                hit = confirm_prim<kTri>(sc, idx, ray[k].o, ray[k].d, t);
            }
#pragma unroll
            for (int i = 0; i < K; ++i) {
                if (i != k) continue;
                if (certain) {
                    if (!st[i].res.occluded) { st[i].res.occluded = true; first[i] = (uint32_t)idx + 1u; }
                    if (!st[i].res.near) st[i].done = true;
                } else if (hit) {
                    if (kTri) st[i].on_t(t, first[i], (uint32_t)idx);
                    else st[i].on_t_view(t, first[i], (uint32_t)idx);
                }
            }
        }
        qn = 0;
        bool all_done = true;
#pragma unroll
        for (int k = 0; k < K; ++k) all_done = all_done && st[k].done;
        if (__all_sync(kFullMask, all_done)) break;
    }
}

// kTri = true: the triangle phase of the same query, through the light's view of the triangles' bounding spheres; `st`
// continues from the sphere and plane phases (canonical order), every candidate is confirmed in FP64 (tri_exact).
template <int K, bool kTri = false>
__device__ __forceinline__ void occluded_spheres_view(const DevScene& sc, const ulonglong2* __restrict__ view, const float* __restrict__ vf, d3 Lp,
                                                      const ShadowRay (&ray)[K], double bias, bool no_filter, OccState (&st)[K],
                                                      RayLists<K>& lists, uint32_t& confirms, uint32_t& executed,
                                                      const CullScratch* cull = nullptr, const CtaCull* cc = nullptr) {
    uint32_t first[K];
    bool scanning[K];
    const int ns_view = view_count<kTri>(sc);
    int ns = ns_view, ns_pad = view_padded<kTri>(sc);
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (!kTri) { lists.n[k] = 0; st[k] = occ_begin(ray[k].maxDist, bias, ray[k].cast); }
        first[k] = (uint32_t)ns_view;
        scanning[k] = kTri ? !st[k].done : ray[k].cast;
    }
    if (ns > 0) {
        const ViewBounds vb = view_bounds(view_extent<kTri>(sc), Lp, bias, true);
        const bool skip_self = !kTri && bias >= 1e-5;
        ViewRay r[K];
        OccRay32 q32[K];
        const float tsafe_hi = __double2float_ru(130.0 * bias) + vb.slack;
        const float bias_hi = __double2float_ru(bias) + vb.slack;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            r[k] = make_view_ray(-ray[k].d);
            q32[k].tref = (float)(ray[k].dist - bias * ray[k].ndl);
            q32[k].tmax_lo = __double2float_rd(ray[k].maxDist) - vb.slack;
            q32[k].tmax_hi = fmaxf(__double2float_ru(ray[k].maxDist) + vb.slack, bias_hi);  // also keeps every t <= bias (`near`)
        }
        const uint16_t* map = nullptr;  // culled: positions of the mini-view -> indices of the view
        if (cc && !no_filter && (ns_pad & 63) == 0) {  // CTA-level cull first (every thread of the CTA is here)
            bool act[K];
#pragma unroll
            for (int k = 0; k < K; ++k) act[k] = !st[k].done;
            int cnt = 0;
            if (!kTri) executed += (uint32_t)(ns_pad >> 3) / (blockDim.x >> 5);
            if (cta_cull_view<K>(view, ns_pad, r, act, *cc, cnt)) { view = cc->mini; map = cc->map; ns = cnt; ns_pad = (cnt + 63) & ~63; }
        }
        if (cull && !no_filter && ns_pad >= kCullMin && (ns_pad & 63) == 0) {
            bool act[K];
#pragma unroll
            for (int k = 0; k < K; ++k) act[k] = !st[k].done;
            int cnt = 0;
            if (!kTri) executed += (uint32_t)(ns_pad >> 4);  // the warp's cull pass: ns_pad/32 cone-cone tests per lane, ~2x a ray test
            if (cull_view<K>(view, ns_pad, r, act, *cull, cnt, map)) { view = cull->mini; map = cull->map; ns = cnt; ns_pad = (cnt + 31) & ~31; }
#ifdef RTE_CULL_DEBUG
            if ((threadIdx.x & 31) == 0) {
                if (RTE_CULL_DEBUG == 1 && cnt == -1) confirms += 1;                 // tile cone too wide
                if (RTE_CULL_DEBUG == 2 && cnt > kCullCap) confirms += 1;            // too many survivors
                if (RTE_CULL_DEBUG == 3 && cnt >= 0 && cnt <= kCullCap) confirms += 1;  // culled
                if (RTE_CULL_DEBUG == 4 && cnt >= 0 && cnt <= kCullCap) confirms += (uint32_t)cnt;  // survivors of culled scans
            }
#endif
        }
        if (SeqScan<K>::value) {
#pragma unroll 1
            for (int k = 0; k < K; ++k) {
                OccState s1[1] = {st[k]};
                if (__all_sync(kFullMask, s1[0].done)) continue;
                const ShadowRay y1[1] = {ray[k]};
                const ViewRay r1[1] = {r[k]};
                const OccRay32 q1[1] = {q32[k]};
                uint32_t f1[1] = {first[k]};
                RayLists<1> l1;
                l1.n[0] = lists.n[k];
                occluded_scan<1, kTri>(sc, view, map, ns, ns_pad, vf, vb, y1, r1, q1, tsafe_hi, skip_self, no_filter, s1, f1, l1, confirms, executed);
                st[k] = s1[0]; first[k] = f1[0];
                if (!kTri) {
                    const int n1 = l1.n[0];
                    lists.n[k] = n1;
                    for (int i = 0; i < n1 && i < kListCap; ++i) lists.idx[k][i] = l1.idx[0][i];
                }
            }
        } else {
            occluded_scan<K, kTri>(sc, view, map, ns, ns_pad, vf, vb, ray, r, q32, tsafe_hi, skip_self, no_filter, st, first, lists, confirms, executed);
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (scanning[k]) {
            if (kTri) st[k].res.tri_tests = first[k];
            else st[k].res.sph_tests = first[k];
        }
    }
}

// Spheres through the light's view, then planes, then triangles (through the light's triangle view when `tri_view`,
// else with the generic FP32 filter) — general kernel.
template <int K>
__device__ __forceinline__ void occluded_view(const DevScene& sc, const ulonglong2* __restrict__ view, const ulonglong2* __restrict__ tview,
                                              const float* cold, const float* tcold, bool tri_view, const float4* __restrict__ s_tri, d3 Lp, const ShadowRay (&ray)[K], double bias,
                                              bool no_filter, OccResult (&out)[K], RayLists<K>& lists, uint32_t& confirms, uint32_t& executed) {
    OccState st[K];
    occluded_spheres_view<K>(sc, view, cold, Lp, ray, bias, no_filter, st, lists, confirms, executed);  // ns == 0: only initialises st
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (st[k].res.near && st[k].res.occluded) st[k].done = true;  // the march decides; planes/triangles are not counted after an occluder
        if (sc.np > 0 || (sc.nt > 0 && !tri_view)) {
            if (!__all_sync(kFullMask, st[k].done)) {
                const Ray32 r32 = make_ray32(sc, ray[k].o, ray[k].d, no_filter);
                occ_planes_triangles(sc, s_tri, r32, ray[k].o, ray[k].d, st[k], confirms, !tri_view);
            }
        }
    }
    if (sc.nt > 0 && tri_view) {
        bool all_done = true;
#pragma unroll
        for (int k = 0; k < K; ++k) all_done = all_done && st[k].done;
        if (!__all_sync(kFullMask, all_done))
            occluded_spheres_view<K, true>(sc, tview, tcold, Lp, ray, bias, no_filter, st, lists, confirms, executed);
    }
#pragma unroll
    for (int k = 0; k < K; ++k) out[k] = st[k].res;
}

}  // namespace rte
