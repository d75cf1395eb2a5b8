// rte_views.cuh — origin-specialised sphere filters ("views") with packed FP32x2 FMAs.
//
// Every ray of the frame starts at the camera (primary rays) or ends at a light (shadow rays, RE/Scene.h:
// 86-104; "RE/" = /root/reference/RaytracingEngine/).  A shadow ray is filtered *reversed*, from the light
// towards the shaded point, so both families share their origin O_j across all pixels.  Everything in the
// line–sphere test that depends only on (sphere i, origin j) is hoisted into a per-frame "view" array
//
//     f_ij  = C_i - O_j                       (FP64, rounded to FP32)
//     ng_ij = -( |f_ij|^2 - r_i^2 ) + M_ij     (FP64, rounded UP to FP32)
//
// and the per-ray work left is   b = f.u  (3 FMA),  h = b*b + ng  (1 FMA):  h >= 0  <=>  the ray's LINE may
// touch the sphere.  Two spheres are processed per instruction with fma.rn.f32x2 (SASS FFMA2, sm_100+), the
// sign bits of h are shifted into a 32-bit mask (1 SHF per test) and only the rare set bits are looked at
// again: FP32 interval pruning, then the FP64 confirmation in the reference's own operation order
// (rte_device.cuh).  4 FMA-pipe lanes-ops + 1 ALU op + 1 LDS.128 per TWO tests' worth of data per test.
//
// The filter is conservative: with u = fl32(d), |d| = 1 (FP64-normalised by the caller),
// F = |f_ij|, eps = 2^-24:   |b~ - b| <= 5 eps F,  |b~^2 - b^2| <= 10.x eps F^2,  so with
//     M_ij = 16 eps F^2   (+ 2 r bias + bias^2 for light views: the true shadow line starts at P + N*bias and
//                          misses the light by at most `bias`, RE/Scene.h:104)
// a sphere whose exact test can succeed always has h~ >= 0.  It may accept extra spheres (cost only).
#pragma once
#include "rte_trace.cuh"

namespace rte {

constexpr int kViewBlock = 32;  // spheres per candidate mask (ns_pad is a multiple of this)

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
// One view = ns_pad/2 sphere pairs of 32 bytes:  {fx0,fx1, fy0,fy1} {fz0,fz1, ng0,ng1}.
__host__ __device__ __forceinline__ size_t view_bytes(int ns_pad) { return (size_t)ns_pad * 16; }

struct ViewEntry { float fx, fy, fz, ng; };
__device__ __forceinline__ ViewEntry view_entry(const float* __restrict__ view, int idx) {
    const float* p = view + (size_t)(idx >> 1) * 8 + (idx & 1);
    ViewEntry e;
    e.fx = p[0]; e.fy = p[2]; e.fz = p[4]; e.ng = p[6];
    return e;
}

// Builds view j (0 = camera, 1.. = lights) for every sphere.  One thread per (view, sphere).
__global__ void __launch_bounds__(256) build_views_kernel(const DevScene sc, const DevCamera cam, double bias, float* __restrict__ views) {
    const int ns_pad = sc.ns_pad, nv = 1 + sc.nl;
    const int total = nv * ns_pad;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
        const int j = t / ns_pad, i = t - j * ns_pad;
        float fx = 0.f, fy = 0.f, fz = 0.f, ng = -INFINITY;  // never-candidate padding: h = -inf
        if (i < sc.ns) {
            const d3 O = j == 0 ? ld3(cam.pos) : ld3(sc.lights[j - 1].pos);
            const double4 s = sc.sph64[i];
            const d3 f = mk(s.x, s.y, s.z) - O;
            const double F2 = dot(f, f);
            const double r = fabs(s.w);
            double M = 16.0 * (double)kEps32 * F2 + 1e-30;
            if (j > 0) M += 2.0 * r * bias + bias * bias;
            fx = (float)f.x; fy = (float)f.y; fz = (float)f.z;
            ng = __double2float_ru((r * r - F2) + M);
            if (!(ng == ng)) ng = INFINITY;  // non-finite input: always a candidate (the FP64 test decides)
        }
        float* p = views + ((size_t)j * ns_pad + (size_t)(i & ~1)) * 4 + (i & 1);
        p[0] = fx; p[2] = fy; p[4] = fz; p[6] = ng;
    }
}

// ------------------------------------------------------------------------------ the scan
// Filters spheres [base, base+32) of one view against the unit direction (ux,uy,uz) given as duplicated
// pairs; returns the candidate bits, bit (31-k) <-> sphere base+k.
__device__ __forceinline__ uint32_t view_block_mask(const ulonglong2* __restrict__ v, f32x2 ux2, f32x2 uy2, f32x2 uz2) {
    uint32_t m = 0;
#pragma unroll
    for (int p = 0; p < kViewBlock / 2; ++p) {
        const ulonglong2 A = v[2 * p], B = v[2 * p + 1];
        f32x2 b = mul2(A.x, ux2);
        b = fma2(A.y, uy2, b);
        b = fma2(B.x, uz2, b);
        const f32x2 h = fma2(b, b, B.y);
        uint32_t h0, h1;
        unpack2_bits(h, h0, h1);
        m = __funnelshift_l(h0, m, 1);  // m = (m << 1) | sign(h0)
        m = __funnelshift_l(h1, m, 1);
    }
    return ~m;  // sign clear (h >= 0, or NaN with the sign clear) -> candidate
}

struct ViewRay {
    f32x2 ux2, uy2, uz2;  // fl32 of the unit direction the view is scanned with, duplicated
    float ux, uy, uz;
    float slack;          // absolute slack on distances along the ray
};

__device__ __forceinline__ ViewRay make_view_ray(d3 u, float S) {
    ViewRay r;
    r.ux = (float)u.x; r.uy = (float)u.y; r.uz = (float)u.z;
    r.ux2 = pack2(r.ux, r.ux); r.uy2 = pack2(r.uy, r.uy); r.uz2 = pack2(r.uz, r.uz);
    r.slack = 16.0f * kEps32 * S + 1e-6f;
    return r;
}

// [s_lo, s_hi] (conservatively widened) where the scan line is inside sphere `idx`
__device__ __forceinline__ void view_interval(const float* __restrict__ view, int idx, const ViewRay& r, float& s_lo, float& s_hi) {
    const ViewEntry e = view_entry(view, idx);
    const float b = __fmaf_rn(e.fz, r.uz, __fmaf_rn(e.fy, r.uy, e.fx * r.ux));
    const float h = __fmaf_rn(b, b, e.ng);
    const float sq = sqrtf(fmaxf(h, 0.0f)) * 1.000001f;
    s_lo = b - sq - r.slack;
    s_hi = b + sq + r.slack;
}

// Scene::IntersectClosest over the spheres for a PRIMARY ray (view 0).  `best`/`found` carry on.
__device__ __forceinline__ void closest_spheres_view(const DevScene& sc, const ulonglong2* __restrict__ view, d3 o, d3 d, bool alive,
                                                     bool no_filter, Hit& best, bool& found, uint32_t& confirms) {
    const float S = sqrtf((float)dot(o, o)) * 1.0000002f + sc.sph_extent;
    const ViewRay r = make_view_ray(d, S);
    const float* vf = reinterpret_cast<const float*>(view);
    const int ns = sc.ns, ns_pad = sc.ns_pad;
    float best_f = INFINITY;
#pragma unroll 1
    for (int base = 0; base < ns_pad; base += kViewBlock) {
        uint32_t cand = view_block_mask(view + base, r.ux2, r.uy2, r.uz2);
        if (no_filter) cand = 0xFFFFFFFFu;
        if (!alive) cand = 0u;
        while (cand) {  // rare, divergent
            const int k = __clz(cand);
            cand &= ~(0x80000000u >> k);
            const int idx = base + k;
            if (idx >= ns) break;
            if (!no_filter) {
                float s_lo, s_hi;
                view_interval(vf, idx, r, s_lo, s_hi);
                if (s_hi < 0.0f || s_lo > best_f) continue;  // behind the camera / cannot beat the best so far
            }
            double t;
            ++confirms;
            if (sphere_exact(sc.sph64[idx], o, d, t) && (!found || t < best.t)) {  // strict '<', RE/Scene.h:223
                found = true; best.t = t; best.kind = RTE_HIT_SPHERE; best.idx = (uint32_t)idx;
                best_f = __double2float_ru(t);
            }
        }
    }
}

// Scene::IntersectClosest for a primary ray: spheres through view 0, then planes and triangles.
__device__ __forceinline__ void closest_hit_view(const DevScene& sc, const ulonglong2* __restrict__ view0, const float4* __restrict__ s_tri,
                                                 d3 o, d3 d, bool alive, bool no_filter, Hit& best, uint32_t& confirms) {
    best.t = 0.0; best.kind = RTE_HIT_NONE; best.idx = 0;
    bool found = false;
    if (sc.ns > 0) closest_spheres_view(sc, view0, o, d, alive, no_filter, best, found, confirms);
    if (sc.np > 0 || sc.nt > 0) {
        const Ray32 r = make_ray32(sc, o, d, no_filter);
        closest_planes_triangles(sc, s_tri, r, o, d, alive, best, found, confirms);
    }
}

// Any-hit shadow query (see occluded() in rte_trace.cuh for the equivalence argument) with the sphere phase
// scanned through the view of light `Lp`: the line is walked from the light, direction -L, parameter s;
// forward distance t from the shadow origin so = P + N*bias is  t = (dist - bias*ndl) - s.
// (o, d, maxDist) are the reference's forward shadow ray; every confirmation uses them untouched.
__device__ __forceinline__ OccResult occluded_view(const DevScene& sc, const ulonglong2* __restrict__ view, const float4* __restrict__ s_tri,
                                                   d3 o, d3 d, double dist, d3 Lp, double maxDist, double bias, bool active,
                                                   bool no_filter, uint32_t& confirms) {
    OccState st = occ_begin(maxDist, bias, active);
    const int ns = sc.ns, ns_pad = sc.ns_pad;
    uint32_t first = (uint32_t)ns;
    if (ns > 0) {
        const float S = sqrtf((float)dot(Lp, Lp)) * 1.0000002f + sc.sph_extent;
        const ViewRay r = make_view_ray(-d, S);
        const float* vf = reinterpret_cast<const float*>(view);
        // occluders (and the `t <= bias` near cases) have s in (-bias, dist + bias): widen by 2*bias
        const float s_max = __double2float_ru(dist + 2.0 * bias), s_min = -__double2float_ru(2.0 * bias);
#pragma unroll 1
        for (int base = 0; base < ns_pad; base += kViewBlock) {
            uint32_t cand = view_block_mask(view + base, r.ux2, r.uy2, r.uz2);
            if (no_filter) cand = 0xFFFFFFFFu;
            if (st.done) cand = 0u;
            while (cand) {  // rare, divergent
                const int k = __clz(cand);
                cand &= ~(0x80000000u >> k);
                const int idx = base + k;
                if (idx >= ns) break;
                if (!no_filter) {
                    float s_lo, s_hi;
                    view_interval(vf, idx, r, s_lo, s_hi);
                    if (s_hi < s_min || s_lo > s_max) continue;  // behind the light / beyond the shaded point
                }
                double t;
                ++confirms;
                if (sphere_exact(sc.sph64[idx], o, d, t)) {
                    st.on_t(t, first, (uint32_t)idx);
                    if (st.done) break;
                }
            }
            if ((base & 63) == 32 && __all_sync(kFullMask, st.done)) break;
        }
    }
    if (active) st.res.sph_tests = first;
    if (sc.np > 0 || sc.nt > 0) {
        if (!__all_sync(kFullMask, st.done)) {
            const Ray32 r32 = make_ray32(sc, o, d, no_filter);
            occ_planes_triangles(sc, s_tri, r32, o, d, st, confirms);
        }
    }
    return st.res;
}

}  // namespace rte
