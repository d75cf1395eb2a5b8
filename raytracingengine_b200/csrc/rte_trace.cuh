// rte_trace.cuh — brute-force ray queries: FP32-FMA filter loops over shared-memory SoA arrays,
// FP64 confirmation of the candidates (rte_device.cuh), any-hit shadow queries with warp votes and
// the exact reference shadow march as the rare slow path.
//
// Reference functions replaced ("RE/" = /root/reference/RaytracingEngine/):
//   closest_hit()  <- Scene::IntersectClosest      RE/Scene.h:218-257
//   occluded()     <- Scene::computeTransmittance  RE/Scene.h:35-77 (opaque fast path, see below)
//   march_exact()  <- Scene::computeTransmittance  RE/Scene.h:35-77 (literal)
//
// Loop structure: the O(primitives) loops only run the 10-instruction FP32 filter and
// push the rare candidates on a small per-thread list; the FP64 confirmations run after the loop (or
// after a 64-primitive chunk for shadow rays) with the warp re-converged, in ascending canonical
// order, so ties resolve exactly like the reference's strict `<` scan.
#pragma once
#include "rte_device.cuh"

namespace rte {

constexpr unsigned kFullMask = 0xFFFFFFFFu;
constexpr int kSphUnroll = 8;   // spheres per filter iteration (ns_pad is a multiple of this)
constexpr int kTriUnroll = 2;   // triangles per filter iteration (nt_pad is a multiple of this)
constexpr int kMaxCand = 8;     // candidate list capacity (flushes in order when full)
constexpr int kSphChunk = 64;   // shadow rays: spheres between two warp votes
constexpr int kTriChunk = 16;   //              triangles between two warp votes

// Per-ray FP32 view + filter margins.  The filters must never reject a primitive
// the FP64 test accepts; they may accept extra ones (cost only).
struct Ray32 {
    float ox, oy, oz, dx, dy, dz;
    float m_sph;   // sphere:   q = (|f|^2 - r^2) - (f.d)^2 <= m_sph
    float m_b;     //           slack on f.d (behind / beyond pruning of candidates)
    float m_uv;    // triangle: slack on the unnormalised barycentrics
    float m_a;     //           slack on the determinant
    float a_min;   //           |a| >= a_min  (1e-6 of RE/Shape.h:211 scaled to the unit direction, minus m_a)
    float m_t;     //           slack on the unnormalised distance
};

__device__ __forceinline__ Ray32 make_ray32(const DevScene& sc, d3 o, d3 d, bool no_filter) {
    Ray32 r;
    // the filters work on the unit direction; the exact tests use the caller's (o, d) untouched
    const double len2 = dot(d, d);
    const double inv = rsqrt(len2);
    r.ox = (float)o.x; r.oy = (float)o.y; r.oz = (float)o.z;
    r.dx = (float)(d.x * inv); r.dy = (float)(d.y * inv); r.dz = (float)(d.z * inv);
    const float on = sqrtf(r.ox * r.ox + r.oy * r.oy + r.oz * r.oz) * 1.0000002f;
    const float Ss = on + sc.sph_extent;          // >= |C - O|, >= coordinate magnitudes
    r.m_sph = 48.0f * kEps32 * Ss * Ss;           // error bound 26*eps*S^2
    r.m_b = 16.0f * kEps32 * Ss;                  // error bound 7*eps*S
    const float St = on + sc.tri_extent, E = sc.tri_edge;
    r.m_uv = 40.0f * kEps32 * St * E;             // bound 20*eps*S*E, x2 safety
    r.m_a = 32.0f * kEps32 * E * E;               // bound 16*eps*E^2, x2
    r.m_t = 40.0f * kEps32 * St * E * E;          // bound 20*eps*S*E^2, x2
    r.a_min = 1e-6f * (float)inv * 0.999f - r.m_a;  // exact test: |a_exact| = |a_unit| * |d| >= 1e-6
    if (no_filter) {  // RTE_FLAG_NO_FILTER: accept every real primitive (padding entries stay rejected)
        r.m_sph = 3.0e38f; r.m_b = 3.0e38f; r.m_uv = 1.0e30f; r.m_a = 1.0e30f; r.m_t = 1.0e30f; r.a_min = -1.0e30f;
    }
    return r;
}

// q <= m  <=>  the ray's line may intersect the sphere.  10 FP32 instructions.
__device__ __forceinline__ float sph_q(const float4 s, const Ray32& r) {
    const float fx = s.x - r.ox, fy = s.y - r.oy, fz = s.z - r.oz;
    float g = __fmaf_rn(fx, fx, -s.w);
    g = __fmaf_rn(fy, fy, g);
    g = __fmaf_rn(fz, fz, g);
    float b = fx * r.dx;
    b = __fmaf_rn(fy, r.dy, b);
    b = __fmaf_rn(fz, r.dz, b);
    return __fmaf_rn(-b, b, g);
}

// Rare path: prune line-candidates that lie entirely behind the origin or entirely beyond `tmax`.
// With |d| = 1 the roots are t = b -+ sqrt(b^2 - g), g = |f|^2 - r^2 (their product is g, their sum 2b):
//   origin outside (g > 0) and b < 0            -> both roots negative -> the exact test misses;
//   origin outside and g > 2*b*tmax (b > 0)     -> near root g/(b+sqrt(..)) >= g/(2b) > tmax.
// tmax is in unit-direction distance; pass +inf for closest-hit queries.
__device__ __forceinline__ bool sph_prune(const float4 s, const Ray32& r, float tmax) {
    const float fx = s.x - r.ox, fy = s.y - r.oy, fz = s.z - r.oz;
    const float g = __fmaf_rn(fz, fz, __fmaf_rn(fy, fy, __fmaf_rn(fx, fx, -s.w)));
    const float b = __fmaf_rn(fz, r.dz, __fmaf_rn(fy, r.dy, fx * r.dx));
    if (!(g > r.m_sph)) return false;            // origin inside / on the sphere: keep
    if (b < -r.m_b) return true;                 // behind
    return (g - r.m_sph) > 2.0002f * (b + r.m_b) * tmax;  // beyond tmax
}

// Conservative Möller–Trumbore in FP32 on unnormalised quantities (no division).
__device__ __forceinline__ bool tri_filter(const float4 A, const float4 E1, const float4 E2, const Ray32& r) {
    const float hx = __fmaf_rn(r.dy, E2.z, -(r.dz * E2.y));
    const float hy = __fmaf_rn(r.dz, E2.x, -(r.dx * E2.z));
    const float hz = __fmaf_rn(r.dx, E2.y, -(r.dy * E2.x));
    const float a = __fmaf_rn(E1.z, hz, __fmaf_rn(E1.y, hy, E1.x * hx));
    const float sx = r.ox - A.x, sy = r.oy - A.y, sz = r.oz - A.z;
    const float U = __fmaf_rn(sz, hz, __fmaf_rn(sy, hy, sx * hx));
    const float qx = __fmaf_rn(sy, E1.z, -(sz * E1.y));
    const float qy = __fmaf_rn(sz, E1.x, -(sx * E1.z));
    const float qz = __fmaf_rn(sx, E1.y, -(sy * E1.x));
    const float V = __fmaf_rn(r.dz, qz, __fmaf_rn(r.dy, qy, r.dx * qx));
    const float T = __fmaf_rn(E2.z, qz, __fmaf_rn(E2.y, qy, E2.x * qx));
    const float sg = copysignf(1.0f, a), aa = fabsf(a);
    const float Us = U * sg, Vs = V * sg;
    return (aa >= r.a_min) & (Us >= -r.m_uv) & (Vs >= -r.m_uv) &
           (Us + Vs <= aa + 2.0f * r.m_uv + r.m_a) & (T * sg >= -r.m_t);
}

struct Hit {
    double t;
    uint32_t kind;  // RTE_HIT_*
    uint32_t idx;   // canonical index inside its class (triangles: flattened index)
};

struct CandList {
    int n;
    int idx[kMaxCand];
};

// Filter spheres [i0, i1) (multiples of kSphUnroll) and push candidates; `flush` empties a full list.
template <class Flush>
__device__ __forceinline__ void scan_spheres(const float4* __restrict__ s_sph, int i0, int i1, const Ray32& r, float m,
                                             float tmax, CandList& L, Flush&& flush) {
#pragma unroll 1
    for (int i = i0; i < i1; i += kSphUnroll) {
        float q[kSphUnroll];
#pragma unroll
        for (int k = 0; k < kSphUnroll; ++k) q[k] = sph_q(s_sph[i + k], r);
        float mn = q[0];
#pragma unroll
        for (int k = 1; k < kSphUnroll; ++k) mn = fminf(mn, q[k]);
        if (mn <= m) {  // rare
#pragma unroll 1
            for (int k = i; k < i + kSphUnroll; ++k) {
                const float4 s = s_sph[k];
                if (sph_q(s, r) <= m && !sph_prune(s, r, tmax)) {
                    if (L.n == kMaxCand) flush();
                    L.idx[L.n++] = k;
                }
            }
        }
    }
}

template <class Flush>
__device__ __forceinline__ void scan_triangles(const float4* __restrict__ s_tri, int i0, int i1, const Ray32& r, bool on,
                                               CandList& L, Flush&& flush) {
#pragma unroll 1
    for (int i = i0; i < i1; i += kTriUnroll) {
        const float4* p = s_tri + 3 * i;
        const bool c0 = tri_filter(p[0], p[1], p[2], r), c1 = tri_filter(p[3], p[4], p[5], r);
        if ((c0 | c1) & on) {
            if (c0) { if (L.n == kMaxCand) flush(); L.idx[L.n++] = i; }
            if (c1) { if (L.n == kMaxCand) flush(); L.idx[L.n++] = i + 1; }
        }
    }
}

// Planes (exact) then triangles (FP32 filter + FP64 confirm) of a closest-hit query, after the spheres.
__device__ __forceinline__ void closest_planes_triangles(const DevScene& sc, const float4* __restrict__ s_tri, const Ray32& r, d3 o, d3 d,
                                                         bool alive, Hit& best, bool& found, uint32_t& confirms, bool tris = true) {
    if (alive) {
        const int np = sc.np;
#pragma unroll 1
        for (int i = 0; i < np; ++i) {
            double t;
            if (plane_exact(sc.planes[i], o, d, t) && (!found || t < best.t)) {
                found = true; best.t = t; best.kind = RTE_HIT_PLANE; best.idx = (uint32_t)i;
            }
        }
    }
    if (sc.nt == 0 || !tris) return;  // !tris: the caller scans the triangles through their view (rte_views.cuh)
    CandList L;
    L.n = 0;
    auto flush = [&]() {
#pragma unroll 1
        for (int j = 0; j < L.n; ++j) {
            const int k = L.idx[j];
            double t;
            ++confirms;
            if (tri_exact(sc.tri64[k], o, d, t) && (!found || t < best.t)) {
                found = true; best.t = t; best.kind = RTE_HIT_TRIANGLE; best.idx = (uint32_t)k;
            }
        }
        L.n = 0;
    };
    scan_triangles(s_tri, 0, sc.nt_pad, r, alive, L, flush);
    flush();
}

// Scene::IntersectClosest.  No warp-synchronous operation inside: callable from divergent code.
__device__ __forceinline__ void closest_hit_body(const DevScene& sc, const float4* __restrict__ s_sph,
                                                 const float4* __restrict__ s_tri, d3 o, d3 d, bool alive, bool no_filter,
                                                 Hit& best, uint32_t& confirms) {
    const Ray32 r = make_ray32(sc, o, d, no_filter);
    best.t = 0.0; best.kind = RTE_HIT_NONE; best.idx = 0;
    bool found = false;
    if (sc.ns > 0) {
        CandList L;
        L.n = 0;
        auto flush = [&]() {
#pragma unroll 1
            for (int j = 0; j < L.n; ++j) {
                const int k = L.idx[j];
                double t;
                ++confirms;
                if (sphere_exact(sc.sph64[k], o, d, t) && (!found || t < best.t)) {  // strict '<', RE/Scene.h:223
                    found = true; best.t = t; best.kind = RTE_HIT_SPHERE; best.idx = (uint32_t)k;
                }
            }
            L.n = 0;
        };
        scan_spheres(s_sph, 0, sc.ns_pad, r, alive ? r.m_sph : -INFINITY, INFINITY, L, flush);
        flush();
    }
    closest_planes_triangles(sc, s_tri, r, o, d, alive, best, found, confirms);
}

static __device__ __noinline__ void closest_hit_cold(const DevScene& sc, const float4* s_sph, const float4* s_tri, d3 o,
                                              d3 d, bool no_filter, Hit& best, uint32_t& confirms) {
    closest_hit_body(sc, s_sph, s_tri, o, d, true, no_filter, best, confirms);
}

__device__ __forceinline__ int material_of(const DevScene& sc, const Hit& h) {
    int m = h.kind == RTE_HIT_SPHERE ? sc.sph_mat[h.idx]
          : h.kind == RTE_HIT_PLANE  ? sc.planes[h.idx].mat
                                     : sc.tri_info[h.idx].x;
    return (m >= 0 && m < sc.nm) ? m : sc.nm;  // index nm = default Material (RE/Shape.h:13-19)
}

// Scene::computeTransmittance, RE/Scene.h:35-77, literally (transparent occluders included).
static __device__ __noinline__ double march_exact(const DevScene& sc, const float4* s_sph, const float4* s_tri, d3 o, d3 d,
                                           double maxDist, double bias, bool no_filter, uint32_t& confirms) {
    double T = 1.0, traveled = 0.0;
    int safety = 64;
    while (safety-- > 0 && T > 1e-4 && traveled < maxDist) {
        Hit h;
        closest_hit_cold(sc, s_sph, s_tri, o, d, no_filter, h, confirms);
        if (h.kind == RTE_HIT_NONE) break;
        const double t = h.t;
        if (t <= 0.0) { o = o + d * bias; traveled += bias; continue; }
        if (t <= bias) { o = (o + d * t) + d * bias; traveled += t + bias; continue; }
        if (traveled + t >= maxDist) break;
        const double tr = std_clamp(sc.materials[material_of(sc, h)].transparency, 0.0, 1.0);
        T *= tr;
        const d3 newOrigin = (o + d * t) + d * bias;
        traveled += t + bias;
        o = newOrigin;
    }
    return std_clamp(T, 0.0, 1.0);
}

struct OccResult {
    bool occluded;  // some primitive reports t in (bias, maxDist)
    bool near;      // the any-hit answer is not decisive: take the literal march (see occluded())
    uint32_t sph_tests, pl_tests, tri_tests;  // canonical-order tests up to and incl. the first occluder
};

// State of one any-hit shadow query while it walks the primitive classes in canonical order.
struct OccState {
    OccResult res;
    bool done;       // decided (occluded) or inactive lane
    double bias, maxDist, t_safe;
    __device__ __forceinline__ void on_t(double t, uint32_t& firstv, uint32_t k) {
        if (t <= bias) res.near = true;
        else if (t < maxDist) {
            res.occluded = true; done = true; firstv = k + 1u;
            if (t <= t_safe) res.near = true;
        }
    }
    // View scans: a ray that turned `near` keeps scanning (never `done`), so that its candidate list is complete and
    // the exact march can run over that list instead of over every primitive (march_list, rte_views.cuh).
    __device__ __forceinline__ void on_t_view(double t, uint32_t& firstv, uint32_t k) {
        if (t <= bias) res.near = true;
        else if (t < maxDist) {
            if (!res.occluded) { res.occluded = true; firstv = k + 1u; }
            if (t <= t_safe) res.near = true;
            if (!res.near) done = true;
        }
    }
};

__device__ __forceinline__ OccState occ_begin(double maxDist, double bias, bool active) {
    OccState st;
    st.res.occluded = false; st.res.near = false;
    st.res.sph_tests = st.res.pl_tests = st.res.tri_tests = 0;
    st.done = !active;
    st.bias = bias; st.maxDist = maxDist; st.t_safe = 130.0 * bias;
    return st;
}

// Planes (exact) and triangles (FP32 filter + FP64 confirm) of an any-hit query; all 32 lanes call it.
__device__ __forceinline__ void occ_planes_triangles(const DevScene& sc, const float4* __restrict__ s_tri, const Ray32& r, d3 o, d3 d,
                                                     OccState& st, uint32_t& confirms, bool tris = true) {
    if (__all_sync(kFullMask, st.done)) return;
    const int np = sc.np;
    if (!st.done) {
        uint32_t firstp = (uint32_t)np;
#pragma unroll 1
        for (int i = 0; i < np; ++i) {
            double t;
            if (plane_exact(sc.planes[i], o, d, t)) {
                st.on_t(t, firstp, (uint32_t)i);
                if (st.done) break;
            }
        }
        st.res.pl_tests = firstp;
    }
    const int nt_pad = sc.nt_pad, nt = sc.nt;
    if (nt == 0 || !tris) return;
    const bool scan_tris = !st.done;
    uint32_t firstt = (uint32_t)nt;
    CandList L;
    L.n = 0;
    auto flush = [&]() {
#pragma unroll 1
        for (int j = 0; j < L.n; ++j) {
            const int k = L.idx[j];
            double t;
            if (st.done) break;
            ++confirms;
            if (tri_exact(sc.tri64[k], o, d, t)) st.on_t(t, firstt, (uint32_t)k);
        }
        L.n = 0;
    };
#pragma unroll 1
    for (int base = 0; base < nt_pad; base += kTriChunk) {
        scan_triangles(s_tri, base, min(base + kTriChunk, nt_pad), r, !st.done, L, flush);
        flush();
        if (__all_sync(kFullMask, st.done)) break;
    }
    if (scan_tris) st.res.tri_tests = firstt;
}

// Any-hit shadow query for opaque scenes.  Must be called by all 32 lanes of the warp (votes inside).
//
// Equivalence with the reference march (RE/Scene.h:35-77) for opaque scenes: if no primitive reports
// t <= bias from the shadow origin, the march's first closest hit has t > bias, and T = 0 iff that closest
// t < maxDist, i.e. iff ANY primitive reports t in (bias, maxDist).  If some primitive does report
// t <= bias the march steps past it (at most 64 steps of at most 2*bias): the any-hit answer stays right
// as long as the occluder it found is farther than those steps can reach, t > t_safe = 130*bias.
// So: lanes that see a t <= bias, or whose first occluder has t <= t_safe, are flagged `near` and
// re-done by march_exact(); all others are decided here.
__device__ __forceinline__ OccResult occluded(const DevScene& sc, const float4* __restrict__ s_sph,
                                              const float4* __restrict__ s_tri, d3 o, d3 d, double maxDist,
                                              double bias, bool active, bool no_filter, uint32_t& confirms) {
    const Ray32 r = make_ray32(sc, o, d, no_filter);
    OccState st = occ_begin(maxDist, bias, active);
    // unit-direction distance bound for pruning (|d| may differ from 1 for user queries)
    const float tmax32 = (float)(maxDist * sqrt(dot(d, d))) * 1.00001f + r.m_b;
    const int ns_pad = sc.ns_pad, ns = sc.ns;
    uint32_t first = (uint32_t)ns;
    CandList L;
    L.n = 0;
    {
        auto flush = [&]() {
#pragma unroll 1
            for (int j = 0; j < L.n; ++j) {
                const int k = L.idx[j];
                double t;
                if (st.done) break;
                ++confirms;
                if (sphere_exact(sc.sph64[k], o, d, t)) st.on_t(t, first, (uint32_t)k);
            }
            L.n = 0;
        };
#pragma unroll 1
        for (int base = 0; base < ns_pad; base += kSphChunk) {
            scan_spheres(s_sph, base, min(base + kSphChunk, ns_pad), r, st.done ? -INFINITY : r.m_sph, tmax32, L, flush);
            flush();
            if (__all_sync(kFullMask, st.done)) break;
        }
    }
    if (active) st.res.sph_tests = first;
    occ_planes_triangles(sc, s_tri, r, o, d, st, confirms);
    return st.res;
}

}  // namespace rte
