// rte_render32r.cu — the FP32 production path for REFLECTIVE scenes with few primitives (sm_100a): the reference's default
// scene (main(): five slightly specular planes, a specular box mesh, two lights — config C1) and its relatives.
//
// Replaces Scene::TraceRay's recursion with reflection (RE/Scene.h:131-198; no transparent material, so every node has at
// most one child and the recursion is a chain), directLightning with the Blinn-Phong term (RE/Scene.h:79-129) and
// tonemap()/toColor for those scenes, to the bar BASELINE.json states: nearest-hit index and shadow visibility of the primary
// hit bit-exact, RGB8 within +-1 LSB.  ("RE/" = /root/reference/RaytracingEngine/.)
//
//  * Values are FP32 (hit points in a handful of cheap FP64 operations), decisions come from FP32 INTERVALS (rte_interval.cuh)
//    over ALL primitives — a reflected ray has no common origin with its neighbours, so the per-origin direction grids of
//    rte_render32.cu do not apply; the path takes scenes of at most kRrPrims spheres + triangles, everything in shared memory.
//  * The CHAIN of rays — origin, direction, hit distance of the chosen primitive, hit point, normal, reflected direction, and
//    per light the vector, distance and N.L of the shadow-ray guards — is computed with the REFERENCE'S OWN FP64 operations
//    (rte_device.cuh: the functions of the exact kernels), ~900 instructions per node against ~3500 for the interval tests: the
//    ray at depth k is the reference's ray bit for bit.  (A reflected ray in FP32 is not the reference's ray any more: its
//    deviation triples with every bounce and after two bounces it is of the order of the shadow bias — measured, every pixel of
//    C1 was inconclusive.)  The interval tests see that ray rounded once, like a primary ray.
//  * Because the chain is the reference's, an inconclusive decision can be settled on the spot by the reference's own test:
//    overlapping nearest-hit intervals by its IntersectClosest over all primitives (closest_exact_brute), an unsure shadow ray
//    by its literal transmittance march (march_exact_brute).  C1 needs that on whole pixel rows: its camera sits on the room's
//    axes and looks exactly along the lines where two planes meet, at every depth.  (The tile-flag mechanism — a tile with an
//    undecidable pixel is rendered again by the exact kernel, rte_render.cu — remains as the safety net; no scene has used it.)
#include <algorithm>

#include "rte_grid.cuh"
#include "rte_interval.cuh"
#include "rte_kernels.h"
#include "rte_tonemap.cuh"

namespace rte {

constexpr int kRrThreads = 256;
constexpr int kRrPrims = 32, kRrPlanes = 8, kRrLights = 16;  // every ray tests every primitive: measured against the exact kernel's filters, 12 primitives win (0.34 vs 0.40 ms), 200 lose (2.8 vs 0.64 ms)
// A chain ends once its weight falls below this: what the reference still adds below it (down to maxRecursion = 10) is less than
// 1e-7 of a radiance per node, under the resolution of the FP32 pixel.  (C1: the planes reflect 1 %: four or five nodes instead of
// ten.)  The ray counters therefore count the rays traced, fewer than the reference's for such scenes.
constexpr float kRrCutWeight = 1e-7f;
constexpr unsigned kFullMask = 0xFFFFFFFFu;

struct Sph32r { float hx, hy, hz, r, lx, ly, lz; int mat; };
struct Light32r { float hi[3], lo[3], E[3]; };
struct SceneRr {
    Sph32r sph[kRrPrims];
    Tri32 tri[kRrPrims];
    Plane32 pl[kRrPlanes];
    Light32r li[kRrLights];
    float4 tb[kRrPrims];  // bounding sphere of every triangle {centroid, farthest vertex (inflated)}: far_miss() in front of tri_interval()
    float4 tb_all;        // ... and of all of them: most rays of a room skip the mesh with one test
    double cam[3];
    int ns, nt, np, nl;
};

#ifdef RTE_RR_DEBUG
#define RR_WHY(code) do { if (out.why == 0u) out.why = (uint32_t)(code) + 16u * (uint32_t)depth_dbg; } while (0)
#else
#define RR_WHY(code) do { } while (0)
#endif
struct PxR {
    uint32_t why;
    float r, g, b;
    uint32_t prim, lmask;
    uint32_t c_shadow, c_secondary, c_sph, c_pl, c_tri;
};

__device__ __forceinline__ void split64(double v, float& hi, float& lo) { hi = (float)v; lo = (float)(v - (double)hi); }

// ------------------------------------------------------------------------------ the reference's own tests, for the rare cases
// Scene::IntersectClosest, RE/Scene.h:218-257, over every primitive with the exact FP64 tests: canonical order (spheres, planes,
// triangles — standalone ones, then every model's), strict '<'.  Runs when the FP32 intervals of two candidates overlap (a room
// corner seen exactly along its edge: C1's camera sits on the room's axes, whole pixel rows look along such lines).
struct HitX { int kind; int idx; double t; };
static __device__ __noinline__ HitX closest_exact_brute(const DevScene& sc, d3 o, d3 d) {
    HitX best;
    best.kind = RTE_HIT_NONE; best.idx = 0; best.t = 0.0;
    for (int i = 0; i < sc.ns; ++i) {
        double t;
        if (sphere_exact(sc.sph64[i], o, d, t) && (best.kind == RTE_HIT_NONE || t < best.t)) { best.kind = RTE_HIT_SPHERE; best.idx = i; best.t = t; }
    }
    for (int i = 0; i < sc.np; ++i) {
        double t;
        if (plane_exact(sc.planes[i], o, d, t) && (best.kind == RTE_HIT_NONE || t < best.t)) { best.kind = RTE_HIT_PLANE; best.idx = i; best.t = t; }
    }
    for (int i = 0; i < sc.nt; ++i) {
        double t;
        if (tri_exact(sc.tri64[i], o, d, t) && (best.kind == RTE_HIT_NONE || t < best.t)) { best.kind = RTE_HIT_TRIANGLE; best.idx = i; best.t = t; }
    }
    return best;
}
// Scene::computeTransmittance, RE/Scene.h:35-77, literally.
static __device__ __noinline__ double march_exact_brute(const DevScene& sc, d3 o, d3 d, double maxDist, double bias) {
    double T = 1.0, traveled = 0.0;
    int safety = 64;
    while (safety-- > 0 && T > 1e-4 && traveled < maxDist) {
        const HitX h = closest_exact_brute(sc, o, d);
        if (h.kind == RTE_HIT_NONE) break;
        const double t = h.t;
        if (t <= 0.0) { o = o + d * bias; traveled += bias; continue; }
        if (t <= bias) { o = (o + d * t) + d * bias; traveled += t + bias; continue; }
        if (traveled + t >= maxDist) break;
        int m = h.kind == RTE_HIT_SPHERE ? sc.sph_mat[h.idx] : h.kind == RTE_HIT_PLANE ? sc.planes[h.idx].mat : sc.tri_info[h.idx].x;
        m = (m >= 0 && m < sc.nm) ? m : sc.nm;
        T *= std_clamp(sc.materials[m].transparency, 0.0, 1.0);
        o = (o + d * t) + d * bias;
        traveled += t + bias;
    }
    return std_clamp(T, 0.0, 1.0);
}

// One sample of one pixel per lane.  Returns false when something could not be decided (never seen; the tile would be rendered
// again by the exact kernel and nothing of `out` used).
// kTree: scenes with transparent materials — TraceRay is a tree (refracted + reflected child, RE/Scene.h:175-195), walked depth
// first with a small per-pixel stack of pending rays {origin, direction, weight, depth}; without them it is a chain and no stack
// exists.  pixel = sum over the nodes of  weight * (1 - transparency) * localLight  (+ weight * sky where a branch ends).
struct NodeR { d3 o, d; float w; int depth; };
template <bool kTree>
__device__ __forceinline__ bool trace32r(const RenderParams& P, const SceneRr& S, int px, int py, bool alive, int sample, PxR& out) {
    const DevScene& sc = P.sc;
    const int ns = S.ns, nt = S.nt, np = S.np, nl = S.nl;
    const double bias64 = P.bias;
    const float bias = (float)P.bias;
    const int max_depth = P.max_depth;
    out.r = out.g = out.b = 0.0f; out.why = 0u;
    int depth_dbg = 0; (void)depth_dbg;
    out.prim = kMiss; out.lmask = 0u; out.c_shadow = out.c_secondary = out.c_sph = out.c_pl = out.c_tri = 0u;
    bool ok = true;
    // ---- Camera::getRay, RE/Math.h:99-121, with the reference's own operations: the chain of rays below is the reference's, bit for bit
    d3 o = ld3(P.cam.pos), d = camera_dir(P.cam, px, py, sample);
    const float eo = 0.0f, ed = 4.0f * kEps32;  // what the interval tests see of a ray: its origin exactly (hi + lo), its direction rounded once
    float w = 1.0f;                             // weight of the node's light in the pixel
    float ar = 0.0f, ag = 0.0f, ab = 0.0f;
    bool live = alive;
    int depth = 0;
    NodeR stk[kTree ? RTE_MAX_DEPTH + 2 : 1];
    int sp = 0;  // pending rays of this pixel
#pragma unroll 1
    while (true) {
        if (kTree && !live && sp > 0) { --sp; o = stk[sp].o; d = stk[sp].d; w = stk[sp].w; depth = stk[sp].depth; live = true; }
        if (!__any_sync(kFullMask, live)) break;
        depth_dbg = depth;
        const float Dx = (float)d.x, Dy = (float)d.y, Dz = (float)d.z;
        auto sky = [&]() {  // Scene::backgroundColor, RE/Scene.h:30-33
            const float t = 0.5f * (Dy + 1.0f);
            ar = __fmaf_rn(w, __fmaf_rn(0.5f, t, 1.0f - t), ar); ag = __fmaf_rn(w, __fmaf_rn(0.7f, t, 1.0f - t), ag); ab = __fmaf_rn(w, 1.0f, ab);
        };
        if (live && depth >= max_depth) { sky(); live = false; }  // TraceRay(recursionAmount >= maxRecursion), RE/Scene.h:132-134
        if (!live) continue;
        float ohx, ohy, ohz, olx, oly, olz;
        split64(o.x, ohx, olx); split64(o.y, ohy, oly); split64(o.z, ohz, olz);
        const float o1 = fabsf(ohx) + fabsf(ohy) + fabsf(ohz);
        // ---- Scene::IntersectClosest, RE/Scene.h:218-257, over every primitive, from intervals
        Near32 nr;
        nr.kind = RTE_HIT_NONE; nr.idx = 0; nr.lo = 0.0f; nr.hi = 0.0f; nr.amb_lo = INFINITY;
#pragma unroll 1
        for (int i = 0; i < ns; ++i) {
            const Sph32r& s = S.sph[i];
            if (far_miss(make_float4(s.hx, s.hy, s.hz, s.r), ohx, ohy, ohz, o1, Dx, Dy, Dz)) continue;
            const Cand32 c = classify_sphere(make_float4(s.hx, s.hy, s.hz, s.r), make_float4(s.lx, s.ly, s.lz, 0.0f), ohx, ohy, ohz, olx, oly, olz, Dx, Dy, Dz, eo, ed);
            float lo, hi;
            const int cls = reported_t(c, lo, hi);
            if (cls == kCandHit) near_hit(nr, RTE_HIT_SPHERE, i, lo, hi);
            else if (cls == kCandUnsure) near_unsure(nr, lo);
        }
#pragma unroll 1
        for (int i = 0; i < np; ++i) {
            float lo, hi;
            const int cls = plane_interval(S.pl[i], ohx, ohy, ohz, o1, Dx, Dy, Dz, lo, hi, eo, ed);
            if (cls == kCandHit) near_hit(nr, RTE_HIT_PLANE, i, lo, hi);
            else if (cls == kCandUnsure) near_unsure(nr, 0.0f);
        }
        if (nt > 0 && !far_miss(S.tb_all, ohx, ohy, ohz, o1, Dx, Dy, Dz)) {
#pragma unroll 1
            for (int i = 0; i < nt; ++i) {
                if (far_miss(S.tb[i], ohx, ohy, ohz, o1, Dx, Dy, Dz)) continue;  // the line misses the triangle's bounding sphere
                float lo, hi;
                const int cls = tri_interval(S.tri[i], ohx, ohy, ohz, olx, oly, olz, Dx, Dy, Dz, eo, ed, lo, hi);
                if (cls == kCandHit) near_hit(nr, RTE_HIT_TRIANGLE, i, lo, hi);
                else if (cls == kCandUnsure) near_unsure(nr, lo);
            }
        }
        out.c_sph += (uint32_t)ns; out.c_pl += (uint32_t)np; out.c_tri += (uint32_t)nt;
        if (depth > 0) out.c_secondary += 1u;
        const int node_depth = depth;
        // ---- the hit: which primitive from the intervals (or, when they overlap, from the reference's own scan), its distance
        // from the reference's own test of that primitive
        HitX hx;
        if (near_ambiguous(nr)) {
            hx = closest_exact_brute(sc, o, d);
        } else {
            hx.kind = nr.kind; hx.idx = nr.idx; hx.t = 0.0;
            bool hit = true;
            if (nr.kind == RTE_HIT_SPHERE) hit = sphere_exact(sc.sph64[nr.idx], o, d, hx.t);
            else if (nr.kind == RTE_HIT_PLANE) hit = plane_exact(sc.planes[nr.idx], o, d, hx.t);
            else if (nr.kind == RTE_HIT_TRIANGLE) hit = tri_exact(sc.tri64[nr.idx], o, d, hx.t);
            if (!hit) { RR_WHY(2); ok = false; live = false; continue; }  // (a certain FP32 hit the exact test denies: a bound would be wrong)
        }
        if (hx.kind == RTE_HIT_NONE) { sky(); live = false; continue; }
        // ---- the hit record and TraceRay's frame with the reference's operations (GetHitInfoAt, RE/Shape.h:104-121,165-180,229-243;
        // RE/Scene.h:144-148), as in trace_pixels (rte_render.cu)
        const d3 Pn = o + d * hx.t;
        d3 ngeo;
        int mat, self = -1;
        uint32_t report;
        if (hx.kind == RTE_HIT_SPHERE) {
            const double4 s = sc.sph64[hx.idx];
            ngeo = normalize(Pn - mk(s.x, s.y, s.z));
            mat = S.sph[hx.idx].mat; report = (uint32_t)hx.idx;
        } else if (hx.kind == RTE_HIT_PLANE) {
            ngeo = ld3(sc.planes[hx.idx].n);
            mat = S.pl[hx.idx].mat; report = (uint32_t)hx.idx;
        } else {
            ngeo = ld3(sc.tri64[hx.idx].n);
            mat = S.tri[hx.idx].mat; report = S.tri[hx.idx].report;
        }
        const d3 incoming = normalize(d);
        const bool front = dot(ngeo, incoming) < 0.0;
        const d3 N = front ? ngeo : -ngeo;
        const d3 normal = normalize(N);  // directLightning re-normalises, RE/Scene.h:81
        const float Nx = (float)normal.x, Ny = (float)normal.y, Nz = (float)normal.z;
        if (hx.kind == RTE_HIT_SPHERE && front && bias >= 1e-5f) self = hx.idx;
        if (node_depth == 0) out.prim = ((uint32_t)hx.kind << 30) | report;
        mat = (mat >= 0 && mat < sc.nm) ? mat : sc.nm;
        const rte_material& M = sc.materials[mat];
        const float m_spec = (float)M.specular, m_shin = (float)M.shininess;
        const bool spec_on = M.transparency <= 0.0 && M.specular > 0.0;  // RE/Scene.h:116
        // ---- Scene::directLightning, RE/Scene.h:79-129: the guards and the shadow ray in the reference's FP64, the light in FP32
        const d3 so = Pn + normal * bias64;  // RE/Scene.h:104
        float shx, shy, shz, slx, sly, slz;
        split64(so.x, shx, slx); split64(so.y, shy, sly); split64(so.z, shz, slz);
        const float so1 = fabsf(shx) + fabsf(shy) + fabsf(shz);
        float dr = 0.0f, dg = 0.0f, db = 0.0f, sr = 0.0f, sg = 0.0f, sb = 0.0f;
#pragma unroll 1
        for (int j = 0; j < nl; ++j) {
            const d3 vecToLight = ld3(sc.lights[j].pos) - Pn;
            const d4 ul = unit_len(vecToLight);  // {lightToHit, distanceToLight}
            const double dist64 = ul.w;
            if (!(dist64 > 0.0)) continue;       // RE/Scene.h:89
            const d3 L64 = mk(ul.x, ul.y, ul.z);
            const double ndl64 = std_max(0.0, dot(normal, L64));
            if (!(ndl64 > 0.0) || !(dist64 > bias64)) continue;  // RE/Scene.h:92-101
            const float Lx = (float)L64.x, Ly = (float)L64.y, Lz = (float)L64.z, dist = (float)dist64, ndl = (float)ndl64;
            const float maxDist = (float)(dist64 - bias64);
            const float dt = 24.0f * kEps32 * (dist + so1) + 1e-7f;  // slack of the range ends (maxDist, bias) against the candidate bounds
            const float tmax_lo = maxDist - dt, tmax_hi = maxDist + dt, bias_hi = bias + dt;
            // computeTransmittance (RE/Scene.h:35-77) for opaque scenes: some primitive reports t in (bias, maxDist) — unless one
            // reports t <= bias (the march steps past it): certain occluder, certain miss, or unsure.  Canonical order.
            bool occl = false, unsure = false;
            int first_s = ns, first_p = np, first_t = nt;
#pragma unroll 1
            for (int i = 0; i < ns && !occl && !unsure; ++i) {
                if (i == self) continue;  // so lies outside the sphere the point is on, the ray leaves it: both roots < 0
                const Sph32r& s = S.sph[i];
                if (far_miss(make_float4(s.hx, s.hy, s.hz, s.r), shx, shy, shz, so1, Lx, Ly, Lz)) continue;
                const Cand32 c = classify_sphere(make_float4(s.hx, s.hy, s.hz, s.r), make_float4(s.lx, s.ly, s.lz, 0.0f), shx, shy, shz, slx, sly, slz, Lx, Ly, Lz, eo, ed);
                float lo, hi;
                const int cls = reported_t(c, lo, hi);
                if (cls == kCandMiss || lo > tmax_hi) continue;
                if (cls == kCandHit && lo > bias_hi && hi < tmax_lo) { occl = true; first_s = i; }
                else unsure = true;
            }
#pragma unroll 1
            for (int i = 0; i < np && !occl && !unsure; ++i) {
                float lo, hi;
                const int cls = plane_interval(S.pl[i], shx, shy, shz, so1, Lx, Ly, Lz, lo, hi, eo, ed);
                if (cls == kCandMiss) continue;
                if (cls == kCandHit && lo > tmax_hi) continue;
                if (cls == kCandHit && lo > bias_hi && hi < tmax_lo) { occl = true; first_p = i; }
                else unsure = true;
            }
            const bool mesh = nt > 0 && !far_miss(S.tb_all, shx, shy, shz, so1, Lx, Ly, Lz);
#pragma unroll 1
            for (int i = 0; mesh && i < nt && !occl && !unsure; ++i) {
                if (far_miss(S.tb[i], shx, shy, shz, so1, Lx, Ly, Lz)) continue;
                float lo, hi;
                const int cls = tri_interval(S.tri[i], shx, shy, shz, slx, sly, slz, Lx, Ly, Lz, eo, ed, lo, hi);
                if (cls == kCandMiss || lo > tmax_hi) continue;
                if (cls == kCandHit && lo > bias_hi && hi < tmax_lo) { occl = true; first_t = i; }
                else unsure = true;
            }
            float Tf = 1.0f;  // transmittance of an unoccluded ray
            if (unsure || (kTree && occl)) {
                // the literal march with the reference's tests decides — and, with transparent materials, every occluded ray takes
                // it: T is the product of the occluders' transparencies (RE/Scene.h:68-69)
                const double T = march_exact_brute(sc, so, L64, dist64 - bias64, bias64);
                occl = !(T > bias64);
                Tf = (float)T;
                if (unsure) { first_s = ns; first_p = np; first_t = nt; }
            }
            out.c_shadow += 1u;
            out.c_sph += first_s < ns ? (uint32_t)first_s + 1u : (uint32_t)ns;
            if (first_s >= ns) {
                out.c_pl += first_p < np ? (uint32_t)first_p + 1u : (uint32_t)np;
                if (first_p >= np) out.c_tri += first_t < nt ? (uint32_t)first_t + 1u : (uint32_t)nt;
            }
            if (node_depth == 0 && j < 16) out.lmask |= 1u << (16 + j);
            if (occl) continue;
            if (node_depth == 0 && j < 16) out.lmask |= 1u << j;
            const float inv = __frcp_rn(dist);
            const float att = inv * inv * Tf;  // transmittance / d^2
            const float wd = att * ndl;
            const float Er = (float)sc.lights[j].E[0], Eg = (float)sc.lights[j].E[1], Eb = (float)sc.lights[j].E[2];
            dr = __fmaf_rn(Er, wd, dr); dg = __fmaf_rn(Eg, wd, dg); db = __fmaf_rn(Eb, wd, db);
            if (spec_on) {  // Blinn-Phong, RE/Scene.h:116-124: halfVector = normalize(lightToHit + viewDir), viewDir = -incoming
                const float hx_ = Lx - Dx, hy_ = Ly - Dy, hz_ = Lz - Dz;
                const float h2 = __fmaf_rn(hz_, hz_, __fmaf_rn(hy_, hy_, hx_ * hx_));
                if (h2 > 1e-24f) {  // (normalize() returns 0 below 1e-12: NdotH = 0, no term)
                    const float ndh = __fmaf_rn(Nz, hz_, __fmaf_rn(Ny, hy_, Nx * hx_)) * rsqrtf(h2);
                    if (ndh > 0.0f) {
                        const float sf = tm32::ex2(m_shin * tm32::lg2(fminf(ndh, 1.0f))) * att;
                        sr = __fmaf_rn(Er, sf, sr); sg = __fmaf_rn(Eg, sf, sg); sb = __fmaf_rn(Eb, sf, sb);
                    }
                }
            }
        }
        // localLight = color * diffuse + specular * material.specular; finalLight += localLight * (1 - transparency), RE/Scene.h:126-128,171-173
        const double transparency = kTree ? std_clamp(M.transparency, 0.0, 1.0) : 0.0;
        const float wl = kTree ? w * (float)(1.0 - transparency) : w;
        if (!kTree || transparency < 1.0) {
            ar = __fmaf_rn(wl, __fmaf_rn((float)M.color[0], dr, sr * m_spec), ar);
            ag = __fmaf_rn(wl, __fmaf_rn((float)M.color[1], dg, sg * m_spec), ag);
            ab = __fmaf_rn(wl, __fmaf_rn((float)M.color[2], db, sb * m_spec), ab);
        }
        // ---- the children, RE/Scene.h:161-195: the refracted ray of a transparent material with weight transparency * (1 - fresnel),
        // the reflected ray with weight fresnel (transparent) or material.specular, when that exceeds bias — rays and decisions in
        // the reference's FP64 (Vec3::refract, Scene::fresnel with pow() like the exact kernel)
        bool has_t = false;
        d3 to = o, td = d;
        float wt = 0.0f;
        double reflectiveness = M.specular;
        if (kTree && transparency > 0.0) {
            const double cosTheta = std_max(0.0, dot(N, -incoming));
            const double etaT = M.refractive_index;
            const double f0 = pow((etaT - 1.0) / (etaT + 1.0), 2.0);
            double fresnelAmount = f0 + (1.0 - f0) * pow(1.0 - cosTheta, 5.0);  // Scene::fresnel, RE/Scene.h:26-28
            const double eta = front ? (1.0 / etaT) : (etaT / 1.0);
            d3 rd = refract(incoming, N, eta);
            if (length(rd) > bias64) {
                rd = normalize(rd);
                to = Pn + rd * (bias64 * 1e2); td = rd;
                wt = w * (float)(transparency * (1.0 - fresnelAmount));
                has_t = wt >= kRrCutWeight;
            } else {
                fresnelAmount = 1.0;  // total internal reflection
            }
            reflectiveness = fresnelAmount;
        }
        bool has_r = reflectiveness > bias64;
        const float wr = w * (float)reflectiveness;
        has_r = has_r && wr >= kRrCutWeight;
        depth = node_depth + 1;
        if (has_r) {
            const d3 R = normalize(reflect(incoming, N));  // reflectDir = incoming.reflect(normal).normalize(); origin hitPoint + reflectDir * bias
            const d3 ro = Pn + R * bias64;
            if (kTree && has_t) {  // the reflected ray waits on the stack, the refracted one is traced next
                stk[sp].o = ro; stk[sp].d = R; stk[sp].w = wr; stk[sp].depth = depth; ++sp;
                o = to; d = td; w = wt;
            } else { o = ro; d = R; w = wr; }
        } else if (kTree && has_t) {
            o = to; d = td; w = wt;
        } else {
            live = false;
        }
    }
    out.r = ar; out.g = ag; out.b = ab;
    return ok || !alive;
}

// ------------------------------------------------------------------------------ the kernel
// Persistent CTAs, one 8 x 4 pixel tile per warp at a time (first tile by position, then a global counter), one pixel per lane.
constexpr int kRrTileW = 8, kRrTileH = 4;

template <bool kTree>
__global__ void __launch_bounds__(kRrThreads, 2) render32r_kernel(const __grid_constant__ RenderParams P) {
    __shared__ SceneRr S;
    __shared__ unsigned long long s_stats[kRrThreads / 32][kStatCount];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    {   // scene constants: one thread per primitive / light
        const DevScene& g = P.sc;
        const int t = (int)threadIdx.x;
        if (t == 0) { S.ns = g.ns; S.nt = g.nt; S.np = g.np; S.nl = g.nl; S.cam[0] = P.cam.pos[0]; S.cam[1] = P.cam.pos[1]; S.cam[2] = P.cam.pos[2]; }
        for (int i = t; i < g.ns; i += kRrThreads) {
            const double4 s = g.sph64[i];
            Sph32r& o = S.sph[i];
            split64(s.x, o.hx, o.lx); split64(s.y, o.hy, o.ly); split64(s.z, o.hz, o.lz);
            o.r = (float)fabs(s.w); o.mat = g.sph_mat[i];
        }
        for (int i = t; i < g.nt; i += kRrThreads) {
            const DevTri64 T = g.tri64[i];
            Tri32& o = S.tri[i];
            for (int c = 0; c < 3; ++c) {
                split64(T.a0[c], o.ah[c], o.al[c]);
                o.e1[c] = (float)T.e1[c]; o.e2[c] = (float)T.e2[c]; o.n[c] = (float)T.n[c];
            }
            o.E1 = (float)(sqrt64_fast(T.e1[0] * T.e1[0] + T.e1[1] * T.e1[1] + T.e1[2] * T.e1[2]) * 1.000001);
            o.E2 = (float)(sqrt64_fast(T.e2[0] * T.e2[0] + T.e2[1] * T.e2[1] + T.e2[2] * T.e2[2]) * 1.000001);
            o.mat = g.tri_info[i].x; o.report = (uint32_t)g.tri_info[i].y;
            const double cx = T.a0[0] + (T.e1[0] + T.e2[0]) * (1.0 / 3.0), cy = T.a0[1] + (T.e1[1] + T.e2[1]) * (1.0 / 3.0), cz = T.a0[2] + (T.e1[2] + T.e2[2]) * (1.0 / 3.0);
            double r2 = 0.0;
            for (int v = 0; v < 3; ++v) {
                const double vx = T.a0[0] + (v == 1 ? T.e1[0] : v == 2 ? T.e2[0] : 0.0) - cx, vy = T.a0[1] + (v == 1 ? T.e1[1] : v == 2 ? T.e2[1] : 0.0) - cy,
                             vz = T.a0[2] + (v == 1 ? T.e1[2] : v == 2 ? T.e2[2] : 0.0) - cz;
                r2 = fmax(r2, vx * vx + vy * vy + vz * vz);
            }
            // (FP32 centre: its rounding, <= 1e-7 of its size, goes into the radius)
            S.tb[i] = make_float4((float)cx, (float)cy, (float)cz, (float)(sqrt64_fast(r2) * 1.00001 + 1e-6 * (fabs(cx) + fabs(cy) + fabs(cz)) + 1e-30));
        }
        for (int i = t; i < g.np; i += kRrThreads) {
            const DevPlane p = g.planes[i];
            S.pl[i].nx = (float)p.n[0]; S.pl[i].ny = (float)p.n[1]; S.pl[i].nz = (float)p.n[2];
            S.pl[i].d = (float)(p.n[0] * p.p0[0] + p.n[1] * p.p0[1] + p.n[2] * p.p0[2]);
            S.pl[i].mat = p.mat;
        }
        for (int j = t; j < g.nl; j += kRrThreads) {
            const DevLight l = g.lights[j];
            for (int c = 0; c < 3; ++c) { split64(l.pos[c], S.li[j].hi[c], S.li[j].lo[c]); S.li[j].E[c] = (float)l.E[c]; }
        }
    }
    if (lane < kStatCount) s_stats[warp][lane] = 0ull;
    __syncthreads();
    if (threadIdx.x == 0) {  // one sphere around all the triangles' spheres
        float cx = 0.f, cy = 0.f, cz = 0.f, r = 0.f;
        const int nt = P.sc.nt;
        for (int i = 0; i < nt; ++i) { cx += S.tb[i].x; cy += S.tb[i].y; cz += S.tb[i].z; }
        if (nt > 0) { const float inv = 1.0f / (float)nt; cx *= inv; cy *= inv; cz *= inv; }
        for (int i = 0; i < nt; ++i) {
            const float dx = S.tb[i].x - cx, dy = S.tb[i].y - cy, dz = S.tb[i].z - cz;
            r = fmaxf(r, sqrtf(dx * dx + dy * dy + dz * dz) * 1.00001f + S.tb[i].w);
        }
        S.tb_all = make_float4(cx, cy, cz, r * 1.00001f);
    }
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
    const int tiles_x = (W + kRrTileW - 1) / kRrTileW, tiles_y = (rows + kRrTileH - 1) / kRrTileH;
    const unsigned int n_tiles = (unsigned int)(tiles_x * tiles_y);
    const int lx = lane & (kRrTileW - 1), ly = lane / kRrTileW;
    const int aa = P.cam.aa;
    const unsigned int n_static = gridDim.x * (kRrThreads / 32);
    unsigned int tile = blockIdx.x * (kRrThreads / 32) + warp;
    while (tile < n_tiles) {
        unsigned int next = 0;
        if (lane == 0) next = n_static + atomicAdd(P.tile_counter, 1u);
        const int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile - (unsigned)ty * (unsigned)tiles_x);
        const int band_y = ty * kRrTileH + ly, px = tx * kRrTileW + lx;
        const int py = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
        const bool inb = px < W && band_y < rows;
        const int64_t pix = (int64_t)band_y * W + px;
        // GeneratePixelAt, RE/Scene.h:283-304: mean of the samples
        float cr = 0.0f, cg = 0.0f, cb = 0.0f;
        uint32_t prim = kMiss, lmask = 0u, n_shadow = 0u, n_sec = 0u, n_sph = 0u, n_pl = 0u, n_tri = 0u;
        bool ok = true;
#pragma unroll 1
        for (int s = 0; s < aa; ++s) {
            PxR o;
            ok = trace32r<kTree>(P, S, px, py, inb && ok, s, o) && ok;
            cr += o.r; cg += o.g; cb += o.b;
            n_shadow += o.c_shadow; n_sec += o.c_secondary; n_sph += o.c_sph; n_pl += o.c_pl; n_tri += o.c_tri;
            if (s == 0) { prim = o.prim; lmask = o.lmask; }
#ifdef RTE_RR_DEBUG
            if (s == 0 && !ok) lmask = 0x80000000u | o.why;
#endif
        }
        if (aa > 1) { const float ia = 1.0f / (float)aa; cr *= ia; cg *= ia; cb *= ia; }
#ifdef RTE_RR_DEBUG
        ok = true;  // (debug build: no exact pass, light_mask carries the reason code of the first inconclusive decision)
#endif
        if (__any_sync(kFullMask, inb && !ok)) {
            // an inconclusive decision in this tile: the exact kernel renders the whole tile (and counts its rays)
            const uint32_t n_px = (uint32_t)__popc(__ballot_sync(kFullMask, inb));
            if (lane == 0) { P.tile_flags[tile] = 1; s_stats[warp][kStatConfirms] += (unsigned long long)n_px * (unsigned)aa; }  // pixel samples left to the exact kernel
        } else {
            if (inb) {
                if (P.out.hdr) __stcs(reinterpret_cast<float4*>(P.out.hdr) + pix, make_float4(cr, cg, cb, 0.f));
                if (P.out.prim_id) P.out.prim_id[pix] = prim;
                if (P.out.light_mask) P.out.light_mask[pix] = lmask;
                if (P.out.rgb8) {
                    const tm32::Shared sh = tm32::prepare<0u>(tm32::mk3(cr, cg, cb), 1u << P.tonemap_op);
                    tm32::f3 c;
                    switch (P.tonemap_op) {
                        case RTE_TM_SIMPLE: c = tm32::apply_op<RTE_TM_SIMPLE>(sh, K); break;
                        case RTE_TM_REINHARD_SIMPLE: c = tm32::apply_op<RTE_TM_REINHARD_SIMPLE>(sh, K); break;
                        case RTE_TM_REINHARD_EXTENDED: c = tm32::apply_op<RTE_TM_REINHARD_EXTENDED>(sh, K); break;
                        case RTE_TM_REINHARD_EXTENDED_LUMINANCE: c = tm32::apply_op<RTE_TM_REINHARD_EXTENDED_LUMINANCE>(sh, K); break;
                        case RTE_TM_REINHARD_JODIE: c = tm32::apply_op<RTE_TM_REINHARD_JODIE>(sh, K); break;
                        case RTE_TM_UNCHARTED2: c = tm32::apply_op<RTE_TM_UNCHARTED2>(sh, K); break;
                        default: c = tm32::apply_op<RTE_TM_ACES>(sh, K); break;
                    }
                    uint8_t* d = P.out.rgb8 + 3 * pix;
                    d[0] = (uint8_t)(tm32::qbyte(c.x) & 255u); d[1] = (uint8_t)(tm32::qbyte(c.y) & 255u); d[2] = (uint8_t)(tm32::qbyte(c.z) & 255u);
                }
            }
            const uint32_t v[kStatCount] = {inb ? (uint32_t)aa : 0u, n_shadow, n_sec, n_sph, n_pl, n_tri, 0u, 0u, 0u};
            uint32_t mine = 0;
#pragma unroll
            for (int k = 0; k < kStatCount; ++k) {
                const uint32_t t = __reduce_add_sync(kFullMask, inb ? v[k] : 0u);
                if (lane == k) mine = t;
            }
            if (lane < kStatCount) s_stats[warp][lane] += mine;
        }
        tile = __shfl_sync(kFullMask, next, 0);
    }
    __syncwarp();
    if (lane < kStatCount && s_stats[warp][lane]) atomicAdd(&P.stats[lane], s_stats[warp][lane]);
}

// Can this render call take the reflective FP32 path?  Specular and / or transparent materials, few
// primitives (every ray tests all of them), outputs hdr / rgb8 / prim_id / light_mask, no validation switches.
bool fp32r_path_eligible(const RenderParams& p) {
    const rte_frame& f = p.out;
    return p.sc.reflective && p.sc.ns <= kRrPrims && p.sc.nt <= kRrPrims && p.sc.ns + p.sc.nt <= kRrPrims && p.sc.ns + p.sc.nt + p.sc.np > 0 &&
           p.sc.np <= kRrPlanes && p.sc.nl <= kRrLights && !p.exact_shadows && !p.no_filter && !p.force_general && !p.no_cull && p.pass_mode == 0 &&
           !p.exact_fp64 && !p.exact_grid && !f.hdr64 && !f.depth && !f.depth64 && !f.depth_norm && !f.normal && !f.albedo && p.bias >= 1e-6 && p.bias <= 1e-1;
}

cudaError_t launch_render32r(const RenderParams& p, int sm_count, cudaStream_t stream) {
    const long long n_tiles = (long long)((p.cam.width + kRrTileW - 1) / kRrTileW) * ((p.n_rows + kRrTileH - 1) / kRrTileH);
    long long blocks = (long long)sm_count * 2;
    const long long needed = (n_tiles + kRrThreads / 32 - 1) / (kRrThreads / 32);
    if (blocks > needed) blocks = needed > 0 ? needed : 1;
    if (p.sc.transparent) render32r_kernel<true><<<(unsigned)blocks, kRrThreads, 0, stream>>>(p);  // refraction: TraceRay is a tree
    else render32r_kernel<false><<<(unsigned)blocks, kRrThreads, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace rte
