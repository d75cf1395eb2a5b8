// rte_exact_grid.cuh — one pixel sample with the reference's exact FP64 operations, in the reference's order, over the
// candidates of the ray's direction-grid cell (rte_grid.cuh) instead of over all spheres.  The FP32 production kernel
// (rte_render32.cu) calls it for the pixels whose FP32 decisions were inconclusive (~1e-3 of a frame: contact shadows within
// `bias`, grazing rays, N.L ~ 0): those pixels are the reference's own bits.  "RE/" = /root/reference/RaytracingEngine/.
#pragma once
#include "rte_kernels.h"
#include "rte_trace.cuh"

namespace rte {

// Scene::backgroundColor, RE/Scene.h:30-33
__device__ __forceinline__ d3 background(d3 dir) {
    const double t = 0.5 * (normalize(dir).y + 1.0);
    return mk(1.0, 1.0, 1.0) * (1.0 - t) + mk(0.5, 0.7, 1.0) * t;
}

struct PixelCounters {
    uint32_t shadow, secondary, sph, pl, tri, marches, confirms, sph_exec;
};

// Scene::IntersectClosest (RE/Scene.h:218-257) over {primitives of the cell + near list of `view`} and all planes.  The lists are
// unordered, so the reference's rule — canonical order spheres, planes, triangles (each by ascending index), replace only when
// strictly closer, i.e. of equal distances the FIRST wins — is applied explicitly: a sphere beats everything it ties with except a
// sphere of lower index, a plane beats the triangles it ties with, a triangle only triangles of higher index.
__device__ __forceinline__ void closest_hit_grid(const DevScene& sc, const DevGrid& g, int view, float ux, float uy, float uz, d3 o, d3 d, Hit& best) {
    best.t = 0.0; best.kind = RTE_HIT_NONE; best.idx = 0;
    const uint32_t ns = (uint32_t)sc.ns;
    uint32_t cell = grid_cell(g, view, ux, uy, uz);
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const uint32_t e0 = g.cell_start[cell], e1 = g.cell_start[cell + 1];
#pragma unroll 1
        for (uint32_t e = e0; e < e1; ++e) {
            const uint32_t k = g.entries[e];
            double t;
            if (k < ns) {
                if (sphere_exact(sc.sph64[k], o, d, t) &&
                    (best.kind == RTE_HIT_NONE || t < best.t || (t == best.t && (best.kind != RTE_HIT_SPHERE || k < best.idx)))) {
                    best.t = t; best.kind = RTE_HIT_SPHERE; best.idx = k;
                }
            } else {
                const uint32_t ti = k - ns;
                if (tri_exact(sc.tri64[ti], o, d, t) && (best.kind == RTE_HIT_NONE || t < best.t || (t == best.t && best.kind == RTE_HIT_TRIANGLE && ti < best.idx))) {
                    best.t = t; best.kind = RTE_HIT_TRIANGLE; best.idx = ti;
                }
            }
        }
        cell = grid_near_cell(g, view);
    }
#pragma unroll 1
    for (int i = 0; i < sc.np; ++i) {
        double t;
        if (plane_exact(sc.planes[i], o, d, t) && (best.kind == RTE_HIT_NONE || t < best.t || (t == best.t && best.kind == RTE_HIT_TRIANGLE))) {
            best.t = t; best.kind = RTE_HIT_PLANE; best.idx = (uint32_t)i;
        }
    }
}

// One pixel sample, exactly: TraceRay + directLightning + computeTransmittance (RE/Scene.h:131-198, 79-129, 35-77) for scenes
// without specular materials (no secondary rays; transparent occluders are handled by the literal march) — in two parts, so that the
// lanes of a warp can share one pixel: every lane evaluates the primary hit, lane j the light j (rte_render32.cu: fix_batch).
struct ExactHit {
    bool hit;
    d3 Pn, normal;   // hit point, viewer-facing unit normal (directLightning re-normalises, RE/Scene.h:81)
    d3 sky;          // Scene::backgroundColor of a miss
    d3 ngeo;         // HitInfo.normal (geometric, un-flipped)
    double t;        // HitInfo.distance
    int mat;
    uint32_t prim;   // (type << 30) | reported index (triangles of a model report the MODEL index, RE/Shape.h:276), kMiss
};
static __device__ __noinline__ void exact_primary(const RenderParams& P, const DevScene& sc, int px, int py, int sample, ExactHit& out) {
    const d3 cam_o = ld3(P.cam.pos);
    const d3 d = camera_dir(P.cam, px, py, sample);
    Hit h;
    closest_hit_grid(sc, P.grid, 0, (float)d.x, (float)d.y, (float)d.z, cam_o, d, h);
    out.hit = h.kind != RTE_HIT_NONE;
    out.prim = kMiss; out.mat = sc.nm;
    out.Pn = mk(0.0, 0.0, 0.0); out.normal = mk(0.0, 0.0, 1.0); out.ngeo = mk(0.0, 0.0, 0.0); out.t = 0.0;
    out.sky = mk(0.0, 0.0, 0.0);
    if (!out.hit) { out.sky = mk(0.0, 0.0, 0.0) + background(d) * 1.0; return; }  // sky (RE/Scene.h:137-141)
    out.t = h.t;
    out.Pn = cam_o + d * h.t;
    d3 ng;
    uint32_t report = h.idx;
    if (h.kind == RTE_HIT_SPHERE) { const double4 s = sc.sph64[h.idx]; ng = normalize(out.Pn - mk(s.x, s.y, s.z)); }
    else if (h.kind == RTE_HIT_PLANE) ng = ld3(sc.planes[h.idx].n);
    else { ng = ld3(sc.tri64[h.idx].n); report = (uint32_t)sc.tri_info[h.idx].y; }
    out.ngeo = ng;
    out.prim = (h.kind << 30) | report;
    out.mat = material_of(sc, h);
    const d3 incoming = normalize(d);
    const bool frontFace = dot(ng, incoming) < 0.0;
    out.normal = normalize(frontFace ? ng : -ng);
}
// Light j at an exact hit: returns contribution * transmittance (what directLightning adds to diffuseAccumulation, RE/Scene.h:113)
// and the light_mask bits (bit 16: a shadow ray was cast, bit 0: the light contributes); counts its tests into `cnt`.
static __device__ __noinline__ d3 exact_light(const RenderParams& P, const DevScene& sc, const ExactHit& H, int j, uint32_t& bits, PixelCounters& cnt) {
    const DevGrid& g = P.grid;
    const double bias = P.bias;
    bits = 0u;
    const d3 zero = mk(0.0, 0.0, 0.0);
    if (!H.hit) return zero;
    const d3 Lp = ld3(sc.lights[j].pos);
    const d3 vecToLight = Lp - H.Pn;
    const double dist = length(vecToLight);
    if (!(dist > 0.0)) return zero;
    const d3 L = vecToLight / dist;
    const double ndl = std_max(0.0, dot(H.normal, L));
    if (!(ndl > 0.0) || !(dist > bias)) return zero;
    cnt.shadow += 1;
    bits = 1u << 16;
    // computeTransmittance, literally; every closest-hit query of the march looks at the spheres of the light's cell in the
    // direction of this ray (the march moves its origin along the same line: no other sphere can be hit before the light)
    d3 o = H.Pn + H.normal * bias;
    const double maxDist = dist - bias;
    const float ux = (float)(-L.x), uy = (float)(-L.y), uz = (float)(-L.z);
    {   // algorithmic test counts: canonical order (spheres, planes, triangles) up to and including the first occluder
        const uint32_t ns = (uint32_t)sc.ns, nt = (uint32_t)sc.nt;
        uint32_t first = ns, first_t = nt;
        uint32_t cell = grid_cell(g, 1 + j, ux, uy, uz);
        for (int pass = 0; pass < 2; ++pass) {
            for (uint32_t e = g.cell_start[cell]; e < g.cell_start[cell + 1]; ++e) {
                const uint32_t k = g.entries[e];
                double t;
                if (k < ns) { if (k < first && sphere_exact(sc.sph64[k], o, L, t) && t > bias && t < maxDist) first = k; }
                else if (k - ns < first_t && tri_exact(sc.tri64[k - ns], o, L, t) && t > bias && t < maxDist) first_t = k - ns;
            }
            cell = grid_near_cell(g, 1 + j);
        }
        if (first < ns) cnt.sph += first + 1u;
        else {
            cnt.sph += ns;
            uint32_t firstp = (uint32_t)sc.np;
            for (int i = 0; i < sc.np; ++i) {
                double t;
                if (plane_exact(sc.planes[i], o, L, t) && t > bias && t < maxDist) { firstp = (uint32_t)i; break; }
            }
            if (firstp < (uint32_t)sc.np) cnt.pl += firstp + 1u;
            else { cnt.pl += (uint32_t)sc.np; cnt.tri += first_t < nt ? first_t + 1u : nt; }
        }
    }
    double T = 1.0, traveled = 0.0;
    int safety = 64;
    bool stepped = false;
    while (safety-- > 0 && T > 1e-4 && traveled < maxDist) {
        Hit sh;
        closest_hit_grid(sc, g, 1 + j, ux, uy, uz, o, L, sh);
        if (sh.kind == RTE_HIT_NONE) break;
        const double t = sh.t;
        if (t <= 0.0) { o = o + L * bias; traveled += bias; stepped = true; continue; }
        if (t <= bias) { o = (o + L * t) + L * bias; traveled += t + bias; stepped = true; continue; }
        if (traveled + t >= maxDist) break;
        T *= std_clamp(sc.materials[material_of(sc, sh)].transparency, 0.0, 1.0);
        const d3 newOrigin = (o + L * t) + L * bias;
        traveled += t + bias;
        o = newOrigin;
    }
    T = std_clamp(T, 0.0, 1.0);
    if (stepped) cnt.marches += 1;
    if (!(T > bias)) return zero;
    bits |= 1u;
    const d3 emitted = ld3(sc.lights[j].E);
    const d3 contribution = (emitted * div64(1.0, dist * dist)) * ndl;
    return contribution * T;
}
// TraceRay's result from the light sum (RE/Scene.h:126, 171-173) or the sky.
__device__ __forceinline__ d3 exact_finish(const DevScene& sc, const ExactHit& H, d3 diffuse) {
    if (!H.hit) return H.sky;
    const rte_material& M = sc.materials[H.mat];
    const d3 local = ld3(M.color) * diffuse + mk(0.0, 0.0, 0.0) * M.specular;
    return mk(0.0, 0.0, 0.0) + (local * (1.0 - 0.0)) * 1.0;  // finalLight += localLight * (1 - transparency)
}

}  // namespace rte
