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

// Scene::IntersectClosest (RE/Scene.h:218-257) over {spheres of the cell + near list of `view`} and all planes.  The lists are
// unordered, so "the first of equal distances wins" (strict '<' in canonical order) is applied by index.
__device__ __forceinline__ void closest_hit_grid(const DevScene& sc, const DevGrid& g, int view, float ux, float uy, float uz, d3 o, d3 d, Hit& best) {
    best.t = 0.0; best.kind = RTE_HIT_NONE; best.idx = 0;
    uint32_t cell = grid_cell(g, view, ux, uy, uz);
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const uint32_t e0 = g.cell_start[cell], e1 = g.cell_start[cell + 1];
#pragma unroll 1
        for (uint32_t e = e0; e < e1; ++e) {
            const uint32_t k = g.entries[e];
            double t;
            if (sphere_exact(sc.sph64[k], o, d, t) && (best.kind == RTE_HIT_NONE || t < best.t || (t == best.t && k < best.idx))) {
                best.t = t; best.kind = RTE_HIT_SPHERE; best.idx = k;
            }
        }
        cell = grid_near_cell(g, view);
    }
#pragma unroll 1
    for (int i = 0; i < sc.np; ++i) {
        double t;
        if (plane_exact(sc.planes[i], o, d, t) && (best.kind == RTE_HIT_NONE || t < best.t)) { best.t = t; best.kind = RTE_HIT_PLANE; best.idx = (uint32_t)i; }
    }
}

// One pixel sample, exactly: TraceRay + directLightning + computeTransmittance (RE/Scene.h:131-198, 79-129, 35-77) for scenes
// without specular / transparent materials and without triangles (what the FP32 path accepts) — in two parts, so that the
// lanes of a warp can share one pixel: every lane evaluates the primary hit, lane j the light j (rte_render32.cu: fix_batch).
struct ExactHit {
    bool hit;
    d3 Pn, normal;   // hit point, viewer-facing unit normal (directLightning re-normalises, RE/Scene.h:81)
    d3 sky;          // Scene::backgroundColor of a miss
    int mat;
    uint32_t prim;   // (type << 30) | index, kMiss
};
static __device__ __noinline__ void exact_primary(const RenderParams& P, const DevScene& sc, int px, int py, int sample, ExactHit& out) {
    const d3 cam_o = ld3(P.cam.pos);
    const d3 d = camera_dir(P.cam, px, py, sample);
    Hit h;
    closest_hit_grid(sc, P.grid, 0, (float)d.x, (float)d.y, (float)d.z, cam_o, d, h);
    out.hit = h.kind != RTE_HIT_NONE;
    out.prim = kMiss; out.mat = sc.nm;
    out.Pn = mk(0.0, 0.0, 0.0); out.normal = mk(0.0, 0.0, 1.0);
    out.sky = mk(0.0, 0.0, 0.0);
    if (!out.hit) { out.sky = mk(0.0, 0.0, 0.0) + background(d) * 1.0; return; }  // sky (RE/Scene.h:137-141)
    out.Pn = cam_o + d * h.t;
    d3 ng;
    if (h.kind == RTE_HIT_SPHERE) { const double4 s = sc.sph64[h.idx]; ng = normalize(out.Pn - mk(s.x, s.y, s.z)); }
    else ng = ld3(sc.planes[h.idx].n);
    out.prim = (h.kind << 30) | h.idx;
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
    {   // algorithmic test counts: canonical order up to and including the first occluder
        uint32_t first = (uint32_t)sc.ns;
        uint32_t cell = grid_cell(g, 1 + j, ux, uy, uz);
        for (int pass = 0; pass < 2; ++pass) {
            for (uint32_t e = g.cell_start[cell]; e < g.cell_start[cell + 1]; ++e) {
                const uint32_t k = g.entries[e];
                double t;
                if (sphere_exact(sc.sph64[k], o, L, t) && t > bias && t < maxDist && k < first) first = k;
            }
            cell = grid_near_cell(g, 1 + j);
        }
        if (first < (uint32_t)sc.ns) cnt.sph += first + 1u;
        else {
            cnt.sph += (uint32_t)sc.ns;
            uint32_t firstp = (uint32_t)sc.np;
            for (int i = 0; i < sc.np; ++i) {
                double t;
                if (plane_exact(sc.planes[i], o, L, t) && t > bias && t < maxDist) { firstp = (uint32_t)i; break; }
            }
            cnt.pl += firstp < (uint32_t)sc.np ? firstp + 1u : (uint32_t)sc.np;
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
