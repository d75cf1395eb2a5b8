// rte_render.cu — the EXACT (FP64, bit-identical) per-pixel render kernels (sm_100a); the FP32 production path is rte_render32.cu.
//
// Replaces Scene::RenderImage / GeneratePixelAt / TraceRay / directLightning
// (RE/Scene.h:311-328, 283-304, 131-198, 79-129; "RE/" = /root/reference/RaytracingEngine/).
//
//  * persistent grid (a multiple of the SM count), every warp owns one 8x4 pixel tile at a time and
//    pulls the next one from a global atomic counter (sky rows are cheap, sphere-dense rows are not);
//  * the FP32 filter arrays of the scene are staged once per CTA into shared memory with 1-D TMA
//    bulk copies (cp.async.bulk + mbarrier complete_tx);
//  * per pixel the ray state is FP64 (bit-identical to the reference's doubles), the O(primitives)
//    loops are FP32 FMA filters (rte_trace.cuh);
//  * reflection / refraction / Fresnel (RE/Scene.h:161-187) run as an explicit per-pixel stack (trace_pixels<K, true>);
//  * diffuse scenes of any size take render_exact_grid_kernel: the same FP64 tests over the direction grids (rte_exact_grid.cuh).
#include <vector>

#include "rte_exact_grid.cuh"
#include "rte_kernels.h"
#include "rte_stage.cuh"
#include "rte_tonemap.cuh"
#include "rte_trace.cuh"
#include "rte_views.cuh"

namespace rte {

constexpr int kMaxTreeDepth = RTE_MAX_DEPTH;  // rte.h: largest opts.max_depth (frames of the per-pixel TraceRay stack)
// Pixels per thread of the view-mode render kernel (a warp's tile is 8 x 4K pixels).  Measured on B200: with resident views one
// pixel per thread is fastest (fewer live registers, tighter ray cones for the tile cull: C2 4.78 -> 4.60 ms, C4 1.40 -> 0.95 ms);
// with streamed views two pixels per thread used to halve the per-tile pass over ALL view blocks (C5 19.1 -> 15.8 ms); since the
// CTA-level cull (rte_views.cuh) that pass walks a few hundred survivors and one pixel per thread wins there too (C5 10.5 -> 9.6 ms).
#ifndef RTE_VIEW_PIXELS
#define RTE_VIEW_PIXELS 1
#endif
#ifndef RTE_STREAM_PIXELS
#define RTE_STREAM_PIXELS 1
#endif
#ifndef RTE_P1_BATCH
#define RTE_P1_BATCH 1  // primary-ray launch of scenes with streamed views: CTA batches of 4 x 8 tiles + CTA-level cull of the camera's view
#endif
#ifndef RTE_P2_STATE
#define RTE_P2_STATE 0  // A/B: hit point + normal of the shading launch in shared memory (kSt); measured 2.69 -> 2.78 ms on C2: off
#endif
#ifndef RTE_P2_BIG
#define RTE_P2_BIG 0  // A/B: the shading launch of the two-launch path at kBigThreads (more registers per thread)
#endif
#ifndef RTE_SMEM_CONSTS
#define RTE_SMEM_CONSTS 1
#endif
constexpr int kSmemPlanes = 8, kSmemLights = 16;
#ifndef RTE_BATCH_W
#define RTE_BATCH_W 4
#endif
#ifndef RTE_FORCE_STREAM
#define RTE_FORCE_STREAM 0  // A/B: stream the views (lock-step CTA, CTA-level cull) even when they would fit shared memory
#endif
constexpr int kViewPixels = RTE_VIEW_PIXELS;
constexpr int kStreamPixels = RTE_STREAM_PIXELS;

// Dynamic shared memory layout of the render kernel: [views (1+nl) x ns_pad x 16 B][generic sph32][tri32].
// The generic sphere array is only staged when something can use it often (reflection rays, or no views);
// otherwise the rare generic queries (exact shadow marches) read it from global memory.
SmemPlan plan_smem(const DevScene& sc, size_t budget) {
    SmemPlan pl{};
    const size_t tri = (size_t)sc.nt_pad * 48, sph = (size_t)sc.ns_pad * 16;
    const size_t views = sc.ns > 0 ? (size_t)(1 + sc.nl) * view_bytes(sc.ns_pad) : 0;
    const size_t tviews = (sc.nt > 0 && sc.tri_views) ? (size_t)(1 + sc.nl) * view_bytes(sc.nt_pad32) : 0;
    pl.one_view_bytes = (uint32_t)view_bytes(sc.ns_pad);
    pl.stage_tri = 1;
    if ((sc.views || sc.ns == 0) && tviews > 0 && views + tviews <= budget) {
        // triangle scenes (C1/C4): sphere and triangle views resident; the generic arrays only where reflection rays use them
        pl.use_views = sc.ns > 0 ? 1 : 0;
        pl.use_tri_views = 1;
        pl.view_bytes = (uint32_t)views;
        pl.tview_bytes = (uint32_t)tviews;
        pl.stage_tri = (sc.reflective && views + tviews + tri <= budget) ? 1 : 0;
        pl.stage_sph = (sc.reflective && sc.ns > 0 && views + tviews + (pl.stage_tri ? tri : 0) + sph <= budget) ? 1 : 0;
    } else if (sc.views && tri + views <= budget && !(RTE_FORCE_STREAM && !sc.reflective && sc.nt == 0 && sc.ns_pad >= 128)) {
        pl.use_views = 1;
        pl.view_bytes = (uint32_t)views;
        pl.stage_sph = (sc.reflective && tri + views + sph <= budget) ? 1 : 0;
    } else if (sc.views && !sc.reflective && sc.nt == 0 && 2 * (size_t)pl.one_view_bytes <= budget) {
        // large diffuse sphere scenes (C5: 9 views of 64 KB): two view buffers, streamed (fast kernel only)
        pl.use_views = 1;
        pl.stream_views = 1;
        pl.view_bytes = 2 * pl.one_view_bytes;
        pl.stage_sph = 0;
    } else {
        pl.stage_sph = 1;
    }
    // the generic arrays are staged only while they fit next to the views: what does not fit is read from global memory (L1 / L2)
    if (pl.stage_tri && (size_t)pl.view_bytes + pl.tview_bytes + tri > budget) pl.stage_tri = 0;
    if (pl.stage_sph && (size_t)pl.view_bytes + pl.tview_bytes + (pl.stage_tri ? tri : 0) + sph > budget) pl.stage_sph = 0;
    pl.sph_bytes = pl.stage_sph ? (uint32_t)sph : 0u;
    pl.tri_bytes = pl.stage_tri ? (uint32_t)tri : 0u;
    pl.total = (size_t)pl.view_bytes + pl.tview_bytes + pl.sph_bytes + pl.tri_bytes;
    return pl;
}

static SmemPlan generic_plan(const DevScene& sc);
size_t scene_smem_bytes(const DevScene& sc) { return (size_t)sc.ns_pad * 16 + (size_t)sc.nt_pad * 48; }

// TraceRay for the K pixels of one thread, all 32 lanes of the warp in lock step (pixels outside the image:
// alive = false).  kReflect = false compiles the diffuse-only path (no material has specular > 0): the depth loop,
// Blinn-Phong and the reflection state disappear.  Per pixel the ray state is FP64, bit-identical to the reference.
template <int K, bool kReflect>
__device__ __forceinline__ void trace_pixels(const RenderParams& P, const DevScene& sc, const ulonglong2* s_views, const ulonglong2* s_tviews,
                                             const float* s_cold, const float* s_tcold, const float4* s_sph, const float4* s_tri, const int (&px)[K], const int (&py)[K], const bool (&inb)[K],
                                             const int64_t (&out_idx)[K], int sample, PixelCounters& cnt, d3 (&result)[K]) {
    const d3 cam_o = ld3(P.cam.pos);
    d3 o[K], d[K], acc[K];
    double w[K];
    bool alive[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        o[k] = cam_o;
        d[k] = camera_dir(P.cam, px[k], py[k], sample);
        acc[k] = mk(0.0, 0.0, 0.0);
        w[k] = 1.0;
        alive[k] = inb[k];
    }
    const double bias = P.bias;
    const int max_depth = kReflect ? P.max_depth : 1, nl = sc.nl;
    const bool nf = P.no_filter != 0;
    const bool utv = P.plan.use_tri_views != 0;
    const bool uv = P.plan.use_views != 0 || utv;  // (triangle views exist only next to sphere views, or without spheres)
    const int view_stride = sc.ns_pad, tview_stride = sc.nt_pad32;  // ulonglong2 elements per view
    // kReflect: TraceRay's recursion (RE/Scene.h:131-198) as an explicit per-pixel stack, evaluated in the reference's own
    // order — local light, then the refracted subtree, then the reflected subtree, each added with its coefficient when
    // it returns — so that the sums round exactly like the recursive calls.  One loop iteration = one TraceRay node per
    // lane (the lanes of a warp sit at different nodes of their trees; the light views serve all of them alike).
    struct TreeFrame {
        d3 acc;          // finalLight of the node so far
        d3 ro, rd;       // the reflected ray, traced after the refracted subtree has returned
        double k_pending, k_refl;  // coefficient of the subtree being traced / of the reflected subtree
        int has_refl;
    };
    TreeFrame stk[K][kReflect ? kMaxTreeDepth : 1];
    int dep[K];  // frames on the stack = recursionAmount of the node being evaluated
#pragma unroll
    for (int k = 0; k < K; ++k) dep[k] = 0;
    auto tree_return = [&](int k, d3 v) {  // the current node of pixel k returns v to its caller
        while (true) {
            if (dep[k] == 0) { acc[k] = v; alive[k] = false; return; }
            TreeFrame& f = stk[k][dep[k] - 1];
            f.acc = f.acc + v * f.k_pending;
            if (f.has_refl) {  // next: the reflected subtree of the same node
                f.has_refl = 0; f.k_pending = f.k_refl;
                o[k] = f.ro; d[k] = f.rd;
                if (dep[k] < max_depth) return;
                v = background(d[k]);  // TraceRay(depth >= maxRecursion), RE/Scene.h:132-134
            } else {
                v = f.acc;
                dep[k] -= 1;
            }
        }
    };
#pragma unroll 1
    for (int depth = 0; kReflect || depth < 1; ++depth) {  // depth = iteration; all lanes are at their primary ray in iteration 0
        if (kReflect) {
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (alive[k] && dep[k] >= max_depth) tree_return(k, background(d[k]));
        }
        bool any_alive = false;
#pragma unroll
        for (int k = 0; k < K; ++k) any_alive = any_alive || alive[k];
        if (!__any_sync(kFullMask, any_alive)) break;
        Hit h[K];
        if (uv && depth == 0) {
            closest_hit_view<K>(sc, s_views, s_tviews, s_cold, s_tcold, utv, s_tri, cam_o, d, alive, nf, h, cnt.confirms);
#pragma unroll
            for (int k = 0; k < K; ++k) if (alive[k]) cnt.sph_exec += (uint32_t)sc.ns_pad;
        } else {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                closest_hit_body(sc, s_sph, s_tri, o[k], d[k], alive[k], nf, h[k], cnt.confirms);
                if (alive[k]) cnt.sph_exec += (uint32_t)sc.ns_pad;
            }
        }
        // hit records (GetHitInfoAt, RE/Shape.h:104-121,165-180,229-243) and TraceRay's frame, RE/Scene.h:144-148
        bool hitv[K];
        d3 Pn[K], normal[K], ngeo[K], viewDir[K], incoming[K];
        int mat[K], self[K];
        uint32_t report[K], lmask[K];
        d3 diffuseAcc[K], specAcc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (alive[k]) {
                cnt.sph += (uint32_t)sc.ns; cnt.pl += (uint32_t)sc.np; cnt.tri += (uint32_t)sc.nt;
                if (dep[k] > 0) cnt.secondary += 1;
            }
            hitv[k] = alive[k] && h[k].kind != RTE_HIT_NONE;
            if (alive[k] && !hitv[k]) {  // miss: sky (RE/Scene.h:137-141)
                if (kReflect) tree_return(k, background(d[k]));
                else { acc[k] = acc[k] + background(d[k]) * w[k]; alive[k] = false; }
            }
            Pn[k] = mk(0, 0, 0); ngeo[k] = mk(0, 0, 0);
            mat[k] = sc.nm; report[k] = 0; self[k] = -1; lmask[k] = 0;
            if (hitv[k]) {
                Pn[k] = o[k] + d[k] * h[k].t;
                mat[k] = material_of(sc, h[k]);
                if (h[k].kind == RTE_HIT_SPHERE) {
                    const double4 s = sc.sph64[h[k].idx];
                    ngeo[k] = normalize(Pn[k] - mk(s.x, s.y, s.z));
                    report[k] = h[k].idx;
                } else if (h[k].kind == RTE_HIT_PLANE) {
                    ngeo[k] = ld3(sc.planes[h[k].idx].n);
                    report[k] = h[k].idx;
                } else {
                    ngeo[k] = ld3(sc.tri64[h[k].idx].n);
                    report[k] = (uint32_t)sc.tri_info[h[k].idx].y;
                }
            }
            incoming[k] = normalize(d[k]);
            const bool frontFace = dot(ngeo[k], incoming[k]) < 0.0;
            const d3 N = frontFace ? ngeo[k] : -ngeo[k];
            if (hitv[k] && h[k].kind == RTE_HIT_SPHERE && frontFace) self[k] = (int)h[k].idx;
            viewDir[k] = -incoming[k];
            normal[k] = normalize(N);  // directLightning re-normalises, RE/Scene.h:81
            diffuseAcc[k] = mk(0, 0, 0); specAcc[k] = mk(0, 0, 0);
        }
        // directLightning, RE/Scene.h:79-129
#pragma unroll 1
        for (int j = 0; j < nl; ++j) {
            const d3 Lp = ld3(sc.lights[j].pos);
            ShadowRay sr[K];
            bool any_cast = false;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const d3 vecToLight = Lp - Pn[k];
                const double dist = length(vecToLight);
                bool cast = hitv[k] && dist > 0.0;
                d3 L = mk(0, 0, 1);
                double ndl = 0.0;
                if (cast) {
                    L = vecToLight / dist;
                    ndl = std_max(0.0, dot(normal[k], L));
                    cast = ndl > 0.0 && dist > bias;
                }
                sr[k].o = Pn[k] + normal[k] * bias;
                sr[k].d = L; sr[k].dist = dist; sr[k].maxDist = dist - bias; sr[k].ndl = ndl;
                sr[k].self = self[k]; sr[k].cast = cast;
                any_cast = any_cast || cast;
            }
            if (!__any_sync(kFullMask, any_cast)) continue;
            double T[K];
#pragma unroll
            for (int k = 0; k < K; ++k) T[k] = 1.0;
            if (P.exact_shadows) {
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (sr[k].cast) { T[k] = march_exact(sc, s_sph, s_tri, sr[k].o, sr[k].d, sr[k].maxDist, bias, nf, cnt.confirms); cnt.marches += 1; }
            } else {
                OccResult r[K];
                RayLists<K> lists;
#pragma unroll
                for (int k = 0; k < K; ++k) lists.n[k] = kListCap + 1;
                if (uv) {
                    occluded_view<K>(sc, s_views + (size_t)(1 + j) * view_stride, s_tviews + (size_t)(1 + j) * tview_stride,
                                     s_cold + (size_t)(1 + j) * view_stride * 4, s_tcold + (size_t)(1 + j) * tview_stride * 4, utv, s_tri, Lp, sr,
                                     bias, nf, r, lists, cnt.confirms, cnt.sph_exec);
                } else {
#pragma unroll
                    for (int k = 0; k < K; ++k)
                        r[k] = occluded(sc, s_sph, s_tri, sr[k].o, sr[k].d, sr[k].maxDist, bias, sr[k].cast, nf, cnt.confirms);
                }
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    cnt.sph += r[k].sph_tests; cnt.pl += r[k].pl_tests; cnt.tri += r[k].tri_tests;
                    if (sr[k].cast) {
                        if (r[k].near || (sc.transparent && r[k].occluded)) {
                            // triangles are outside march_list's scope; with transparent materials every occluded ray takes the
                            // literal march (T *= transparency per crossing, RE/Scene.h:68-69) and its list may be incomplete
                            if (lists.n[k] <= kListCap && sc.nt == 0 && !nf && !sc.transparent) T[k] = march_list(sc, lists.idx[k], lists.n[k], sr[k].o, sr[k].d, sr[k].maxDist, bias);
                            else T[k] = march_exact(sc, s_sph, s_tri, sr[k].o, sr[k].d, sr[k].maxDist, bias, nf, cnt.confirms);
                            cnt.marches += 1;
                        }
                        else T[k] = r[k].occluded ? 0.0 : 1.0;
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (!sr[k].cast) continue;
                cnt.shadow += 1;
                if (j < 16) lmask[k] |= 1u << (16 + j);
                if (T[k] > bias) {
                    if (j < 16) lmask[k] |= 1u << j;
                    const d3 emitted = ld3(sc.lights[j].E);
                    const double dist = sr[k].dist;
                    const d3 contribution = (emitted * div64(1.0, dist * dist)) * sr[k].ndl;
                    diffuseAcc[k] = diffuseAcc[k] + contribution * T[k];
                    if (kReflect) {
                        const rte_material& M = sc.materials[mat[k]];
                        if (M.transparency <= 0.0 && M.specular > 0.0) {
                            const d3 H = normalize(sr[k].d + viewDir[k]);
                            const double ndh = std_max(0.0, dot(normal[k], H));
                            if (ndh > 0.0) {
                                const double specFactor = pow(ndh, M.shininess);
                                specAcc[k] = specAcc[k] + ((emitted * div64(1.0, dist * dist)) * specFactor) * T[k];
                            }
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const rte_material& M = sc.materials[mat[k]];
            const d3 local = ld3(M.color) * diffuseAcc[k] + specAcc[k] * M.specular;
            if (hitv[k] && !kReflect) {
                // finalLight += localLight * (1.0 - transparency), transparency == 0 (RE/Scene.h:171-173)
                acc[k] = acc[k] + (local * (1.0 - 0.0)) * w[k];
            }
            if (depth == 0 && sample == 0 && inb[k]) {  // AOVs of the primary hit (of the unjittered sample)
                const int64_t oi = out_idx[k];
                if (P.out.prim_id) P.out.prim_id[oi] = hitv[k] ? ((h[k].kind << 30) | report[k]) : kMiss;
                if (P.out.depth) P.out.depth[oi] = hitv[k] ? (float)h[k].t : INFINITY;
                if (P.out.depth64) P.out.depth64[oi] = hitv[k] ? h[k].t : (double)INFINITY;
                if (P.out.normal) reinterpret_cast<float4*>(P.out.normal)[oi] = make_float4((float)ngeo[k].x, (float)ngeo[k].y, (float)ngeo[k].z, 0.f);
                if (P.out.albedo) {
                    const d3 a = hitv[k] ? ld3(M.color) : background(d[k]);
                    reinterpret_cast<float4*>(P.out.albedo)[oi] = make_float4((float)a.x, (float)a.y, (float)a.z, 0.f);
                }
                if (P.out.light_mask) P.out.light_mask[oi] = lmask[k];
            }
            if (!kReflect) {
                alive[k] = false;
            } else if (hitv[k]) {
                // refraction and reflection, RE/Scene.h:161-195
                const bool frontFace = dot(ngeo[k], incoming[k]) < 0.0;
                const d3 N = frontFace ? ngeo[k] : -ngeo[k];
                const double cosTheta = std_max(0.0, dot(N, viewDir[k]));
                const double etaT = M.refractive_index;
                const double transparency = std_clamp(M.transparency, 0.0, 1.0);
                // Scene::fresnel, RE/Scene.h:26-28 — its value only reaches an output through transparent materials (below): opaque
                // hits skip the two pow() calls (a third of the FP64 instructions of an opaque node)
                double fresnelAmount = 0.0;
                if (transparency > 0.0) {
                    const double f0 = pow((etaT - 1.0) / (etaT + 1.0), 2.0);
                    fresnelAmount = f0 + (1.0 - f0) * pow(1.0 - cosTheta, 5.0);
                }
                d3 fin = mk(0.0, 0.0, 0.0);
                if (transparency < 1.0) fin = fin + local * (1.0 - transparency);
                bool has_t = false;
                d3 to = mk(0, 0, 0), td = mk(0, 0, 1);
                double kt = 0.0;
                if (transparency > 0.0) {
                    const double eta = frontFace ? (1.0 / etaT) : (etaT / 1.0);
                    d3 rd = refract(incoming[k], N, eta);
                    if (length(rd) > bias) {
                        rd = normalize(rd);
                        to = Pn[k] + rd * (bias * 1e2); td = rd;
                        has_t = true;
                        kt = transparency * (1.0 - fresnelAmount);
                    } else {
                        fresnelAmount = 1.0;
                    }
                }
                const double reflectiveness = (transparency > 0.0) ? fresnelAmount : M.specular;
                const bool has_r = reflectiveness > bias;
                if (!has_t && !has_r) {
                    tree_return(k, fin);
                } else {
                    TreeFrame f;
                    f.acc = fin; f.ro = mk(0, 0, 0); f.rd = mk(0, 0, 1); f.k_refl = reflectiveness;
                    if (has_r) {
                        const d3 R = normalize(reflect(incoming[k], N));
                        f.ro = Pn[k] + R * bias; f.rd = R;
                    }
                    if (has_t) { f.k_pending = kt; f.has_refl = has_r ? 1 : 0; o[k] = to; d[k] = td; }
                    else { f.k_pending = reflectiveness; f.has_refl = 0; o[k] = f.ro; d[k] = f.rd; }
                    stk[k][dep[k]] = f;
                    dep[k] += 1;
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) result[k] = acc[k];
}

// ------------------------------------------------------------------------------ the kernel
constexpr int kRenderThreads = 256;     // block size of the query kernels and of the render kernel at 2 CTAs/SM
constexpr int kRenderMaxThreads = 512;  // render kernel at 1 CTA/SM (large scenes: more shared memory per CTA)
constexpr int kTileW = 8, kTileH = 4;   // one warp = K stacked 8x4 pixel groups (coherent primary and shadow rays)

// K = pixels per thread: lane (lx, ly) owns pixels (x0 + lx, y0 + ly + 4k), k < K, of its warp's 8 x 4K tile.
#ifndef RTE_RK_T
#define RTE_RK_T 256
#define RTE_RK_B 2
#endif
constexpr int kReflThreads = RTE_RK_T, kReflBlocks = RTE_RK_B;
template <int K, bool kReflect>
__global__ void __launch_bounds__(kReflect ? kReflThreads : kRenderMaxThreads, kReflect ? kReflBlocks : 1) render_kernel(const __grid_constant__ RenderParams P) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ DevScene s_sc;
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ unsigned long long s_stats[kRenderMaxThreads / 32][kStatCount];

    // second pass of rte_render32r.cu: that kernel left the number of pixel samples in flagged tiles in the statistics — none: nothing to do
    if (P.tile_mask && P.stats[kStatConfirms] == 0ull) return;
    if (threadIdx.x == 0) s_sc = P.sc;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < kStatCount) s_stats[warp][lane] = 0ull;
    StagedScene S;
    stage_scene(P.sc, P.plan, dyn_smem, &s_bar, S);
    const float4 *s_sph = S.sph, *s_tri = S.tri;
    const ulonglong2 *s_views = S.views, *s_tviews = S.tviews;
    const float *s_cold = S.cold, *s_tcold = S.tcold;
    const DevScene& sc = s_sc;

    const int W = P.cam.width, rows = P.n_rows;
    const int sr = P.stripe_rows, sp = P.stripe_period;
    constexpr int TH = kTileH * K;
    const int tiles_x = (W + kTileW - 1) / kTileW, tiles_y = (rows + TH - 1) / TH;
    const unsigned int n_tiles = (unsigned int)(tiles_x * tiles_y);
    const int lx = lane & (kTileW - 1), ly = lane >> 3;

    unsigned int tile = 0;
    if (lane == 0) tile = atomicAdd(P.tile_counter, 1u);
    tile = __shfl_sync(kFullMask, tile, 0);
    while (tile < n_tiles) {
        unsigned int next = 0;
        if (lane == 0) next = atomicAdd(P.tile_counter, 1u);  // prefetch the next tile id behind the work
        if (P.tile_mask && !P.tile_mask[tile]) { tile = __shfl_sync(kFullMask, next, 0); continue; }  // second pass of rte_render32r.cu: flagged tiles only
        const int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile % (unsigned)tiles_x);
        int px[K], py[K];
        bool inb[K];
        int64_t out_idx[K];
        uint32_t n_in = 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int band_y = ty * TH + k * kTileH + ly;
            px[k] = tx * kTileW + lx;
            py[k] = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
            inb[k] = px[k] < W && band_y < rows;
            out_idx[k] = (int64_t)band_y * W + px[k];
            n_in += inb[k] ? 1u : 0u;
        }
        PixelCounters cnt = {0, 0, 0, 0, 0, 0, 0, 0};
        d3 c[K];
        {   // GeneratePixelAt, RE/Scene.h:283-304: accumulatedColor += sample; accumulatedColor /= samples
            const int aa = P.cam.aa;
            d3 sum[K];
#pragma unroll
            for (int k = 0; k < K; ++k) sum[k] = mk(0.0, 0.0, 0.0);
#pragma unroll 1
            for (int s = 0; s < aa; ++s) {
                trace_pixels<K, kReflect>(P, sc, s_views, s_tviews, s_cold, s_tcold, s_sph, s_tri, px, py, inb, out_idx, s, cnt, c);
                if (aa > 1) {
#pragma unroll
                    for (int k = 0; k < K; ++k) sum[k] = sum[k] + c[k];
                }
            }
            if (aa > 1) {  // (one sample: (0 + c) / 1 == c)
#pragma unroll
                for (int k = 0; k < K; ++k) c[k] = sum[k] / (double)aa;
            }
            n_in *= (uint32_t)aa;  // primary rays
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (!inb[k]) continue;
            const int64_t oi = out_idx[k];
            if (P.out.hdr) reinterpret_cast<float4*>(P.out.hdr)[oi] = make_float4((float)c[k].x, (float)c[k].y, (float)c[k].z, 0.f);
            if (P.out.hdr64) { P.out.hdr64[3 * oi] = c[k].x; P.out.hdr64[3 * oi + 1] = c[k].y; P.out.hdr64[3 * oi + 2] = c[k].z; }
            if (P.tile_mask && P.out.rgb8) {  // (the full-frame exact path leaves RGB8 to tonemap64_kernel; a few flagged tiles write their own bytes)
                const uchar3 u = tm64::to_color(tm64::apply(P.tonemap_op, c[k], P.max_white, P.jodie_a));
                uint8_t* d = P.out.rgb8 + 3 * oi;
                d[0] = u.x; d[1] = u.y; d[2] = u.z;
            }
        }
        // per-warp statistics: one REDUX per counter, lane 0 accumulates in shared memory
        // (second pass of rte_render32r.cu: fp64_confirms already holds the pixel samples of the flagged tiles)
        const uint32_t v[kStatCount] = {n_in, cnt.shadow, cnt.secondary, cnt.sph, cnt.pl, cnt.tri, cnt.marches, P.tile_mask ? 0u : cnt.confirms, cnt.sph_exec};
#pragma unroll
        for (int k = 0; k < kStatCount; ++k) {
            const uint32_t s = __reduce_add_sync(kFullMask, v[k]);
            if (lane == 0) s_stats[warp][k] += s;
        }
        tile = __shfl_sync(kFullMask, next, 0);
    }
    __syncwarp();
    if (lane < kStatCount && s_stats[warp][lane]) atomicAdd(&P.stats[lane], s_stats[warp][lane]);
}

// ------------------------------------------------------------------------------ the exact path on the direction grids
// Diffuse scenes (no specular / transparent material: no secondary rays) of ANY size: every primary and shadow ray looks up one
// cell of its origin's direction grid (rte_grid.cuh) and runs the reference's exact FP64 tests, in the reference's order with its
// tie rule, over that cell's spheres and triangles (rte_exact_grid.cuh) — O(cell) instead of O(primitives) per ray, nothing staged
// in shared memory, so there is no capacity limit (the view-based kernels above need (1 + lights) x primitives x 16 B per CTA).
// One pixel per lane; all outputs of rte_frame except rgb8 (the FP64 operator pass follows, like behind the other exact kernels).
__global__ void __launch_bounds__(kRenderThreads, 2) render_exact_grid_kernel(const __grid_constant__ RenderParams P) {
    __shared__ DevScene s_sc;
    __shared__ unsigned long long s_stats[kRenderThreads / 32][kStatCount];
    if (threadIdx.x == 0) s_sc = P.sc;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < kStatCount) s_stats[warp][lane] = 0ull;
    __syncthreads();
    const DevScene& sc = s_sc;
    const int W = P.cam.width, rows = P.n_rows, sr = P.stripe_rows, sp = P.stripe_period, aa = P.cam.aa, nl = sc.nl;
    const int tiles_x = (W + kTileW - 1) / kTileW, tiles_y = (rows + kTileH - 1) / kTileH;
    const unsigned int n_tiles = (unsigned int)(tiles_x * tiles_y);
    const int lx = lane & (kTileW - 1), ly = lane >> 3;
    unsigned int tile = 0;
    if (lane == 0) tile = atomicAdd(P.tile_counter, 1u);
    tile = __shfl_sync(kFullMask, tile, 0);
    while (tile < n_tiles) {
        unsigned int next = 0;
        if (lane == 0) next = atomicAdd(P.tile_counter, 1u);
        const int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile % (unsigned)tiles_x);
        const int band_y = ty * kTileH + ly, px = tx * kTileW + lx;
        const int py = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
        const bool inb = px < W && band_y < rows;
        const int64_t oi = (int64_t)band_y * W + px;
        PixelCounters cnt = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t n_prim = 0;
        if (inb) {
            d3 v = mk(0.0, 0.0, 0.0);
#pragma unroll 1
            for (int s = 0; s < aa; ++s) {  // GeneratePixelAt, RE/Scene.h:283-304
                ExactHit H;
                exact_primary(P, sc, px, py, s, H);
                cnt.sph += (uint32_t)sc.ns; cnt.pl += (uint32_t)sc.np; cnt.tri += (uint32_t)sc.nt;
                d3 diffuse = mk(0.0, 0.0, 0.0);
                uint32_t lmask = 0u;
#pragma unroll 1
                for (int j = 0; j < nl; ++j) {  // directLightning, RE/Scene.h:86-127
                    uint32_t bits;
                    const d3 c = exact_light(P, sc, H, j, bits, cnt);
                    if (bits & 1u) diffuse = diffuse + c;
                    if (j < 16) lmask |= (((bits >> 16) & 1u) << (16 + j)) | ((bits & 1u) << j);
                }
                const d3 e = exact_finish(sc, H, diffuse);
                if (aa > 1) v = v + e; else v = e;
                if (s == 0) {  // AOVs of the primary hit of the unjittered sample
                    if (P.out.prim_id) P.out.prim_id[oi] = H.prim;
                    if (P.out.depth) P.out.depth[oi] = H.hit ? (float)H.t : INFINITY;
                    if (P.out.depth64) P.out.depth64[oi] = H.hit ? H.t : (double)INFINITY;
                    if (P.out.normal) reinterpret_cast<float4*>(P.out.normal)[oi] = make_float4((float)H.ngeo.x, (float)H.ngeo.y, (float)H.ngeo.z, 0.f);
                    if (P.out.albedo) {
                        const d3 a = H.hit ? ld3(sc.materials[H.mat].color) : H.sky;
                        reinterpret_cast<float4*>(P.out.albedo)[oi] = make_float4((float)a.x, (float)a.y, (float)a.z, 0.f);
                    }
                    if (P.out.light_mask) P.out.light_mask[oi] = lmask;
                }
                n_prim += 1u;
            }
            if (aa > 1) v = v / (double)aa;
            if (P.out.hdr) __stcs(reinterpret_cast<float4*>(P.out.hdr) + oi, make_float4((float)v.x, (float)v.y, (float)v.z, 0.f));
            if (P.out.hdr64) { __stcs(P.out.hdr64 + 3 * oi, v.x); __stcs(P.out.hdr64 + 3 * oi + 1, v.y); __stcs(P.out.hdr64 + 3 * oi + 2, v.z); }
        }
        const uint32_t st[kStatCount] = {n_prim, cnt.shadow, 0u, cnt.sph, cnt.pl, cnt.tri, cnt.marches, 0u, 0u};
#pragma unroll
        for (int k = 0; k < kStatCount; ++k) {
            const uint32_t t = __reduce_add_sync(kFullMask, st[k]);
            if (lane == 0) s_stats[warp][k] += t;
        }
        tile = __shfl_sync(kFullMask, next, 0);
    }
    __syncwarp();
    if (lane < kStatCount && s_stats[warp][lane]) atomicAdd(&P.stats[lane], s_stats[warp][lane]);
}

// ------------------------------------------------------------------------------ view providers
// Resident: all (1+nl) views sit in shared memory for the lifetime of the CTA.
struct ViewsResident {
    const ulonglong2* base;
    int stride;  // ulonglong2 elements per view
    __device__ __forceinline__ const ulonglong2* acquire(int v) const { return base + (size_t)v * stride; }
    __device__ __forceinline__ void release() const {}
    static constexpr bool kStream = false;
};
// Streamed: the views of a large scene (C5: 9 x 64 KB) pass through two shared-memory buffers.  Every warp of the CTA
// walks the same cyclic view sequence 0,1,..,nl,0,1,.. (one tile batch per cycle); while the CTA scans the view in one
// buffer, TMA bulk copies fill the other one with the next view of the sequence (mbarrier complete_tx), and
// release() = __syncthreads() + refill of the buffer just left.  acquire()/release() are called by ALL threads.
struct ViewsStream {
    unsigned char* base;  // buffer b = base + b * bytes (computed, not looked up: the scans must compile to LDS, not generic LD)
    uint64_t* bar;  // two mbarriers
    const unsigned char* gsrc;
    uint32_t bytes;  // of one view
    int nv;
    uint32_t step;  // views consumed so far by this CTA
    static constexpr bool kStream = true;
    __device__ __forceinline__ void fill(uint32_t s) {  // one elected thread: view (s mod nv) -> buffer (s & 1)
        uint64_t* b = bar + (s & 1u);
        mbar_expect_tx(b, bytes);
        tma_chunks(base + (s & 1u) * bytes, gsrc + (size_t)(s % (uint32_t)nv) * bytes, bytes, b);
    }
    __device__ __forceinline__ void start() {
        if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar + 1, 1); }
        __syncthreads();
        if (threadIdx.x == 0) { fill(0); fill(1); }
        step = 0;
    }
    __device__ __forceinline__ const ulonglong2* acquire(int) {
        mbar_wait(bar + (step & 1u), (step >> 1) & 1u);
        return reinterpret_cast<const ulonglong2*>(base + (step & 1u) * bytes);
    }
    __device__ __forceinline__ void release() {
        __syncthreads();  // every warp has left the buffer
        if (threadIdx.x == 0) fill(step + 2);
        ++step;
    }
    __device__ __forceinline__ void drain() {  // two refills are always in flight: wait for them before the CTA exits
        mbar_wait(bar + (step & 1u), (step >> 1) & 1u);
        mbar_wait(bar + ((step + 1) & 1u), ((step + 1) >> 1) & 1u);
    }
};

// ------------------------------------------------------------------------------ the fast kernel
// Production path of diffuse sphere/plane scenes (C2/C3/C5 families): views resident in shared memory, no
// material with specular > 0, no triangles, no validation switch, no AOV output.  Same arithmetic as
// trace_pixels (bit-identical results, tests/test_gpu_render.py::test_fast_kernel_*), but laid out for the
// instruction caches: everything that runs once per pixel sits in `#pragma unroll 1` loops over the K pixels of
// the thread (one copy of the FP64 code, per-pixel state in local memory), only the two sphere scans are
// unrolled over K.  The general kernel's per-pixel code is ~4x larger and stalls on instruction fetch.
// kTris: the scene has triangles and their views are resident (s_tviews): a third scan after the planes.
// kPhase: 0 = the whole of TraceRay; 1 = primary rays only (nearest hit of every pixel -> P.hits[pix]); 2 = everything after the
// nearest hit (reads P.hits[pix]).  Two launches of half the code each: the kernel waits on instruction fetch.
// kSt: hit point and normal of the pixel (six doubles that live across the whole light loop, and across every out-of-line call
// in it) sit in shared memory, st[c * blockDim.x + threadIdx.x], instead of registers / local memory (K = 1 only).
template <int K, bool kTris, class Views, int kPhase = 0, bool kSt = false>
__device__ __forceinline__ void trace_tile_fast(const RenderParams& P, const DevScene& sc, Views& views, const ulonglong2* s_tviews,
                                                const float* s_cold, const float* s_tcold, const float4* s_sph, const float4* s_tri, const CullScratch* cull,
                                                const CtaCull* cc, const int (&pxs)[K], const int (&pys)[K], const bool (&inb)[K], const int64_t (&pix)[K], int sample,
                                                PixelCounters& cnt, d3 (&result)[K], double* __restrict__ st = nullptr) {
    const d3 cam_o = ld3(P.cam.pos);
    const double bias = P.bias;
    const int nl = sc.nl, np = sc.np, ns = sc.ns, nt = kTris ? sc.nt : 0;
    const int tview_stride = sc.nt_pad32;
    d3 dir[K];
    bool alive[K], found[K];
    Hit h[K];
#pragma unroll 1
    for (int k = 0; k < K; ++k) {
        dir[k] = camera_dir(P.cam, pxs[k], pys[k], sample);
        alive[k] = inb[k];
        found[k] = false;
        h[k].t = 0.0; h[k].kind = RTE_HIT_NONE; h[k].idx = 0;
    }
    if (kPhase != 2) {
        const ulonglong2* v0 = views.acquire(0);
        if (ns > 0) closest_spheres_view<K>(sc, v0, s_cold, cam_o, dir, alive, false, h, found, cnt.confirms, cull, &cnt.sph_exec, cc);
        views.release();
    }
    if (kPhase == 1) {  // planes after the spheres, strict '<' (RE/Scene.h:218-257); the hit goes to the shading launch
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            if (!alive[k]) continue;
            const d3 d = dir[k];
            Hit hk = h[k];
            bool fk = found[k];
#pragma unroll 1
            for (int i = 0; i < np; ++i) {
                double t;
                if (plane_exact(sc.planes[i], cam_o, d, t) && (!fk || t < hk.t)) {
                    fk = true; hk.t = t; hk.kind = RTE_HIT_PLANE; hk.idx = (uint32_t)i;
                }
            }
            const unsigned long long packed = fk ? (((unsigned long long)hk.kind << 32) | hk.idx) : 0ull;  // RTE_HIT_NONE = 0
            P.hits[pix[k]] = make_double2(hk.t, __longlong_as_double((long long)packed));
        }
        return;
    }
    if (kPhase == 2) {
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            if (!alive[k]) continue;
            const double2 v = P.hits[pix[k]];
            const unsigned long long packed = (unsigned long long)__double_as_longlong(v.y);
            h[k].t = v.x; h[k].kind = (uint32_t)(packed >> 32); h[k].idx = (uint32_t)packed;
            found[k] = h[k].kind != RTE_HIT_NONE;
        }
    }

    d3 Pn[K], normal[K], diffuse[K];
    int mat[K], self[K];
    bool hitv[K];
    if (kTris) {  // planes, then the triangles through their camera view, then (below) the hit records
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            if (!alive[k]) continue;
            const d3 d = dir[k];
            Hit hk = h[k];
            bool fk = found[k];
#pragma unroll 1
            for (int i = 0; i < np; ++i) {
                double t;
                if (plane_exact(sc.planes[i], cam_o, d, t) && (!fk || t < hk.t)) {
                    fk = true; hk.t = t; hk.kind = RTE_HIT_PLANE; hk.idx = (uint32_t)i;
                }
            }
            h[k] = hk; found[k] = fk;
        }
        closest_spheres_view<K, true>(sc, s_tviews, s_tcold, cam_o, dir, alive, false, h, found, cnt.confirms, cull);
    }
#pragma unroll 1
    for (int k = 0; k < K; ++k) {
        const d3 d = dir[k];
        Hit hk = h[k];
        bool fk = found[k];
        if (alive[k]) {
            cnt.sph += (uint32_t)ns; cnt.pl += (uint32_t)np; cnt.tri += (uint32_t)nt;
#pragma unroll 1
            for (int i = 0; i < ((kTris || kPhase == 2) ? 0 : np); ++i) {  // planes after the spheres, strict '<' (RE/Scene.h:218-257)
                double t;
                if (plane_exact(sc.planes[i], cam_o, d, t) && (!fk || t < hk.t)) {
                    fk = true; hk.t = t; hk.kind = RTE_HIT_PLANE; hk.idx = (uint32_t)i;
                }
            }
        }
        const bool hv = alive[k] && fk;
        hitv[k] = hv;
        result[k] = mk(0.0, 0.0, 0.0);
        if (alive[k] && !hv) result[k] = result[k] + background(d) * 1.0;  // miss: sky (RE/Scene.h:137-141)
        d3 P_ = mk(0, 0, 0), ng = mk(0, 0, 0);
        int m = sc.nm, sf = -1;
        if (hv) {
            P_ = cam_o + d * hk.t;
            m = material_of(sc, hk);
            if (hk.kind == RTE_HIT_SPHERE) {
                const double4 s = sc.sph64[hk.idx];
                ng = normalize(P_ - mk(s.x, s.y, s.z));
            } else if (!kTris || hk.kind == RTE_HIT_PLANE) {
                ng = ld3(sc.planes[hk.idx].n);
            } else {
                ng = ld3(sc.tri64[hk.idx].n);
            }
        }
        const d3 incoming = normalize(d);
        const bool frontFace = dot(ng, incoming) < 0.0;
        const d3 N = frontFace ? ng : -ng;
        if (hv && hk.kind == RTE_HIT_SPHERE && frontFace) sf = (int)hk.idx;
        if (kSt) {
            const d3 nn = normalize(N);
            const int T = (int)blockDim.x, tid = (int)threadIdx.x;
            st[tid] = P_.x; st[T + tid] = P_.y; st[2 * T + tid] = P_.z;
            st[3 * T + tid] = nn.x; st[4 * T + tid] = nn.y; st[5 * T + tid] = nn.z;
        } else {
            Pn[k] = P_; normal[k] = normalize(N);
        }
        mat[k] = m; self[k] = sf;
        diffuse[k] = mk(0.0, 0.0, 0.0);
    }
    // directLightning, RE/Scene.h:79-129 (diffuse term; specAcc stays 0 without specular materials)
#pragma unroll 1
    for (int j = 0; j < nl; ++j) {
        const d3 Lp = ld3(sc.lights[j].pos);
        ShadowRay sr[K];
        bool any_cast = false;
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            d3 Pk, Nk;
            if (kSt) {
                const int T = (int)blockDim.x, tid = (int)threadIdx.x;
                Pk = mk(st[tid], st[T + tid], st[2 * T + tid]);
                Nk = mk(st[3 * T + tid], st[4 * T + tid], st[5 * T + tid]);
            } else {
                Pk = Pn[k]; Nk = normal[k];
            }
            const d3 vecToLight = Lp - Pk;
            const d4 vl = unit_len(vecToLight);
            const double dist = vl.w;
            bool cast = hitv[k] && dist > 0.0;
            d3 L = mk(0, 0, 1);
            double ndl = 0.0;
            if (cast) {
                L = mk(vl.x, vl.y, vl.z);  // vecToLight / dist
                ndl = std_max(0.0, dot(Nk, L));
                cast = ndl > 0.0 && dist > bias;
            }
            ShadowRay s;
            s.o = Pk + Nk * bias;
            s.d = L; s.dist = dist; s.maxDist = dist - bias; s.ndl = ndl; s.self = self[k]; s.cast = cast;
            sr[k] = s;
            any_cast = any_cast || cast;
        }
        const bool scan = __any_sync(kFullMask, any_cast);
        if (!Views::kStream && !scan) continue;
        OccState st[K];
        RayLists<K> lists;
        {
            const ulonglong2* vj = views.acquire(1 + j);  // streamed views: every warp of the CTA passes here, scanning or not
            // (with a CTA-level cull every warp of the CTA takes part, casting or not)
            if (scan || cc) occluded_spheres_view<K>(sc, vj, s_cold + (size_t)(1 + j) * sc.ns_pad * 4, Lp, sr, bias, false, st, lists, cnt.confirms, cnt.sph_exec, cull, cc);
            views.release();
        }
        if (!scan) continue;
        bool all_done = true;
#pragma unroll
        for (int k = 0; k < K; ++k) all_done = all_done && st[k].done;
        const bool skip_planes = __all_sync(kFullMask, all_done);
        if (kTris && !skip_planes) {  // planes, then the triangles through the light's view; T is computed below
            bool all_done2 = true;
#pragma unroll 1
            for (int k = 0; k < K; ++k) {
                OccState s = st[k];
                if (s.res.near && s.res.occluded) s.done = true;  // the march decides
                if (!s.done) {
                    const ShadowRay r = sr[k];
                    uint32_t firstp = (uint32_t)np;
#pragma unroll 1
                    for (int i = 0; i < np; ++i) {
                        double t;
                        if (plane_exact(sc.planes[i], r.o, r.d, t)) {
                            s.on_t(t, firstp, (uint32_t)i);
                            if (s.done) break;
                        }
                    }
                    s.res.pl_tests = firstp;
                }
                st[k] = s;
                all_done2 = all_done2 && s.done;
            }
            if (!__all_sync(kFullMask, all_done2))
                occluded_spheres_view<K, true>(sc, s_tviews + (size_t)(1 + j) * tview_stride, s_tcold + (size_t)(1 + j) * sc.nt_pad32 * 4, Lp, sr, bias,
                                               false, st, lists, cnt.confirms, cnt.sph_exec, cull);
        }
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            OccState s = st[k];
            const ShadowRay r = sr[k];
            if (!kTris && !skip_planes && !s.done && !s.res.occluded) {  // planes, any-hit in canonical order (a `near` ray with an occluder: the march decides)
                uint32_t firstp = (uint32_t)np;
#pragma unroll 1
                for (int i = 0; i < np; ++i) {
                    double t;
                    if (plane_exact(sc.planes[i], r.o, r.d, t)) {
                        s.on_t(t, firstp, (uint32_t)i);
                        if (s.done) break;
                    }
                }
                s.res.pl_tests = firstp;
            }
            cnt.sph += s.res.sph_tests; cnt.pl += s.res.pl_tests; cnt.tri += s.res.tri_tests;
            if (!r.cast) continue;
            double T;
            if (s.res.near) {
                const int ln = lists.n[k];
                if (!kTris && ln <= kListCap) T = march_list(sc, lists.idx[k], ln, r.o, r.d, r.maxDist, bias);
                else T = march_exact(sc, s_sph, s_tri, r.o, r.d, r.maxDist, bias, false, cnt.confirms);
                cnt.marches += 1;
            }
            else T = s.res.occluded ? 0.0 : 1.0;
            cnt.shadow += 1;
            if (T > bias) {
                const d3 emitted = ld3(sc.lights[j].E);
                const d3 contribution = (emitted * div64(1.0, r.dist * r.dist)) * r.ndl;
                diffuse[k] = diffuse[k] + contribution * T;
            }
        }
    }
#pragma unroll 1
    for (int k = 0; k < K; ++k) {
        if (!hitv[k]) continue;
        const rte_material& M = sc.materials[mat[k]];
        const d3 local = ld3(M.color) * diffuse[k] + mk(0.0, 0.0, 0.0) * M.specular;
        result[k] = result[k] + (local * (1.0 - 0.0)) * 1.0;  // finalLight += localLight * (1 - transparency), w = 1
    }
}

// Block shape of the fast kernel with resident views, measured on B200 (C2 at 4K, one pixel per thread): 320 x 2 CTAs 4.59 ms,
// 512 x 1 4.11, 768 x 1 4.17, 1024 x 1 3.82.  The kernel waits on instruction fetch and local-memory latency, so 32 resident
// warps at 64 registers beat 16-20 warps at 96-128 registers despite the extra spills, and one CTA stages the views once per SM.
#ifndef RTE_FAST_THREADS
#define RTE_FAST_THREADS 1024
#endif
#ifndef RTE_FAST_MINB
#define RTE_FAST_MINB 1
#endif
#ifndef RTE_STATS_SEL
#define RTE_STATS_SEL 1
#endif
constexpr int kFastThreads = RTE_FAST_THREADS;  // block size of the fast kernel when kFastMinBlocks CTAs fit one SM
constexpr int kFastMinBlocks = RTE_FAST_MINB;
#ifndef RTE_TRIS_THREADS
#define RTE_TRIS_THREADS 768
#endif
constexpr int kTrisThreads = RTE_TRIS_THREADS;  // scenes with triangle views (C4 at 1080p: 1024 x 64 registers 0.96 ms, 768 x 80 0.84 ms)
#ifndef RTE_BIG_THREADS
#define RTE_BIG_THREADS 512
#endif
constexpr int kBigThreads = RTE_BIG_THREADS;  // block size of the streamed-view variant (one pixel per thread, CTA-level cull, C5 at 4K:
                                              // 384 x 128 registers 10.4 ms, 512 x 128 9.6, 768 x 80 10.5, 1024 x 64 9.9)

// kBig = false: kFastThreads x kFastMinBlocks CTAs per SM; kBig = true: one 512-thread CTA per SM (views of large scenes)
// kStream = true (implies kBig): the views pass through two shared-memory buffers (ViewsStream), tile batches are CTA-uniform
// kPhase (resident sphere views, one sample per pixel): 0 = one launch; 1 / 2 = the primary-ray launch and the shading launch
template <int K, bool kBig, bool kStream, bool kTris = false, int kPhase = 0, bool kSt = false>
__global__ void __launch_bounds__(kBig ? kBigThreads : (kTris ? kTrisThreads : kFastThreads), kBig ? 1 : kFastMinBlocks)
render_fast_kernel(const __grid_constant__ RenderParams P) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ DevScene s_sc;
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ __align__(8) uint64_t s_vbar[2];
    __shared__ unsigned long long s_stats[(kBig ? kBigThreads : (kTris ? kTrisThreads : kFastThreads)) / 32][kStatCount];

#if RTE_SMEM_CONSTS
    // the planes and lights every ray reads (a few hundred bytes) live in shared memory: the L1 is busy with local memory
    __shared__ DevPlane s_planes[kSmemPlanes];
    __shared__ DevLight s_lights[kSmemLights];
    if (threadIdx.x == 0) {
        DevScene t = P.sc;
        if (t.np <= kSmemPlanes) { for (int i = 0; i < t.np; ++i) s_planes[i] = P.sc.planes[i]; t.planes = s_planes; }
        if (t.nl <= kSmemLights) { for (int i = 0; i < t.nl; ++i) s_lights[i] = P.sc.lights[i]; t.lights = s_lights; }
        s_sc = t;
    }
#else
    if (threadIdx.x == 0) s_sc = P.sc;
#endif
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < kStatCount) s_stats[warp][lane] = 0ull;
    StagedScene S;
    stage_scene(P.sc, P.plan, dyn_smem, &s_bar, S);
    const float4 *s_sph = S.sph, *s_tri = S.tri;
    const ulonglong2 *s_views = S.views, *s_tviews = S.tviews;
    const float *s_cold = S.cold, *s_tcold = S.tcold;
    const DevScene& sc = s_sc;
    ViewsResident views_res{s_views, sc.ns_pad};
    CullScratch cull_scratch{nullptr, nullptr};
    if (P.cull) {
        unsigned char* base = dyn_smem + ((P.plan.total + 15) & ~(size_t)15) + (size_t)warp * kCullBytesPerWarp;
        cull_scratch.mini = reinterpret_cast<ulonglong2*>(base);
        cull_scratch.map = reinterpret_cast<uint16_t*>(base + kCullCap * 16);
    }
    const CullScratch* cull = P.cull ? &cull_scratch : nullptr;
    CtaCull cta_scratch{nullptr, nullptr, nullptr};
    // kBatch: the CTA works on batches of adjacent tiles and culls a view once per CTA (streamed views; the primary-ray launch)
    constexpr bool kBatch = kStream || (kPhase == 1 && RTE_P1_BATCH != 0);
    if (kBatch && P.cta_cull) {
        unsigned char* base = dyn_smem + ((P.plan.total + 15) & ~(size_t)15) + (size_t)(blockDim.x >> 5) * kCullBytesPerWarp;
        cta_scratch.mini = reinterpret_cast<ulonglong2*>(base);
        cta_scratch.map = reinterpret_cast<uint16_t*>(base + kCtaCap * 16);
        cta_scratch.part = reinterpret_cast<int*>(base + kCtaCap * 18);
    }
    const CtaCull* cc = (kBatch && P.cta_cull) ? &cta_scratch : nullptr;
    const bool batched = kStream || (kBatch && P.cta_cull);  // (the primary-ray launch without CTA scratch keeps per-warp tiles)
    // kSt: six doubles per thread behind the per-warp cull scratch (the launch grants kernel variants with kSt only when they fit)
    double* px_state = kSt ? reinterpret_cast<double*>(dyn_smem + ((P.plan.total + 15) & ~(size_t)15) + (size_t)(blockDim.x >> 5) * kCullBytesPerWarp) : nullptr;
    ViewsStream views_str;
    if (kStream) {
        views_str.base = dyn_smem;
        views_str.bar = s_vbar; views_str.gsrc = reinterpret_cast<const unsigned char*>(P.sc.views);
        views_str.bytes = P.plan.one_view_bytes; views_str.nv = 1 + P.sc.nl;
        if (kPhase == 2) { views_str.gsrc += P.plan.one_view_bytes; views_str.nv = P.sc.nl; }  // the shading launch walks the lights' views only
        views_str.start();
    }

    const int W = P.cam.width, rows = P.n_rows;
    const int sr = P.stripe_rows, sp = P.stripe_period;
    constexpr int TH = kTileH * K;
    const int tiles_x = (W + kTileW - 1) / kTileW, tiles_y = (rows + TH - 1) / TH;
    const unsigned int n_tiles = (unsigned int)(tiles_x * tiles_y);
    const int lx = lane & (kTileW - 1), ly = lane >> 3;

    // Resident views: every warp pulls its next tile from the global counter.  Streamed views: the CTA pulls a BATCH (one tile
    // per warp; the warps walk the views together) whose tiles lie side by side, 4 x 4 (flatter when the rows are striped: a batch
    // stays inside one stripe), so that the CTA-level cull sees a narrow cone.
    __shared__ unsigned int s_batch;
    const int nw = (int)(blockDim.x >> 5);
    int bh = nw / RTE_BATCH_W;                                  // rows of tiles per batch ...
    if (sr > 0) bh = min(bh, max(1, sr / TH));                   // ... within one stripe of rows (stripe_rows is a multiple of 4)
    while (bh > 1 && (nw % bh != 0 || (sr > 0 && (sr / TH) % bh != 0))) --bh;
    const int bw = nw / bh;
    const int batches_x = (tiles_x + bw - 1) / bw, batches_y = (tiles_y + bh - 1) / bh;
    const unsigned int n_batches = (unsigned int)(batches_x * batches_y);
    unsigned int tile = 0;
    if (!batched) {
        if (lane == 0) tile = atomicAdd(P.tile_counter, 1u);
        tile = __shfl_sync(kFullMask, tile, 0);
    }
    while (true) {
        bool have;
        int tx, ty;
        unsigned int next = 0;
        if (batched) {
            if (threadIdx.x == 0) s_batch = atomicAdd(P.tile_counter, 1u);
            __syncthreads();
            const unsigned int b = s_batch;
            __syncthreads();
            if (b >= n_batches) break;
            tx = (int)(b % (unsigned)batches_x) * bw + warp % bw;
            ty = (int)(b / (unsigned)batches_x) * bh + warp / bw;
            have = tx < tiles_x && ty < tiles_y;  // idle warps still walk the views
        } else {
            have = tile < n_tiles;
            if (!have) break;
            if (lane == 0) next = atomicAdd(P.tile_counter, 1u);
            ty = (int)(tile / (unsigned)tiles_x); tx = (int)(tile % (unsigned)tiles_x);
        }
        int px[K], py[K];
        bool inb[K];
        int64_t pix[K];  // band-local pixel index
        uint32_t n_in = 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int band_y = ty * TH + k * kTileH + ly;
            px[k] = tx * kTileW + lx;
            py[k] = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
            inb[k] = have && px[k] < W && band_y < rows;
            pix[k] = (int64_t)band_y * W + px[k];
            n_in += (inb[k] && kPhase != 2) ? 1u : 0u;
        }
        PixelCounters cnt = {0, 0, 0, 0, 0, 0, 0, 0};
        d3 c[K];
        {   // GeneratePixelAt, RE/Scene.h:283-304: accumulatedColor += sample; accumulatedColor /= samples
            const int aa = P.cam.aa;
            d3 sum[K];
#pragma unroll
            for (int k = 0; k < K; ++k) sum[k] = mk(0.0, 0.0, 0.0);
#pragma unroll 1
            for (int s = 0; s < aa; ++s) {
                if (kStream) trace_tile_fast<K, false, ViewsStream, kPhase>(P, sc, views_str, s_tviews, s_cold, s_tcold, s_sph, s_tri, cull, cc, px, py, inb, pix, s, cnt, c);
                else trace_tile_fast<K, kTris, ViewsResident, kPhase, kSt>(P, sc, views_res, s_tviews, s_cold, s_tcold, s_sph, s_tri, cull, kBatch ? cc : nullptr, px, py, inb, pix, s, cnt, c, px_state);
                if (aa > 1) {
#pragma unroll 1
                    for (int k = 0; k < K; ++k) sum[k] = sum[k] + c[k];
                }
            }
            if (aa > 1) {  // (one sample: (0 + c) / 1 == c)
#pragma unroll 1
                for (int k = 0; k < K; ++k) c[k] = sum[k] / (double)aa;
            }
            n_in *= (uint32_t)aa;  // primary rays
        }
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            if (!inb[k] || kPhase == 1) continue;
            const int64_t oi = pix[k];
            const d3 v = c[k];
            // streaming stores (st.global.cs): the framebuffers are write-once and must not evict the threads' local memory from L2
            if (P.out.hdr) __stcs(reinterpret_cast<float4*>(P.out.hdr) + oi, make_float4((float)v.x, (float)v.y, (float)v.z, 0.f));
            if (P.out.hdr64) { __stcs(P.out.hdr64 + 3 * oi, v.x); __stcs(P.out.hdr64 + 3 * oi + 1, v.y); __stcs(P.out.hdr64 + 3 * oi + 2, v.z); }
        }
        const uint32_t v[kStatCount] = {n_in, cnt.shadow, cnt.secondary, cnt.sph, cnt.pl, cnt.tri, cnt.marches, cnt.confirms, cnt.sph_exec};
#if RTE_STATS_SEL
        uint32_t mine = 0;  // lane k keeps the warp's sum of counter k: one shared-memory update per tile instead of kStatCount
#pragma unroll
        for (int k = 0; k < kStatCount; ++k) {
            const uint32_t s = __reduce_add_sync(kFullMask, v[k]);
            if (lane == k) mine = s;
        }
        if (lane < kStatCount) s_stats[warp][lane] += mine;
#else
#pragma unroll
        for (int k = 0; k < kStatCount; ++k) {
            const uint32_t s = __reduce_add_sync(kFullMask, v[k]);
            if (lane == 0) s_stats[warp][k] += s;
        }
#endif
        tile = __shfl_sync(kFullMask, next, 0);
    }
    if (kStream) views_str.drain();
    __syncwarp();
    if (lane < kStatCount && s_stats[warp][lane]) atomicAdd(&P.stats[lane], s_stats[warp][lane]);
}

static cudaError_t ensure_smem(const void* fn, size_t bytes) {
    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

cudaError_t launch_build_views(const DevScene& sc, const DevCamera& cam, double bias, cudaStream_t stream) {
    if (sc.views && sc.ns > 0) {
        const int total = (1 + sc.nl) * sc.ns_pad;
        build_views_kernel<false><<<(total + 255) / 256, 256, 0, stream>>>(sc, cam, bias, const_cast<float*>(sc.views), const_cast<float*>(sc.views_cold));
    }
    if (sc.tri_views && sc.nt > 0) {
        const int total = (1 + sc.nl) * sc.nt_pad32;
        build_views_kernel<true><<<(total + 255) / 256, 256, 0, stream>>>(sc, cam, bias, const_cast<float*>(sc.tri_views), const_cast<float*>(sc.tri_views_cold));
    }
    return cudaGetLastError();
}

// Launch configuration of one (kernel pair, staged bytes, flags) on one device.  The occupancy queries and attribute calls
// behind it take tens of microseconds of host time, during which an otherwise empty stream idles inside the caller's
// timed region — they run once per configuration, not once per frame.
struct LaunchCfg {
    const void *key_fn, *key_big;
    size_t key_smem, key_limit;
    int key_dev, key_threads, key_minb, key_bigt, key_flags;
    const void* fn;
    int threads, per_sm, cull, cta_cull;
    size_t smem;
};
static thread_local std::vector<LaunchCfg> t_launch_cfgs;

// `cull_ok`: the kernel can use per-warp cull scratch (the fast kernel); it is granted when it fits next to the staged arrays
// without lowering the number of CTAs per SM.
// `state_bytes`: shared memory behind the cull scratch that this kernel variant cannot do without (kSt); when it does not fit,
// the configuration is recorded as unusable (c.fn = NULL) and the caller launches the plain variant.
static cudaError_t resolve_launch(LaunchCfg& c, const void* fn, const void* fn_big, int small_threads, int small_min_blocks, const RenderParams& p,
                                  bool cull_ok, size_t smem_limit, int big_threads, size_t state_bytes, bool cta_ok) {
    size_t smem = p.plan.total;
    cudaError_t e = ensure_smem(fn, smem);
    if (e != cudaSuccess) return e;
    // several CTAs per SM when the staged arrays fit that often, else one CTA of 512 threads
    int threads = small_threads, per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < small_min_blocks) {
        fn = fn_big;
        threads = big_threads;
        if ((e = ensure_smem(fn, smem)) != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem);
        if (e != cudaSuccess) return e;
    }
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    c.cull = 0; c.cta_cull = 0;
    if (cull_ok && !p.no_cull) {
        const size_t with_cull = ((smem + 15) & ~(size_t)15) + (size_t)(threads / 32) * kCullBytesPerWarp;
        int per_sm2 = 0;
        if (with_cull <= smem_limit && ensure_smem(fn, with_cull) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm2, fn, threads, with_cull) == cudaSuccess && per_sm2 >= per_sm) {
            smem = with_cull;
            c.cull = 1;
            // streamed views: the CTA-level cull scratch behind the per-warp scratch, when it fits as well
            const size_t with_cta = with_cull + kCtaCullBytes;
            if ((p.plan.stream_views || cta_ok) && with_cta <= smem_limit && ensure_smem(fn, with_cta) == cudaSuccess &&
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm2, fn, threads, with_cta) == cudaSuccess && per_sm2 >= per_sm) {
                smem = with_cta;
                c.cta_cull = 1;
            }
        }
        if ((e = ensure_smem(fn, smem)) != cudaSuccess) return e;  // (a failed probe above may have left a larger request)
    }
    c.fn = fn; c.threads = threads; c.per_sm = per_sm; c.smem = smem;
    if (state_bytes) {
        int per_sm2 = 0;
        const size_t with_state = smem + state_bytes;
        if (!c.cull || with_state > smem_limit || ensure_smem(fn, with_state) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm2, fn, threads, with_state) != cudaSuccess || per_sm2 < per_sm) {
            cudaGetLastError();
            c.fn = nullptr;
        } else {
            c.smem = with_state;
        }
    }
    return cudaSuccess;
}

static cudaError_t launch_render_fn(const void* fn, const void* fn_big, int small_threads, int small_min_blocks, int K,
                                    const RenderParams& p0, int sm_count, cudaStream_t stream, bool cull_ok = false, size_t smem_limit = 0,
                                    int big_threads = kRenderMaxThreads, size_t state_bytes = 0, bool cta_ok = false) {
    RenderParams p = p0;
    int dev = 0;
    cudaGetDevice(&dev);
    const int flags = (cull_ok ? 1 : 0) | (p.no_cull ? 2 : 0) | (p.plan.stream_views ? 4 : 0) | (cta_ok ? 8 : 0);
    const LaunchCfg* cfg = nullptr;
    for (const LaunchCfg& c : t_launch_cfgs)
        if (c.key_fn == fn && c.key_big == fn_big && c.key_smem == p.plan.total && c.key_limit == smem_limit && c.key_dev == dev &&
            c.key_threads == small_threads && c.key_minb == small_min_blocks && c.key_bigt == big_threads && c.key_flags == flags) { cfg = &c; break; }
    if (!cfg) {
        LaunchCfg c{};
        c.key_fn = fn; c.key_big = fn_big; c.key_smem = p.plan.total; c.key_limit = smem_limit; c.key_dev = dev;
        c.key_threads = small_threads; c.key_minb = small_min_blocks; c.key_bigt = big_threads; c.key_flags = flags;
        const cudaError_t e = resolve_launch(c, fn, fn_big, small_threads, small_min_blocks, p, cull_ok, smem_limit, big_threads, state_bytes, cta_ok);
        if (e != cudaSuccess) return e;
        if (t_launch_cfgs.size() >= 64) t_launch_cfgs.clear();
        t_launch_cfgs.push_back(c);
        cfg = &t_launch_cfgs.back();
    }
    if (!cfg->fn) return cudaErrorInvalidConfiguration;  // (a kSt variant that does not fit: not an error of the device)
    p.cull = cfg->cull;
    p.cta_cull = cfg->cta_cull;
    const int threads = cfg->threads;
    const int W = p.cam.width, rows = p.n_rows;
    const long long n_tiles = (long long)((W + kTileW - 1) / kTileW) * ((rows + kTileH * K - 1) / (kTileH * K));
    long long blocks = (long long)sm_count * cfg->per_sm;  // persistent: a multiple of the SM count
    const long long needed = (n_tiles + threads / 32 - 1) / (threads / 32);
    if (blocks > needed) blocks = needed > 0 ? needed : 1;
    const cudaError_t e = ensure_smem(cfg->fn, cfg->smem);  // (per function and device, and other configurations of this kernel lower it)
    if (e != cudaSuccess) return e;
    void* args[] = {(void*)&p};
    return cudaLaunchKernel(cfg->fn, dim3((unsigned)blocks), dim3((unsigned)threads), args, cfg->smem, stream);
}

bool fast_path_eligible(const RenderParams& p) {
    const rte_frame& f = p.out;
    return (p.plan.use_views || p.plan.use_tri_views) && !p.sc.reflective && (p.sc.nt == 0 || p.plan.use_tri_views) && !p.exact_shadows &&
           !p.no_filter && !f.prim_id && !f.depth && !f.depth64 && !f.normal && !f.albedo && !f.light_mask;
}

// Two launches of half the code each pay on big frames of heavy scenes (C2 at 4K: 2.93 -> 2.68 ms); every launch of the
// persistent kernel has a floor of ~0.1 ms (ramp-up + the tail of 50 us tiles), which light scenes (C3: 0.85 -> 0.91 ms) and
// small row bands (C2: under ~1.3 Mpx) do not earn back.
bool two_pass_wanted(const RenderParams& p) {
    if (!fast_path_eligible(p) || p.force_general || p.sc.nt > 0 || p.cam.aa != 1 || kViewPixels != 1) return false;
    if (p.plan.stream_views && (kStreamPixels != 1 || p.sc.nl < 1)) return false;  // (streamed views: the primary launch keeps view 0 resident)
    if (p.pass_mode == 1) return false;
    if (p.pass_mode == 2) return true;
    // (streamed views: the primary launch keeps the camera's view resident and culls it per CTA batch: C5 at 4K 9.35 -> 9.19 ms;
    // with per-warp culls of the whole view it was 9.72 ms)
    const long long entries = (long long)(1 + p.sc.nl) * p.sc.ns_pad, pixels = (long long)p.n_rows * p.cam.width;
    return entries >= 2048 && pixels >= (5ll << 18);  // C2, 1/N of the 4K frame (tools/pass_probe.py): two launches win down to 1/6 (1.4 Mpx), lose at 1/8
}

// Does this exact-path call take the grid kernel?  Diffuse scenes when asked for (RTE_FLAG_EXACT_GRID), when the per-origin views
// do not fit one CTA's shared memory, or with more lights than the view kernels keep in shared memory.
bool exact_grid_wanted(const RenderParams& p) {
    if (p.sc.reflective || p.exact_shadows || p.no_filter || p.sc.ns + p.sc.nt == 0) return false;
    if (p.exact_grid) return true;
    if (p.sc.nl > kSmemLights) return true;
    if (p.sc.ns > 0 && !p.plan.use_views) return true;
    if (p.sc.nt > 0 && !p.plan.use_tri_views) return true;
    return false;
}

cudaError_t launch_render(const RenderParams& p0, int sm_count, cudaStream_t stream) {
    RenderParams p = p0;
    if (p.grid.cell_start && exact_grid_wanted(p)) {
        const long long n_tiles = (long long)((p.cam.width + kTileW - 1) / kTileW) * ((p.n_rows + kTileH - 1) / kTileH);
        long long blocks = (long long)sm_count * 2;
        const long long needed = (n_tiles + kRenderThreads / 32 - 1) / (kRenderThreads / 32);
        if (blocks > needed) blocks = needed > 0 ? needed : 1;
        void* args[] = {(void*)&p};
        return cudaLaunchKernel((const void*)render_exact_grid_kernel, dim3((unsigned)blocks), dim3(kRenderThreads), args, 0, stream);
    }
    if (p.plan.stream_views && !(fast_path_eligible(p) && !p.force_general)) {
        // only the fast kernel streams views: AOV / validation renders of a large scene use the generic filters
        p.plan = generic_plan(p.sc);
    }
    if (fast_path_eligible(p) && !p.force_general)
    {
        int dev = 0, optin = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        const size_t limit = optin > 4096 ? (size_t)optin - 4096 : 0;  // static shared memory of the kernel
        if (p.plan.stream_views) {
            if (p.hits && two_pass_wanted(p)) {
                // primary rays need the camera's view only: it stays resident in a launch of its own (no lock step), and the
                // shading launch streams the views of the lights
                RenderParams p1 = p;
                p1.plan = SmemPlan{};
                p1.plan.use_views = 1;
                p1.plan.view_bytes = p.plan.one_view_bytes; p1.plan.one_view_bytes = p.plan.one_view_bytes;
                p1.plan.total = p.plan.one_view_bytes;
                const void* f1 = (const void*)render_fast_kernel<1, false, false, false, 1>;
                cudaError_t e = launch_render_fn(f1, f1, kFastThreads, kFastMinBlocks, 1, p1, sm_count, stream, true, limit, kFastThreads, 0, RTE_P1_BATCH != 0);
                if (e != cudaSuccess) return e;
                if ((e = cudaMemsetAsync(p.tile_counter, 0, sizeof(unsigned int), stream)) != cudaSuccess) return e;
                const void* f2 = (const void*)render_fast_kernel<1, true, true, false, 2>;
                return launch_render_fn(f2, f2, kBigThreads, 1, 1, p, sm_count, stream, true, limit, kBigThreads);
            }
            const void* fn = (const void*)render_fast_kernel<kStreamPixels, true, true>;
            return launch_render_fn(fn, fn, kBigThreads, 1, kStreamPixels, p, sm_count, stream, true, limit, kBigThreads);
        }
        if (p.sc.nt > 0)
            return launch_render_fn((const void*)render_fast_kernel<kViewPixels, false, false, true>,
                                    (const void*)render_fast_kernel<kViewPixels, true, false, true>, kTrisThreads, kFastMinBlocks, kViewPixels, p,
                                    sm_count, stream, true, limit, kBigThreads);
        if (p.hits && two_pass_wanted(p)) {
            // two launches of half the code each: primary rays (nearest hit -> p.hits), then shading
            const void* f1 = (const void*)render_fast_kernel<1, false, false, false, 1>;
            const void* f2 = (const void*)render_fast_kernel<1, RTE_P2_BIG != 0, false, false, 2>;
            const int t2 = RTE_P2_BIG ? kBigThreads : kFastThreads;
            cudaError_t e = launch_render_fn(f1, f1, kFastThreads, kFastMinBlocks, 1, p, sm_count, stream, true, limit, kFastThreads);  // (CTA batches + CTA-level cull measured 2.69 -> 2.73 ms here)
            if (e != cudaSuccess) return e;
            if ((e = cudaMemsetAsync(p.tile_counter, 0, sizeof(unsigned int), stream)) != cudaSuccess) return e;
#if RTE_P2_STATE
            if (!RTE_P2_BIG) {  // hit point and normal in shared memory when 48 B per thread fit behind the cull scratch
                const void* f2s = (const void*)render_fast_kernel<1, false, false, false, 2, true>;
                e = launch_render_fn(f2s, f2s, t2, 1, 1, p, sm_count, stream, true, limit, t2, (size_t)t2 * 48);
                if (e != cudaErrorInvalidConfiguration) return e;
            }
#endif
            return launch_render_fn(f2, f2, t2, 1, 1, p, sm_count, stream, true, limit, t2);
        }
        return launch_render_fn((const void*)render_fast_kernel<kViewPixels, false, false>, (const void*)render_fast_kernel<kViewPixels, true, false>,
                                kFastThreads, kFastMinBlocks, kViewPixels, p, sm_count, stream, true, limit, kBigThreads);
    }
    if (p.sc.reflective) {
        const void* fr = (const void*)render_kernel<1, true>;
        return launch_render_fn(fr, fr, kReflThreads, kReflBlocks, 1, p, sm_count, stream, false, 0, kReflThreads);
    }
    const void* fn = (const void*)render_kernel<1, false>;
    return launch_render_fn(fn, fn, kRenderThreads, 2, 1, p, sm_count, stream);
}

// ------------------------------------------------------------------------------ ray queries
struct QueryParams {
    DevScene sc;
    SmemPlan plan;  // generic arrays only (no views: query rays have arbitrary origins)
    int64_t n;
    const double* origins;
    const double* dirs;
    const double* max_dist;
    rte_hit* hits;
    uint8_t* any_out;
};

// Scene::IntersectClosest, RE/Scene.h:218-257, one ray per thread
__global__ void __launch_bounds__(kRenderThreads, 2) closest_kernel(const __grid_constant__ QueryParams Q) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ DevScene s_sc;
    __shared__ __align__(8) uint64_t s_bar;
    if (threadIdx.x == 0) s_sc = Q.sc;
    StagedScene S;
    stage_scene(Q.sc, Q.plan, dyn_smem, &s_bar, S);
    const float4 *s_sph = S.sph, *s_tri = S.tri;
    const DevScene& sc = s_sc;
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < Q.n; base += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = base + threadIdx.x;
        const bool alive = i < Q.n;
        const d3 o = alive ? ld3(Q.origins + 3 * i) : mk(0, 0, 0);
        const d3 d = alive ? ld3(Q.dirs + 3 * i) : mk(0, 0, 1);
        Hit h;
        uint32_t confirms = 0;
        closest_hit_body(sc, s_sph, s_tri, o, d, alive, false, h, confirms);
        if (!alive) continue;
        rte_hit out;
        out.distance = INFINITY; out.prim_id = kMiss; out.material = -1;
        out.normal[0] = out.normal[1] = out.normal[2] = 0.0;
        out.point[0] = out.point[1] = out.point[2] = 0.0;
        if (h.kind != RTE_HIT_NONE) {
            const d3 Pn = o + d * h.t;
            d3 ng;
            uint32_t report = h.idx;
            if (h.kind == RTE_HIT_SPHERE) { const double4 s = sc.sph64[h.idx]; ng = normalize(Pn - mk(s.x, s.y, s.z)); }
            else if (h.kind == RTE_HIT_PLANE) ng = ld3(sc.planes[h.idx].n);
            else { ng = ld3(sc.tri64[h.idx].n); report = (uint32_t)sc.tri_info[h.idx].y; }
            out.distance = h.t;
            out.normal[0] = ng.x; out.normal[1] = ng.y; out.normal[2] = ng.z;
            out.point[0] = Pn.x; out.point[1] = Pn.y; out.point[2] = Pn.z;
            out.prim_id = (h.kind << 30) | report;
            const int m = material_of(sc, h);
            out.material = m < sc.nm ? m : -1;
        }
        Q.hits[i] = out;
    }
}

// Scene::IntersectAnyBefore, RE/Scene.h:259-276: any primitive reporting 0 < t < maxDist
__global__ void __launch_bounds__(kRenderThreads, 2) any_before_kernel(const __grid_constant__ QueryParams Q) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ DevScene s_sc;
    __shared__ __align__(8) uint64_t s_bar;
    if (threadIdx.x == 0) s_sc = Q.sc;
    StagedScene S;
    stage_scene(Q.sc, Q.plan, dyn_smem, &s_bar, S);
    const float4 *s_sph = S.sph, *s_tri = S.tri;
    const DevScene& sc = s_sc;
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < Q.n; base += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = base + threadIdx.x;
        const bool alive = i < Q.n;
        const d3 o = alive ? ld3(Q.origins + 3 * i) : mk(0, 0, 0);
        const d3 d = alive ? ld3(Q.dirs + 3 * i) : mk(0, 0, 1);
        const double md = alive ? Q.max_dist[i] : 0.0;
        uint32_t confirms = 0;
        // bias = 0: "t <= bias" can only be a plane at t == 0, which within() rejects as well
        const OccResult r = occluded(sc, s_sph, s_tri, o, d, md, 0.0, alive, false, confirms);
        if (alive) Q.any_out[i] = r.occluded ? 1 : 0;
    }
}

static size_t smem_budget() {  // dynamic shared memory one CTA of this device may have (minus the kernels' static part)
    static thread_local int dev_cached = -1;
    static thread_local size_t budget = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev != dev_cached) {
        int optin = 0;
        cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        budget = optin > 8192 ? (size_t)optin - 8192 : 0;
        dev_cached = dev;
    }
    return budget;
}
// generic FP32 filter arrays only: staged when they fit (two CTAs per SM: half the budget), else read from global memory
static SmemPlan generic_plan(const DevScene& sc) {
    SmemPlan pl{};
    const size_t sph = (size_t)sc.ns_pad * 16, tri = (size_t)sc.nt_pad * 48, budget = smem_budget();
    pl.stage_tri = tri <= budget ? 1 : 0;
    pl.stage_sph = (pl.stage_tri ? tri : 0) + sph <= budget ? 1 : 0;
    pl.sph_bytes = pl.stage_sph ? (uint32_t)sph : 0u;
    pl.tri_bytes = pl.stage_tri ? (uint32_t)tri : 0u;
    pl.total = (size_t)pl.sph_bytes + pl.tri_bytes;
    return pl;
}

static cudaError_t launch_query(const void* fn, const QueryParams& q, int sm_count, cudaStream_t stream) {
    const size_t smem = q.plan.total;
    cudaError_t e = ensure_smem(fn, smem);
    if (e != cudaSuccess) return e;
    long long blocks = (q.n + kRenderThreads - 1) / kRenderThreads;
    if (blocks > 2ll * sm_count) blocks = 2ll * sm_count;
    if (blocks < 1) blocks = 1;
    void* args[] = {(void*)&q};
    return cudaLaunchKernel(fn, dim3((unsigned)blocks), dim3(kRenderThreads), args, smem, stream);
}

cudaError_t launch_intersect_closest(const DevScene& sc, int64_t n, const double* origins, const double* dirs,
                                     rte_hit* hits, int sm_count, cudaStream_t stream) {
    QueryParams q{sc, generic_plan(sc), n, origins, dirs, nullptr, hits, nullptr};
    return launch_query((const void*)closest_kernel, q, sm_count, stream);
}
cudaError_t launch_intersect_any_before(const DevScene& sc, int64_t n, const double* origins, const double* dirs,
                                        const double* max_dist, uint8_t* out, int sm_count, cudaStream_t stream) {
    QueryParams q{sc, generic_plan(sc), n, origins, dirs, max_dist, nullptr, out};
    return launch_query((const void*)any_before_kernel, q, sm_count, stream);
}

}  // namespace rte
