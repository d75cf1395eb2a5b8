// rte_render.cu — the per-pixel render kernel (sm_100a).
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
//  * reflection (material.specular) is an iterative loop; refraction is rejected at upload time.
#include "rte_kernels.h"
#include "rte_tonemap.cuh"
#include "rte_trace.cuh"
#include "rte_views.cuh"

namespace rte {

// ------------------------------------------------------------------------------ TMA / mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(phase)
            : "memory");
    }
}

constexpr uint32_t kTmaChunk = 32768;

// Dynamic shared memory layout of the render kernel: [views (1+nl) x ns_pad x 16 B][generic sph32][tri32].
// The generic sphere array is only staged when something can use it often (reflection rays, or no views);
// otherwise the rare generic queries (exact shadow marches) read it from global memory.
SmemPlan plan_smem(const DevScene& sc, size_t budget) {
    SmemPlan pl{};
    const size_t tri = (size_t)sc.nt_pad * 48, sph = (size_t)sc.ns_pad * 16;
    const size_t views = sc.ns > 0 ? (size_t)(1 + sc.nl) * view_bytes(sc.ns_pad) : 0;
    pl.tri_bytes = (uint32_t)tri;
    if (sc.views && tri + views <= budget) {
        pl.use_views = 1;
        pl.view_bytes = (uint32_t)views;
        pl.stage_sph = (sc.reflective && tri + views + sph <= budget) ? 1 : 0;
    } else {
        pl.stage_sph = 1;
    }
    pl.sph_bytes = pl.stage_sph ? (uint32_t)sph : 0u;
    pl.total = (size_t)pl.view_bytes + pl.sph_bytes + pl.tri_bytes;
    return pl;
}

size_t scene_smem_bytes(const DevScene& sc) { return (size_t)sc.ns_pad * 16 + (size_t)sc.nt_pad * 48; }

__device__ __forceinline__ void tma_chunks(unsigned char* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    for (uint32_t off = 0; off < bytes; off += kTmaChunk)
        tma_bulk_g2s(dst + off, reinterpret_cast<const unsigned char*>(src) + off, min(kTmaChunk, bytes - off), bar);
}

// Stage the filter arrays into dynamic shared memory with TMA bulk copies.  Called by every thread of the CTA.
__device__ __forceinline__ void stage_scene(const DevScene& sc, const SmemPlan& pl, unsigned char* dyn, uint64_t* bar,
                                            const ulonglong2*& s_views, const float4*& s_sph, const float4*& s_tri) {
    s_views = reinterpret_cast<const ulonglong2*>(dyn);
    s_sph = pl.stage_sph ? reinterpret_cast<const float4*>(dyn + pl.view_bytes) : sc.sph32;
    s_tri = reinterpret_cast<const float4*>(dyn + pl.view_bytes + pl.sph_bytes);
    if (threadIdx.x == 0) mbar_init(bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, pl.view_bytes + pl.sph_bytes + pl.tri_bytes);
        tma_chunks(dyn, sc.views, pl.view_bytes, bar);
        tma_chunks(dyn + pl.view_bytes, sc.sph32, pl.sph_bytes, bar);
        tma_chunks(dyn + pl.view_bytes + pl.sph_bytes, sc.tri32, pl.tri_bytes, bar);
    }
    mbar_wait(bar, 0);
}

// Scene::backgroundColor, RE/Scene.h:30-33
__device__ __forceinline__ d3 background(d3 dir) {
    const double t = 0.5 * (normalize(dir).y + 1.0);
    return mk(1.0, 1.0, 1.0) * (1.0 - t) + mk(0.5, 0.7, 1.0) * t;
}

struct PixelCounters {
    uint32_t shadow, secondary, sph, pl, tri, marches, confirms;
};

// TraceRay for one pixel, all 32 lanes of the warp in lock step (lanes outside the image: alive=false).
__device__ __forceinline__ d3 trace_pixel(const RenderParams& P, const DevScene& sc, const ulonglong2* s_views, const float4* s_sph,
                                          const float4* s_tri, int px, int py, bool inb, int64_t out_idx,
                                          PixelCounters& cnt) {
    d3 o = ld3(P.cam.pos);
    d3 d = camera_dir(P.cam, px, py);
    d3 acc = mk(0.0, 0.0, 0.0);
    double w = 1.0;
    bool alive = inb;
    const double bias = P.bias;
    const int max_depth = P.max_depth, nl = sc.nl;
    const bool nf = P.no_filter != 0;
    const bool uv = P.plan.use_views != 0;
    const int view_stride = sc.ns_pad;  // ulonglong2 elements per view
#pragma unroll 1
    for (int depth = 0; depth < max_depth; ++depth) {
        if (!__any_sync(kFullMask, alive)) break;
        Hit h;
        if (uv && depth == 0) closest_hit_view(sc, s_views, s_tri, o, d, alive, nf, h, cnt.confirms);
        else closest_hit_body(sc, s_sph, s_tri, o, d, alive, nf, h, cnt.confirms);
        if (alive) {
            cnt.sph += (uint32_t)sc.ns; cnt.pl += (uint32_t)sc.np; cnt.tri += (uint32_t)sc.nt;
            if (depth > 0) cnt.secondary += 1;
        }
        const bool hitv = alive && h.kind != RTE_HIT_NONE;
        if (alive && !hitv) {
            acc = acc + background(d) * w;
            alive = false;
        }
        // hit record (GetHitInfoAt, RE/Shape.h:104-121,165-180,229-243)
        d3 Pn = mk(0, 0, 0), ng = mk(0, 0, 0);
        int mat = sc.nm;
        uint32_t report = 0;
        if (hitv) {
            Pn = o + d * h.t;
            mat = material_of(sc, h);
            if (h.kind == RTE_HIT_SPHERE) {
                const double4 s = sc.sph64[h.idx];
                ng = normalize(Pn - mk(s.x, s.y, s.z));
                report = h.idx;
            } else if (h.kind == RTE_HIT_PLANE) {
                ng = ld3(sc.planes[h.idx].n);
                report = h.idx;
            } else {
                ng = ld3(sc.tri64[h.idx].n);
                report = (uint32_t)sc.tri_info[h.idx].y;
            }
        }
        const rte_material& M = sc.materials[mat];
        // TraceRay, RE/Scene.h:144-148
        const d3 incoming = normalize(d);
        const bool frontFace = dot(ng, incoming) < 0.0;
        const d3 N = frontFace ? ng : -ng;
        const d3 viewDir = -incoming;
        // directLightning, RE/Scene.h:79-129
        const d3 normal = normalize(N);
        d3 diffuseAcc = mk(0, 0, 0), specAcc = mk(0, 0, 0);
        uint32_t lmask = 0;
        const bool want_spec = hitv && M.transparency <= 0.0 && M.specular > 0.0;
#pragma unroll 1
        for (int j = 0; j < nl; ++j) {
            const d3 Lp = ld3(sc.lights[j].pos);
            const d3 vecToLight = Lp - Pn;
            const double dist = length(vecToLight);
            bool cast = hitv && dist > 0.0;
            d3 L = mk(0, 0, 1);
            double ndl = 0.0;
            if (cast) {
                L = vecToLight / dist;
                ndl = std_max(0.0, dot(normal, L));
                cast = ndl > 0.0 && dist > bias;
            }
            if (!__any_sync(kFullMask, cast)) continue;
            const d3 so = Pn + normal * bias;
            const double maxDist = dist - bias;
            double T = 1.0;
            if (P.exact_shadows) {
                if (cast) { T = march_exact(sc, s_sph, s_tri, so, L, maxDist, bias, nf, cnt.confirms); cnt.marches += 1; }
            } else {
                const OccResult r = uv ? occluded_view(sc, s_views + (size_t)(1 + j) * view_stride, s_tri, so, L, dist, Lp, maxDist, bias, cast, nf, cnt.confirms)
                                       : occluded(sc, s_sph, s_tri, so, L, maxDist, bias, cast, nf, cnt.confirms);
                cnt.sph += r.sph_tests; cnt.pl += r.pl_tests; cnt.tri += r.tri_tests;
                if (cast) {
                    if (r.near) { T = march_exact(sc, s_sph, s_tri, so, L, maxDist, bias, nf, cnt.confirms); cnt.marches += 1; }
                    else T = r.occluded ? 0.0 : 1.0;
                }
            }
            if (cast) {
                cnt.shadow += 1;
                if (j < 16) lmask |= 1u << (16 + j);
                if (T > bias) {
                    if (j < 16) lmask |= 1u << j;
                    const d3 emitted = ld3(sc.lights[j].E);
                    const d3 contribution = (emitted * (1.0 / (dist * dist))) * ndl;
                    diffuseAcc = diffuseAcc + contribution * T;
                    if (want_spec) {
                        const d3 H = normalize(L + viewDir);
                        const double ndh = std_max(0.0, dot(normal, H));
                        if (ndh > 0.0) {
                            const double specFactor = pow(ndh, M.shininess);
                            specAcc = specAcc + ((emitted * (1.0 / (dist * dist))) * specFactor) * T;
                        }
                    }
                }
            }
        }
        if (hitv) {
            const d3 local = ld3(M.color) * diffuseAcc + specAcc * M.specular;
            // finalLight += localLight * (1.0 - transparency), transparency == 0 (RE/Scene.h:171-173)
            acc = acc + (local * (1.0 - 0.0)) * w;
        }
        if (depth == 0 && inb) {  // AOVs of the primary hit
            if (P.out.prim_id) P.out.prim_id[out_idx] = hitv ? ((h.kind << 30) | report) : kMiss;
            if (P.out.depth) P.out.depth[out_idx] = hitv ? (float)h.t : INFINITY;
            if (P.out.depth64) P.out.depth64[out_idx] = hitv ? h.t : (double)INFINITY;
            if (P.out.normal) reinterpret_cast<float4*>(P.out.normal)[out_idx] = make_float4((float)ng.x, (float)ng.y, (float)ng.z, 0.f);
            if (P.out.albedo) {
                const d3 a = hitv ? ld3(M.color) : background(d);
                reinterpret_cast<float4*>(P.out.albedo)[out_idx] = make_float4((float)a.x, (float)a.y, (float)a.z, 0.f);
            }
            if (P.out.light_mask) P.out.light_mask[out_idx] = lmask;
        }
        // reflection, RE/Scene.h:189-195 (transparency == 0: reflectiveness = material.specular)
        if (hitv && M.specular > bias) {
            const d3 R = normalize(reflect(incoming, N));
            w = w * M.specular;
            o = Pn + R * bias;
            d = R;
        } else {
            alive = false;
        }
    }
    if (alive) acc = acc + background(d) * w;  // TraceRay(depth >= maxRecursion), RE/Scene.h:132-134
    return acc;
}

// ------------------------------------------------------------------------------ the kernel
constexpr int kRenderThreads = 256;     // block size of the query kernels and of the render kernel at 2 CTAs/SM
constexpr int kRenderMaxThreads = 512;  // render kernel at 1 CTA/SM (large scenes: more shared memory per CTA)
constexpr int kTileW = 8, kTileH = 4;  // one warp = 8x4 pixels (coherent primary and shadow rays)

__global__ void __launch_bounds__(kRenderMaxThreads, 1) render_kernel(const __grid_constant__ RenderParams P) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ DevScene s_sc;
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ unsigned long long s_stats[kRenderMaxThreads / 32][kStatCount];

    if (threadIdx.x == 0) s_sc = P.sc;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < kStatCount) s_stats[warp][lane] = 0ull;
    const float4 *s_sph, *s_tri;
    const ulonglong2* s_views;
    stage_scene(P.sc, P.plan, dyn_smem, &s_bar, s_views, s_sph, s_tri);  // contains __syncthreads: s_sc visible after it
    const DevScene& sc = s_sc;

    const int W = P.cam.width, rows = P.n_rows;
    const int sr = P.stripe_rows, sp = P.stripe_period;
    const int tiles_x = (W + kTileW - 1) / kTileW, tiles_y = (rows + kTileH - 1) / kTileH;
    const unsigned int n_tiles = (unsigned int)(tiles_x * tiles_y);
    const int lx = lane & (kTileW - 1), ly = lane >> 3;

    unsigned int tile = 0;
    if (lane == 0) tile = atomicAdd(P.tile_counter, 1u);
    tile = __shfl_sync(kFullMask, tile, 0);
    while (tile < n_tiles) {
        unsigned int next = 0;
        if (lane == 0) next = atomicAdd(P.tile_counter, 1u);  // prefetch the next tile id behind the work
        const int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile % (unsigned)tiles_x);
        const int px = tx * kTileW + lx, band_y = ty * kTileH + ly;
        const int py = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
        const bool inb = px < W && band_y < rows;
        const int64_t out_idx = (int64_t)band_y * W + px;
        PixelCounters cnt = {0, 0, 0, 0, 0, 0, 0};
        const d3 c = trace_pixel(P, sc, s_views, s_sph, s_tri, px, py, inb, out_idx, cnt);
        if (inb) {
            if (P.out.hdr) reinterpret_cast<float4*>(P.out.hdr)[out_idx] = make_float4((float)c.x, (float)c.y, (float)c.z, 0.f);
            if (P.out.hdr64) { P.out.hdr64[3 * out_idx] = c.x; P.out.hdr64[3 * out_idx + 1] = c.y; P.out.hdr64[3 * out_idx + 2] = c.z; }
            if (P.out.rgb8 && P.tonemap_op >= 0) {
                const uchar3 q = tm64::to_color(tm64::apply(P.tonemap_op, c, P.max_white, P.jodie_a));
                P.out.rgb8[3 * out_idx] = q.x; P.out.rgb8[3 * out_idx + 1] = q.y; P.out.rgb8[3 * out_idx + 2] = q.z;
            }
        }
        // per-warp statistics: one REDUX per counter, lane 0 accumulates in shared memory
        const uint32_t v[kStatCount] = {inb ? 1u : 0u, cnt.shadow, cnt.secondary, cnt.sph, cnt.pl, cnt.tri, cnt.marches, cnt.confirms};
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

static cudaError_t ensure_smem(const void* fn, size_t bytes) {
    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

cudaError_t launch_build_views(const DevScene& sc, const DevCamera& cam, double bias, cudaStream_t stream) {
    if (!sc.views || sc.ns == 0) return cudaSuccess;
    const int total = (1 + sc.nl) * sc.ns_pad;
    build_views_kernel<<<(total + 255) / 256, 256, 0, stream>>>(sc, cam, bias, const_cast<float*>(sc.views));
    return cudaGetLastError();
}

cudaError_t launch_render(const RenderParams& p, int sm_count, cudaStream_t stream) {
    const size_t smem = p.plan.total;
    cudaError_t e = ensure_smem((const void*)render_kernel, smem);
    if (e != cudaSuccess) return e;
    // 2 CTAs of 256 threads per SM when the arrays fit twice, else 1 CTA of 512 threads
    int threads = kRenderThreads, per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, render_kernel, threads, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 2) {
        threads = kRenderMaxThreads;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, render_kernel, threads, smem);
        if (e != cudaSuccess) return e;
    }
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    const int W = p.cam.width, rows = p.n_rows;
    const long long n_tiles = (long long)((W + kTileW - 1) / kTileW) * ((rows + kTileH - 1) / kTileH);
    long long blocks = (long long)sm_count * per_sm;  // persistent: a multiple of the SM count
    const long long needed = (n_tiles + threads / 32 - 1) / (threads / 32);
    if (blocks > needed) blocks = needed > 0 ? needed : 1;
    render_kernel<<<(unsigned)blocks, threads, smem, stream>>>(p);
    return cudaGetLastError();
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
    const float4 *s_sph, *s_tri;
    const ulonglong2* s_views;
    stage_scene(Q.sc, Q.plan, dyn_smem, &s_bar, s_views, s_sph, s_tri);
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
    const float4 *s_sph, *s_tri;
    const ulonglong2* s_views;
    stage_scene(Q.sc, Q.plan, dyn_smem, &s_bar, s_views, s_sph, s_tri);
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

static SmemPlan generic_plan(const DevScene& sc) {
    SmemPlan pl{};
    pl.stage_sph = 1;
    pl.sph_bytes = (uint32_t)((size_t)sc.ns_pad * 16);
    pl.tri_bytes = (uint32_t)((size_t)sc.nt_pad * 48);
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
