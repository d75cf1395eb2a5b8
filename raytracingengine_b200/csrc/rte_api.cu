// rte_api.cu — the extern "C" boundary of librte_b200.so (include/rte.h): context, scene packing,
// launches, host-buffer variants, PPM output.  No torch types, no CPU compute path: every compute entry
// point needs a CUDA device and fails with RTE_ERR_NO_DEVICE / RTE_ERR_CUDA otherwise.
//
// Scene packing replaces what the reference does implicitly in Scene::Add* / the shape constructors
// ("RE/" = /root/reference/RaytracingEngine/): Plane ctor normalisation (RE/Shape.h:141-147), Triangle
// tv0..tv2 / edges / flat normal (RE/Shape.h:198-227), Model -> triangles in index order reporting the
// MODEL index (RE/Shape.h:258-283), Light E = color * intensity (RE/Scene.h:110).  All of that is FP64
// host arithmetic done once per upload, in the reference's operation order.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "rte_kernels.h"

using namespace rte;

namespace {

thread_local std::string tl_error;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

}  // namespace

constexpr int kHostBandsMax = 3;    // row bands of one rte_render_host call
constexpr int kHostBandRows = 1024;  // calls of fewer rows are one band

struct rte_ctx {
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    std::string err;
    bool sticky = false;

    // scene
    bool has_scene = false, has_camera = false;
    DevScene sc{};
    DevCamera cam{};
    rte_camera cam_host{};
    DevBuf b_sph32, b_tri32, b_sph64, b_sphhl, b_sphmat, b_planes, b_tri64, b_triiv, b_triinfo, b_mats, b_lights, b_views, b_tviews, b_views_cold, b_tviews_cold;
    bool views_dirty = true;
    double views_bias = -1.0;
    // FP32 production path: per-origin direction grids (rte_grid.cuh), rebuilt when the scene, the camera or the bias changes
    DevBuf b_grid_count, b_grid_start, b_grid_entries, b_grid_aggr;
    uint32_t grid_epoch = 0;
    long long grid_shape = -1;        // (ns, nl, G) of the build the entries buffer was sized for
    bool grid_check_pending = false;  // the last build was not followed by a read-back of its entry count yet
    bool grid_dirty = true;
    size_t grid_call_px = 0;  // pixels of the API call being served (ensure_grid sizes the cells by it)
    double grid_bias = -1.0;
    int grid_G = 0;
    unsigned int* h_grid_total = nullptr;  // pinned: number of (sphere, cell) entries of the last build

    // rte_set_rgb8_peers
    uint8_t* peer_rgb8[kMaxPeers] = {};
    int n_peers = 0;
    // rte_set_fence_peers
    uint32_t* fence_flags[kMaxPeers] = {};
    int n_fence = 0, fence_rank = 0;
    uint32_t fence_seq = 0;
    unsigned int* d_fence_counter = nullptr;
    unsigned int* h_fence_timeout = nullptr;  // mapped pinned word, set by the wait kernel
    // rte_upload_scene: pinned staging of the flattened scene arrays (the H2D copies are asynchronous, no stream sync at the end)
    unsigned char* h_stage = nullptr;
    size_t h_stage_cap = 0, h_stage_off = 0;
    cudaEvent_t ev_stage = nullptr;

    // launch scratch
    unsigned int* d_tile_counter = nullptr;
    unsigned long long* d_stats = nullptr;
    unsigned int* d_maxlum = nullptr;
    unsigned long long* h_stats = nullptr;  // pinned
    cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
    // rte_render_host: device->host copies of a finished row band run on their own stream under the next band's kernels
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_band[kHostBandsMax] = {};
    cudaEvent_t ev_copied = nullptr;
    bool render_timed = false, tonemap_timed = false;
    int render_launches = 0;  // kernels enqueued by the last render call
    rte_stats stats{};

    // staging for the *_host variants (grow-only)
    DevBuf s_hdr, s_hdr64, s_rgb8, s_prim, s_depth, s_depth64, s_normal, s_albedo, s_lmask, s_tm64, s_hits, s_fix, s_dnorm, s_dn64, s_tileflags;
    DevBuf s_tm_in, s_tm_out[RTE_TM_COUNT], s_q_o, s_q_d, s_q_m, s_q_out;
};

namespace {

int fail(rte_ctx* ctx, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    tl_error = buf;
    if (ctx) ctx->err = buf;
    return code;
}

// Errors that leave the CUDA context healthy (an oversized allocation, a launch configuration the device cannot run, a bad
// pointer handed to a copy) fail the call only; everything else (launch failures, illegal addresses, ...) poisons the CUDA
// context for good and makes the ctx sticky: every later call reports it until rte_destroy.
int cuda_fail(rte_ctx* ctx, cudaError_t e, const char* what) {
    const bool limit = e == cudaErrorMemoryAllocation || e == cudaErrorLaunchOutOfResources || e == cudaErrorInvalidConfiguration;
    const bool recoverable = limit || e == cudaErrorInvalidValue;
    if (recoverable) (void)cudaGetLastError();  // (clears the per-thread last-error slot)
    else if (ctx) ctx->sticky = true;
    return fail(ctx, limit ? RTE_ERR_LIMIT : RTE_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
}

#define RTE_CUDA(ctx, call)                                       \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call); \
    } while (0)

#define RTE_ENTER(ctx)                                                                        \
    do {                                                                                      \
        if (!(ctx)) return fail(nullptr, RTE_ERR_INVALID, "%s: ctx is NULL", __func__);        \
        if ((ctx)->sticky) return fail(nullptr, RTE_ERR_CUDA, "ctx is in a failed CUDA state: %s", (ctx)->err.c_str()); \
        cudaError_t e__ = cudaSetDevice((ctx)->device);                                       \
        if (e__ != cudaSuccess) return cuda_fail(ctx, e__, "cudaSetDevice");                  \
    } while (0)

int reserve(rte_ctx* ctx, DevBuf& b, size_t bytes) {
    if (bytes == 0) bytes = 16;
    if (b.cap >= bytes) return RTE_OK;
    if (b.p) {
        RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        RTE_CUDA(ctx, cudaFree(b.p));
        b.p = nullptr; b.cap = 0;
    }
    RTE_CUDA(ctx, cudaMalloc(&b.p, bytes));
    b.cap = bytes;
    return RTE_OK;
}

// The staging buffer holds `bytes` more (grow-only); waits for the copies of the previous upload that still read it.
int stage_begin(rte_ctx* ctx, size_t bytes) {
    if (!ctx->ev_stage) RTE_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_stage, cudaEventDisableTiming));
    else RTE_CUDA(ctx, cudaEventSynchronize(ctx->ev_stage));
    if (ctx->h_stage_cap < bytes) {
        if (ctx->h_stage) RTE_CUDA(ctx, cudaFreeHost(ctx->h_stage));
        ctx->h_stage = nullptr; ctx->h_stage_cap = 0;
        RTE_CUDA(ctx, cudaMallocHost(&ctx->h_stage, bytes + bytes / 2));
        ctx->h_stage_cap = bytes + bytes / 2;
    }
    ctx->h_stage_off = 0;
    return RTE_OK;
}
template <class T>
size_t stage_bytes(const std::vector<T>& v) { return (v.size() * sizeof(T) + 255) & ~(size_t)255; }

template <class T>
int upload(rte_ctx* ctx, DevBuf& b, const std::vector<T>& v) {
    int rc = reserve(ctx, b, v.size() * sizeof(T));
    if (rc) return rc;
    if (v.empty()) return RTE_OK;
    const size_t bytes = v.size() * sizeof(T);
    if (ctx->h_stage && ctx->h_stage_off + bytes <= ctx->h_stage_cap) {  // through the pinned staging buffer: truly asynchronous
        unsigned char* h = ctx->h_stage + ctx->h_stage_off;
        memcpy(h, v.data(), bytes);
        ctx->h_stage_off += (bytes + 255) & ~(size_t)255;
        RTE_CUDA(ctx, cudaMemcpyAsync(b.p, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    } else {
        RTE_CUDA(ctx, cudaMemcpyAsync(b.p, v.data(), bytes, cudaMemcpyHostToDevice, ctx->stream));
        RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // pageable source about to go out of scope
    }
    return RTE_OK;
}

// ---- FP64 host Vec3 in the reference's operation order (RE/Math.h:14-37), compiled without contraction
struct H3 { double x, y, z; };
inline H3 h3(const double* p) { return H3{p[0], p[1], p[2]}; }
inline H3 hadd(H3 a, H3 b) { return H3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline H3 hsub(H3 a, H3 b) { return H3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline H3 hmul(H3 a, double s) { return H3{a.x * s, a.y * s, a.z * s}; }
inline double hdot(H3 a, H3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline H3 hcross(H3 a, H3 b) { return H3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline double hlen(H3 a) { return std::sqrt(hdot(a, a)); }
inline H3 hnorm(H3 a) {
    const double l = hlen(a);
    if (l <= 1e-12) return H3{0, 0, 0};
    return H3{a.x / l, a.y / l, a.z / l};
}
inline void put3(double* d, H3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
inline float up(float v) { return std::nextafter(v, INFINITY); }
inline bool finite3(const double* p) { return std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]); }

struct Defaults {
    int max_depth; int tonemap_op; double bias, max_white, jodie_a; int row0, row1; uint32_t flags;
    int stripe_rows, stripe_period, n_rows;
};

int resolve_opts(rte_ctx* ctx, const rte_render_opts* o, Defaults& d) {
    const int H = (int)ctx->cam_host.height;
    d.row0 = o ? o->row0 : 0;
    d.row1 = (o && o->row1 > 0) ? o->row1 : H;
    d.max_depth = (o && o->max_depth > 0) ? o->max_depth : 10;
    if (d.max_depth > RTE_MAX_DEPTH) return fail(ctx, RTE_ERR_LIMIT, "opts.max_depth = %d exceeds RTE_MAX_DEPTH = %d", d.max_depth, RTE_MAX_DEPTH);
    d.tonemap_op = o ? o->tonemap_op : RTE_TM_NONE;
    d.bias = (o && o->bias > 0) ? o->bias : 1e-3;
    d.max_white = (o && o->max_white > 0) ? o->max_white : 5.0;
    d.jodie_a = (o && o->jodie_a > 0) ? o->jodie_a : 0.18;
    d.flags = o ? (uint32_t)o->flags : 0u;
    if (d.row0 < 0 || d.row1 > H || d.row0 > d.row1)
        return fail(ctx, RTE_ERR_INVALID, "row band [%d,%d) outside the camera image (height %d)", d.row0, d.row1, H);
    d.stripe_rows = o ? o->stripe_rows : 0;
    d.stripe_period = o ? o->stripe_period : 0;
    d.n_rows = d.row1 - d.row0;
    if (d.stripe_rows != 0) {
        if (d.stripe_rows < 0 || d.stripe_rows % 4 || d.stripe_period < d.stripe_rows)
            return fail(ctx, RTE_ERR_INVALID, "stripe_rows %d must be a positive multiple of 4 and <= stripe_period %d", d.stripe_rows, d.stripe_period);
        const int span = d.row1 - d.row0, full = span / d.stripe_period, rem = span % d.stripe_period;
        d.n_rows = full * d.stripe_rows + (rem < d.stripe_rows ? rem : d.stripe_rows);
    }
    if (d.tonemap_op < RTE_TM_NONE || d.tonemap_op >= RTE_TM_COUNT)
        return fail(ctx, RTE_ERR_INVALID, "tonemap_op %d is not an RTE_TM_* value", d.tonemap_op);
    return RTE_OK;
}

}  // namespace

// ====================================================================================== basics
extern "C" int rte_version(void) { return RTE_VERSION; }

extern "C" int rte_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" const char* rte_last_error(const rte_ctx* ctx) { return ctx ? ctx->err.c_str() : tl_error.c_str(); }

extern "C" int rte_create(int device, rte_ctx** out) {
    if (!out) return fail(nullptr, RTE_ERR_INVALID, "rte_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(nullptr, RTE_ERR_NO_DEVICE, "no CUDA device visible (%s): librte_b200 has no CPU path",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= n) return fail(nullptr, RTE_ERR_INVALID, "device %d out of range [0,%d)", device, n);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceProperties");
    if (prop.major != 10)
        return fail(nullptr, RTE_ERR_NO_DEVICE, "device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
    rte_ctx* ctx = new (std::nothrow) rte_ctx();
    if (!ctx) return fail(nullptr, RTE_ERR_INVALID, "out of host memory");
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
#define CK(call) do { cudaError_t e2 = (call); if (e2 != cudaSuccess) { int rc = cuda_fail(nullptr, e2, #call); rte_destroy(ctx); return rc; } } while (0)
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CK(cudaMalloc(&ctx->d_stats, sizeof(unsigned long long) * (kStatCount + 2)));  // counters, tile counter, fix-up list length
    ctx->d_tile_counter = reinterpret_cast<unsigned int*>(ctx->d_stats + kStatCount);
    CK(cudaMalloc(&ctx->d_maxlum, sizeof(unsigned int)));
    CK(cudaMalloc(&ctx->d_fence_counter, sizeof(unsigned int)));
    CK(cudaMemset(ctx->d_fence_counter, 0, sizeof(unsigned int)));
    CK(cudaHostAlloc(&ctx->h_fence_timeout, sizeof(unsigned int), cudaHostAllocMapped));
    *ctx->h_fence_timeout = 0u;
    CK(cudaMallocHost(&ctx->h_stats, sizeof(unsigned long long) * (kStatCount + 1)));
    CK(cudaMallocHost(&ctx->h_grid_total, sizeof(unsigned int)));
    CK(cudaEventCreate(&ctx->ev_r0)); CK(cudaEventCreate(&ctx->ev_r1));
    CK(cudaEventCreate(&ctx->ev_t0)); CK(cudaEventCreate(&ctx->ev_t1));
    CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (cudaEvent_t& e : ctx->ev_band) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ctx->ev_copied, cudaEventDisableTiming));
#undef CK
    *out = ctx;
    return RTE_OK;
}

extern "C" void rte_destroy(rte_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    DevBuf* bufs[] = {&ctx->b_sph32, &ctx->b_tri32, &ctx->b_sph64, &ctx->b_sphhl, &ctx->b_sphmat, &ctx->b_planes, &ctx->b_tri64, &ctx->b_triiv, &ctx->b_triinfo,
                      &ctx->b_mats, &ctx->b_lights, &ctx->b_views, &ctx->b_tviews, &ctx->b_views_cold, &ctx->b_tviews_cold, &ctx->b_grid_count, &ctx->b_grid_start, &ctx->b_grid_entries, &ctx->b_grid_aggr, &ctx->s_hits, &ctx->s_fix, &ctx->s_hdr, &ctx->s_hdr64, &ctx->s_tm64, &ctx->s_rgb8, &ctx->s_prim, &ctx->s_depth,
                      &ctx->s_depth64, &ctx->s_dnorm, &ctx->s_dn64, &ctx->s_tileflags, &ctx->s_normal, &ctx->s_albedo, &ctx->s_lmask, &ctx->s_tm_in, &ctx->s_q_o, &ctx->s_q_d,
                      &ctx->s_q_m, &ctx->s_q_out};
    for (DevBuf* b : bufs) if (b->p) cudaFree(b->p);
    for (DevBuf& b : ctx->s_tm_out) if (b.p) cudaFree(b.p);
    if (ctx->d_stats) cudaFree(ctx->d_stats);
    if (ctx->d_maxlum) cudaFree(ctx->d_maxlum);
    if (ctx->d_fence_counter) cudaFree(ctx->d_fence_counter);
    if (ctx->h_fence_timeout) cudaFreeHost(ctx->h_fence_timeout);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    if (ctx->ev_stage) cudaEventDestroy(ctx->ev_stage);
    if (ctx->h_stats) cudaFreeHost(ctx->h_stats);
    if (ctx->h_grid_total) cudaFreeHost(ctx->h_grid_total);
    cudaEvent_t evs[] = {ctx->ev_r0, ctx->ev_r1, ctx->ev_t0, ctx->ev_t1};
    for (cudaEvent_t e : evs) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->ev_band) if (e) cudaEventDestroy(e);
    if (ctx->ev_copied) cudaEventDestroy(ctx->ev_copied);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    cudaGetLastError();
    delete ctx;
}

extern "C" int rte_set_stream(rte_ctx* ctx, void* cuda_stream) {
    RTE_ENTER(ctx);
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (cuda_stream) {
        if (ctx->own_stream) RTE_CUDA(ctx, cudaStreamDestroy(ctx->stream));
        ctx->stream = (cudaStream_t)cuda_stream;
        ctx->own_stream = false;
    } else if (!ctx->own_stream) {
        RTE_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        ctx->own_stream = true;
    }
    return RTE_OK;
}

static int check_grid_overflow(rte_ctx* ctx);
static int check_fence_timeout(rte_ctx* ctx) {
    if (ctx->h_fence_timeout && *ctx->h_fence_timeout) {
        *ctx->h_fence_timeout = 0u;
        return fail(ctx, RTE_ERR_TIMEOUT, "rte_frame_fence: a peer did not deliver frame %u in time", ctx->fence_seq);
    }
    return RTE_OK;
}

extern "C" int rte_synchronize(rte_ctx* ctx) {
    RTE_ENTER(ctx);
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (int rc = check_grid_overflow(ctx)) return rc;
    return check_fence_timeout(ctx);
}

// ====================================================================================== scene
extern "C" int rte_upload_scene(rte_ctx* ctx, const rte_scene_desc* d) {
    RTE_ENTER(ctx);
    if (!d) return fail(ctx, RTE_ERR_INVALID, "rte_upload_scene: scene is NULL");
    const int64_t counts[] = {d->n_materials, d->n_spheres, d->n_planes, d->n_triangles, d->n_models, d->n_model_vertices, d->n_model_indices, d->n_lights};
    const void* ptrs[] = {d->materials, d->spheres, d->planes, d->triangles, d->models, d->model_vertices, d->model_indices, d->lights};
    static const char* names[] = {"materials", "spheres", "planes", "triangles", "models", "model_vertices", "model_indices", "lights"};
    for (int i = 0; i < 8; ++i) {
        if (counts[i] < 0) return fail(ctx, RTE_ERR_INVALID, "n_%s is negative", names[i]);
        if (counts[i] > 0 && !ptrs[i]) return fail(ctx, RTE_ERR_INVALID, "%s is NULL with n_%s = %lld", names[i], names[i], (long long)counts[i]);
    }
    if (d->n_lights > 4096) return fail(ctx, RTE_ERR_LIMIT, "at most 4096 lights, got %lld", (long long)d->n_lights);  // (light_mask covers the first 16)
    if (d->n_spheres >= (1ll << 30) || d->n_planes >= (1ll << 30)) return fail(ctx, RTE_ERR_LIMIT, "primitive index does not fit 30 bits");

    // materials (+ the default Material at index nm, RE/Shape.h:13-19)
    std::vector<rte_material> mats(d->materials, d->materials + d->n_materials);
    rte_material def{};
    def.shininess = 128.0; def.refractive_index = 1.0;
    mats.push_back(def);

    // spheres
    const int ns = (int)d->n_spheres, ns_pad = ns > 96 ? (ns + 63) & ~63 : (ns + 31) & ~31;  // kViewBlock; views that get culled (>= kCullMin): whole groups of 64
    std::vector<float4> sph32((size_t)ns_pad);
    std::vector<double4> sph64((size_t)ns_pad);
    std::vector<int32_t> sphmat((size_t)ns_pad, -1);
    std::vector<float4> sphhl((size_t)ns_pad * 2, make_float4(0.f, 0.f, 0.f, 0.f));
    double s_ext = 0.0;
    for (int i = 0; i < ns; ++i) {
        const rte_sphere& s = d->spheres[i];
        if (!finite3(s.center) || !std::isfinite(s.radius)) return fail(ctx, RTE_ERR_INVALID, "sphere %d is not finite", i);
        sph64[i] = make_double4(s.center[0], s.center[1], s.center[2], s.radius);
        const float r = up((float)std::fabs(s.radius));
        sph32[i] = make_float4((float)s.center[0], (float)s.center[1], (float)s.center[2], up(r * r));
        sphmat[i] = s.material;
        {   // centre as hi + lo floats (FP32 production path, rte_render32.cu)
            const float hx = (float)s.center[0], hy = (float)s.center[1], hz = (float)s.center[2];
            sphhl[2 * (size_t)i] = make_float4(hx, hy, hz, (float)std::fabs(s.radius));
            sphhl[2 * (size_t)i + 1] = make_float4((float)(s.center[0] - (double)hx), (float)(s.center[1] - (double)hy), (float)(s.center[2] - (double)hz), 0.f);
        }
        s_ext = std::fmax(s_ext, hlen(h3(s.center)) + std::fabs(s.radius));
    }
    for (int i = ns; i < ns_pad; ++i) {  // never-hit padding: q = +inf
        sph32[i] = make_float4(0.f, 0.f, 0.f, -INFINITY);
        sph64[i] = make_double4(0, 0, 0, 0);
    }

    // planes — Plane ctor normalises the normal (RE/Shape.h:142)
    const int np = (int)d->n_planes;
    std::vector<DevPlane> planes((size_t)np);
    for (int i = 0; i < np; ++i) {
        const rte_plane& p = d->planes[i];
        if (!finite3(p.point) || !finite3(p.normal)) return fail(ctx, RTE_ERR_INVALID, "plane %d is not finite", i);
        put3(planes[i].n, hnorm(h3(p.normal)));
        put3(planes[i].p0, h3(p.point));
        planes[i].mat = p.material; planes[i].pad = 0;
    }

    // triangles: standalone ones, then every model's triples in order (canonical order of RE/Scene.h:218-257)
    int64_t nt64 = d->n_triangles;
    for (int64_t m = 0; m < d->n_models; ++m) {
        const rte_model& mo = d->models[m];
        if (mo.index_count < 0 || mo.index_count % 3) return fail(ctx, RTE_ERR_INVALID, "model %lld: index_count must be a non-negative multiple of 3", (long long)m);
        if (mo.first_vertex < 0 || mo.vertex_count < 0 || mo.first_vertex + mo.vertex_count > d->n_model_vertices ||
            mo.first_index < 0 || mo.first_index + mo.index_count > d->n_model_indices)
            return fail(ctx, RTE_ERR_INVALID, "model %lld: vertex/index range outside the shared arrays", (long long)m);
        nt64 += mo.index_count / 3;
    }
    if (nt64 >= (1ll << 30)) return fail(ctx, RTE_ERR_LIMIT, "too many triangles");
    const int nt = (int)nt64, nt_pad = (nt + 1) & ~1;
    std::vector<float4> tri32((size_t)nt_pad * 3);
    std::vector<DevTri64> tri64((size_t)nt_pad);
    std::vector<int2> triinfo((size_t)nt_pad, make_int2(-1, 0));
    std::vector<Tri32> triiv((size_t)nt_pad);
    double t_ext = 0.0, t_edge = 0.0;
    int k = 0;
    auto add_tri = [&](H3 v0, H3 v1, H3 v2, H3 T, int mat, int report) {
        const H3 a0 = hadd(v0, T);                 // tv0, RE/Shape.h:198
        const H3 e1 = hsub(hadd(v1, T), a0);       // edge1 = tv1 - tv0, :204-205
        const H3 e2 = hsub(hadd(v2, T), a0);       // edge2 = tv2 - tv0, :206
        const H3 n = hnorm(hcross(hsub(v1, v0), hsub(v2, v0)));  // GetNormalAt, :222-227 (untranslated vertices)
        put3(tri64[k].a0, a0); put3(tri64[k].e1, e1); put3(tri64[k].e2, e2); put3(tri64[k].n, n);
        tri32[3 * (size_t)k + 0] = make_float4((float)a0.x, (float)a0.y, (float)a0.z, 0.f);
        tri32[3 * (size_t)k + 1] = make_float4((float)e1.x, (float)e1.y, (float)e1.z, 0.f);
        tri32[3 * (size_t)k + 2] = make_float4((float)e2.x, (float)e2.y, (float)e2.z, 0.f);
        triinfo[k] = make_int2(mat, report);
        {
            Tri32& q = triiv[k];
            const double A0[3] = {a0.x, a0.y, a0.z}, E1[3] = {e1.x, e1.y, e1.z}, E2[3] = {e2.x, e2.y, e2.z}, Nn[3] = {n.x, n.y, n.z};
            for (int c = 0; c < 3; ++c) {
                q.ah[c] = (float)A0[c]; q.al[c] = (float)(A0[c] - (double)q.ah[c]);
                q.e1[c] = (float)E1[c]; q.e2[c] = (float)E2[c]; q.n[c] = (float)Nn[c];
            }
            q.E1 = (float)(hlen(e1) * 1.000001); q.E2 = (float)(hlen(e2) * 1.000001);
            q.mat = mat; q.report = (uint32_t)report; q.pad = 0.f;
        }
        t_ext = std::fmax(t_ext, std::fmax(hlen(a0), std::fmax(hlen(hadd(v1, T)), hlen(hadd(v2, T)))));
        t_edge = std::fmax(t_edge, std::fmax(hlen(e1), hlen(e2)));
        ++k;
    };
    for (int64_t i = 0; i < d->n_triangles; ++i) {
        const rte_triangle& t = d->triangles[i];
        if (!finite3(t.v0) || !finite3(t.v1) || !finite3(t.v2) || !finite3(t.translate)) return fail(ctx, RTE_ERR_INVALID, "triangle %lld is not finite", (long long)i);
        add_tri(h3(t.v0), h3(t.v1), h3(t.v2), h3(t.translate), t.material, (int)i);
    }
    for (int64_t m = 0; m < d->n_models; ++m) {
        const rte_model& mo = d->models[m];
        const double* vb = d->model_vertices + 3 * mo.first_vertex;
        const int32_t* ib = d->model_indices + mo.first_index;
        if (!finite3(mo.translate)) return fail(ctx, RTE_ERR_INVALID, "model %lld translate is not finite", (long long)m);
        for (int64_t i = 0; i + 2 < mo.index_count; i += 3) {
            for (int c = 0; c < 3; ++c) {
                if (ib[i + c] < 0 || ib[i + c] >= mo.vertex_count)
                    return fail(ctx, RTE_ERR_INVALID, "model %lld: index %d out of range [0,%lld)", (long long)m, ib[i + c], (long long)mo.vertex_count);
                if (!finite3(vb + 3 * ib[i + c])) return fail(ctx, RTE_ERR_INVALID, "model %lld has a non-finite vertex", (long long)m);
            }
            add_tri(h3(vb + 3 * ib[i]), h3(vb + 3 * ib[i + 1]), h3(vb + 3 * ib[i + 2]), h3(mo.translate), mo.material, (int)m);
        }
    }
    const float qnan = std::nanf("");
    for (int i = nt; i < nt_pad; ++i) {  // never-candidate padding: every filter comparison is false on NaN
        for (int c = 0; c < 3; ++c) tri32[3 * (size_t)i + c] = make_float4(qnan, qnan, qnan, 0.f);
        std::memset(&tri64[i], 0, sizeof(DevTri64));
    }

    // lights — E = color * intensity (RE/Scene.h:110)
    const int nl = (int)d->n_lights;
    std::vector<DevLight> lights((size_t)nl);
    for (int i = 0; i < nl; ++i) {
        put3(lights[i].pos, h3(d->lights[i].position));
        put3(lights[i].E, hmul(h3(d->lights[i].color), d->lights[i].intensity));
    }

    int rc;
    const int nt_pad32 = nt > 96 ? (nt + 63) & ~63 : (nt + 31) & ~31;  // like ns_pad
    if ((rc = reserve(ctx, ctx->b_views, (size_t)(1 + nl) * (size_t)ns_pad * 16))) return rc;
    if ((rc = reserve(ctx, ctx->b_tviews, (size_t)(1 + nl) * (size_t)nt_pad32 * 16))) return rc;
    if ((rc = reserve(ctx, ctx->b_views_cold, (size_t)(1 + nl) * (size_t)ns_pad * 16))) return rc;
    if ((rc = reserve(ctx, ctx->b_tviews_cold, (size_t)(1 + nl) * (size_t)nt_pad32 * 16))) return rc;
    if ((rc = stage_begin(ctx, stage_bytes(sph32) + stage_bytes(sph64) + stage_bytes(sphhl) + stage_bytes(sphmat) + stage_bytes(planes) + stage_bytes(tri32) + stage_bytes(tri64) + stage_bytes(triiv) +
                               stage_bytes(triinfo) + stage_bytes(mats) + stage_bytes(lights))))
        return rc;
    if ((rc = upload(ctx, ctx->b_sph32, sph32)) || (rc = upload(ctx, ctx->b_sph64, sph64)) || (rc = upload(ctx, ctx->b_sphhl, sphhl)) || (rc = upload(ctx, ctx->b_sphmat, sphmat)) ||
        (rc = upload(ctx, ctx->b_planes, planes)) || (rc = upload(ctx, ctx->b_tri32, tri32)) || (rc = upload(ctx, ctx->b_tri64, tri64)) || (rc = upload(ctx, ctx->b_triiv, triiv)) ||
        (rc = upload(ctx, ctx->b_triinfo, triinfo)) || (rc = upload(ctx, ctx->b_mats, mats)) || (rc = upload(ctx, ctx->b_lights, lights)))
        return rc;
    RTE_CUDA(ctx, cudaEventRecord(ctx->ev_stage, ctx->stream));  // (the copies read the staging buffer, not the host vectors: no sync)

    DevScene& sc = ctx->sc;
    sc.sph32 = (const float4*)ctx->b_sph32.p; sc.tri32 = (const float4*)ctx->b_tri32.p;
    sc.views = ns > 0 ? (const float*)ctx->b_views.p : nullptr;
    sc.tri_views = nt > 0 ? (const float*)ctx->b_tviews.p : nullptr;
    sc.views_cold = (const float*)ctx->b_views_cold.p; sc.tri_views_cold = (const float*)ctx->b_tviews_cold.p;
    sc.reflective = 0; sc.nt_pad32 = nt_pad32;
    sc.transparent = 0;
    for (int64_t i = 0; i < d->n_materials; ++i) {  // (the default Material appended at index nm is opaque and diffuse)
        if (mats[i].specular > 0.0 || mats[i].transparency > 0.0) sc.reflective = 1;  // secondary rays will be traced
        if (mats[i].transparency > 0.0) sc.transparent = 1;
    }
    ctx->views_dirty = true;
    ctx->grid_dirty = true;
    sc.ns = ns; sc.ns_pad = ns_pad; sc.nt = nt; sc.nt_pad = nt_pad;
    sc.sph64 = (const double4*)ctx->b_sph64.p; sc.sph_hl = (const float4*)ctx->b_sphhl.p; sc.sph_mat = (const int32_t*)ctx->b_sphmat.p;
    sc.planes = (const DevPlane*)ctx->b_planes.p; sc.tri64 = (const DevTri64*)ctx->b_tri64.p; sc.tri_iv = (const Tri32*)ctx->b_triiv.p;
    sc.tri_info = (const int2*)ctx->b_triinfo.p; sc.materials = (const rte_material*)ctx->b_mats.p;
    sc.lights = (const DevLight*)ctx->b_lights.p;
    sc.np = np; sc.nm = (int)d->n_materials; sc.nl = nl;
    sc.sph_extent = up((float)s_ext) * 1.000001f;
    sc.tri_extent = up((float)t_ext) * 1.000001f;
    sc.tri_edge = up((float)t_edge) * 1.000001f;
    sc.pad1 = 0.f;
    ctx->has_scene = true;
    return RTE_OK;
}

extern "C" int rte_set_camera(rte_ctx* ctx, const rte_camera* c) {
    RTE_ENTER(ctx);
    if (!c) return fail(ctx, RTE_ERR_INVALID, "rte_set_camera: camera is NULL");
    if (c->width == 0 || c->height == 0 || c->width > (1u << 20) || c->height > (1u << 20))
        return fail(ctx, RTE_ERR_INVALID, "camera size %llux%llu not supported", (unsigned long long)c->width, (unsigned long long)c->height);
    if (!finite3(c->position) || !std::isfinite(c->focal)) return fail(ctx, RTE_ERR_INVALID, "camera is not finite");
    if (c->anti_aliasing < 1 || c->anti_aliasing > RTE_MAX_AA)
        return fail(ctx, RTE_ERR_LIMIT, "anti_aliasing = %d: 1..%d samples per pixel", c->anti_aliasing, RTE_MAX_AA);
    ctx->cam_host = *c;
    DevCamera& dc = ctx->cam;
    dc.pos[0] = c->position[0]; dc.pos[1] = c->position[1]; dc.pos[2] = c->position[2];
    dc.focal = c->focal;
    dc.half_w = (double)c->width / 2.0;   // RE/Math.h:100
    dc.half_h = (double)c->height / 2.0;  // RE/Math.h:101
    dc.width = (int)c->width; dc.height = (int)c->height;
    dc.aa = c->anti_aliasing; dc.jitter_seed = (uint32_t)c->jitter_seed;
    ctx->has_camera = true;
    ctx->views_dirty = true;
    ctx->grid_dirty = true;
    return RTE_OK;
}

// ====================================================================================== render
// Direction grids of the FP32 path: count + scan on the device, then fill.  The entry count is only known on the device:
// the first build of a scene shape (sphere / light count, G) reads it back (one synchronisation) and sizes the buffer at
// twice that; later builds of the same shape (a re-uploaded or moved scene) reuse the buffer without synchronising — the fill
// pass never writes beyond it, the count travels to a pinned word and is checked at the next synchronisation point
// (check_grid_overflow).  RTE_GRID_G (environment, developer knob): cells per cube-face edge.
static int ensure_grid(rte_ctx* ctx, double bias, DevGrid& out) {
    // Finer cells = fewer candidates per ray, but the build grows with G^2 and is paid per frame.  Measured on B200, render +
    // rebuild, C2 / C5(8 lights) at 4K: G = 48: 0.946 / -, 64: 0.928 / 2.62 ms, 96: 0.914 / 2.53, 128: 0.914 / 2.53; one eighth
    // of the rows (one rank of eight): 48: 0.183 / -, 64: 0.190 / 0.458, 96: 0.194 / 0.476, 128: 0.200 / 0.518.
    const size_t px = ctx->grid_call_px;
    int G = px >= (4u << 20) ? 128 : px >= (2u << 20) ? 96 : px >= (768u << 10) ? 64 : 48;
    if (const char* e = std::getenv("RTE_GRID_G")) { const int v = std::atoi(e); if (v >= 4 && v <= 512) G = v; }
    const int n_views = 1 + ctx->sc.nl, cpv = 6 * G * G + 1;
    const size_t n_cells = (size_t)n_views * (size_t)cpv;
    if (ctx->grid_dirty || ctx->grid_bias != bias || ctx->grid_G != G) {
        int rc;
        const size_t aggr_bytes = (size_t)grid_scan_blocks((int)n_cells) * 8;
        const size_t count_cap_was = ctx->b_grid_count.cap;
        if ((rc = reserve(ctx, ctx->b_grid_count, n_cells * 4)) || (rc = reserve(ctx, ctx->b_grid_start, (n_cells + 1) * 4))) return rc;
        if (ctx->b_grid_count.cap != count_cap_was) {  // the counters are all zero between builds: the fill pass counts them back down
            RTE_CUDA(ctx, cudaMemsetAsync(ctx->b_grid_count.p, 0, ctx->b_grid_count.cap, ctx->stream));
        }
        if (ctx->b_grid_aggr.cap < aggr_bytes) {
            if ((rc = reserve(ctx, ctx->b_grid_aggr, aggr_bytes))) return rc;
            RTE_CUDA(ctx, cudaMemsetAsync(ctx->b_grid_aggr.p, 0, ctx->b_grid_aggr.cap, ctx->stream));  // (epoch 0 is never used)
        }
        ctx->grid_epoch = ctx->grid_epoch == 0xFFFFFFFFu ? 1u : ctx->grid_epoch + 1u;
        RTE_CUDA(ctx, launch_grid_count_scan(ctx->sc, ctx->cam, bias, G, (uint32_t*)ctx->b_grid_count.p, (uint32_t*)ctx->b_grid_start.p,
                                             (unsigned long long*)ctx->b_grid_aggr.p, ctx->grid_epoch, ctx->h_grid_total, ctx->stream));
        const long long shape = ((long long)(ctx->sc.ns + ctx->sc.nt) << 24) ^ ((long long)ctx->sc.nl << 12) ^ (long long)G;
        if (ctx->grid_shape != shape || ctx->b_grid_entries.cap == 0) {
            RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            const size_t total = *ctx->h_grid_total;
            if ((rc = reserve(ctx, ctx->b_grid_entries, (2 * total + 1024) * 4))) return rc;
            ctx->grid_shape = shape;
            ctx->grid_check_pending = false;
        } else {
            ctx->grid_check_pending = true;
        }
        RTE_CUDA(ctx, launch_grid_fill(ctx->sc, ctx->cam, bias, G, (uint32_t*)ctx->b_grid_count.p, (uint32_t*)ctx->b_grid_start.p,
                                       (uint32_t*)ctx->b_grid_entries.p, (uint32_t)(ctx->b_grid_entries.cap / 4), ctx->stream));
        ctx->grid_dirty = false; ctx->grid_bias = bias; ctx->grid_G = G;
        ctx->render_launches += 3;
    }
    out.cell_start = (const uint32_t*)ctx->b_grid_start.p;
    out.entries = (const uint32_t*)ctx->b_grid_entries.p;
    out.G = G; out.cells_per_view = cpv;
    return RTE_OK;
}
// After a stream synchronisation: did the last unsynchronised grid build fit its buffer?  If not, the frames rendered with it
// are invalid: the next build sizes the buffer anew, and the caller is told (rte_render_host renders again by itself).
static int check_grid_overflow(rte_ctx* ctx) {
    if (!ctx->grid_check_pending) return RTE_OK;
    ctx->grid_check_pending = false;
    if ((size_t)*ctx->h_grid_total * 4 <= ctx->b_grid_entries.cap) return RTE_OK;
    ctx->grid_dirty = true; ctx->grid_shape = -1;
    return fail(ctx, RTE_ERR_LIMIT, "the direction grid of the last render outgrew its buffer (%u entries): that frame is incomplete, render it again",
                *ctx->h_grid_total);
}

// `first` / `last`: one API call may render its rows as several bands (rte_render_host); the statistics and the
// ev_r0..ev_r1 interval then cover all of them.
static int do_render(rte_ctx* ctx, const Defaults& o, const rte_frame* frame, bool first = true, bool last = true) {
    RenderParams p{};
    p.sc = ctx->sc; p.cam = ctx->cam;
    const size_t smem_budget = ctx->smem_optin > 4096 ? ctx->smem_optin - 4096 : 0;  // static smem + headroom
    p.plan = plan_smem(ctx->sc, smem_budget);
    if (first) ctx->render_launches = 0;
    p.row0 = o.row0; p.row1 = o.row1; p.n_rows = o.n_rows; p.stripe_rows = o.stripe_rows; p.stripe_period = o.stripe_period; p.max_depth = o.max_depth; p.tonemap_op = o.tonemap_op;
    p.exact_shadows = (o.flags & RTE_FLAG_EXACT_SHADOWS) ? 1 : 0;
    p.no_filter = (o.flags & RTE_FLAG_NO_FILTER) ? 1 : 0;
    p.force_general = (o.flags & RTE_FLAG_GENERAL_KERNEL) ? 1 : 0;
    p.no_cull = (o.flags & RTE_FLAG_NO_CULL) ? 1 : 0;
    p.bias = o.bias; p.max_white = o.max_white; p.jodie_a = o.jodie_a;
    p.out = *frame;
    p.pass_mode = (o.flags & RTE_FLAG_SINGLE_PASS) ? 1 : (o.flags & RTE_FLAG_TWO_PASS) ? 2 : 0;
    p.exact_fp64 = (o.flags & RTE_FLAG_EXACT_FP64) ? 1 : 0;
    p.exact_grid = (o.flags & RTE_FLAG_EXACT_GRID) ? 1 : 0;
    p.tile_counter = ctx->d_tile_counter; p.stats = ctx->d_stats;
    if (p.out.rgb8 && p.tonemap_op == RTE_TM_NONE) return fail(ctx, RTE_ERR_INVALID, "frame.rgb8 requested but opts.tonemap_op is RTE_TM_NONE");
    // FP32 production path (rte_render32.cu): diffuse sphere / plane scenes with resident views, outputs hdr / rgb8 / prim_id /
    // light_mask.  The operator runs in the kernel; pixels with an inconclusive FP32 decision are rendered again, exactly, by fix_batch in the same kernel.
    if (o.n_rows > 0 && fp32_path_eligible(p)) {
        if (p.out.rgb8 && ((uintptr_t)p.out.rgb8 & 3u) != 0 && (ctx->cam.width & 3) == 0) return fail(ctx, RTE_ERR_INVALID, "frame.rgb8 must be 4-byte aligned");
        int rc;
        if ((rc = ensure_grid(ctx, o.bias, p.grid))) return rc;
        // multi-GPU output: the kernel writes RGB8 to a band buffer (the caller's, or scratch); the exchange launch behind it
        // copies the band into every registered frame and signals the fence
        Tonemap64Params xp{};
        xp.n_peers = p.tonemap_op == RTE_TM_NONE ? 0 : ctx->n_peers;
        if (xp.n_peers > 0) {
            const size_t n_px = (size_t)o.n_rows * (size_t)ctx->cam.width;
            if (!p.out.rgb8) {
                if ((rc = reserve(ctx, ctx->s_rgb8, n_px * 3 + 16))) return rc;
                p.out.rgb8 = (uint8_t*)ctx->s_rgb8.p;
            }
            xp.rgb8 = p.out.rgb8; xp.n_pixels = (int64_t)n_px;
            for (int i = 0; i < xp.n_peers; ++i) xp.peer_rgb8[i] = ctx->peer_rgb8[i];
            xp.W = ctx->cam.width; xp.row0 = o.row0; xp.stripe_rows = o.stripe_rows; xp.stripe_period = o.stripe_period;
            if (ctx->n_fence > 0) {
                ctx->fence_seq += 1;
                for (int i = 0; i < ctx->n_fence; ++i) xp.fence_flags[i] = ctx->fence_flags[i];
                xp.n_fence = ctx->n_fence; xp.fence_rank = ctx->fence_rank; xp.fence_seq = ctx->fence_seq; xp.fence_counter = ctx->d_fence_counter;
            }
        }
        if (first) {
            RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_stats, 0, sizeof(unsigned long long) * (kStatCount + 2), ctx->stream));
            RTE_CUDA(ctx, cudaEventRecord(ctx->ev_r0, ctx->stream));
        } else {
            RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_tile_counter, 0, sizeof(unsigned long long) * 2, ctx->stream));
        }
        RTE_CUDA(ctx, launch_render32(p, ctx->sm_count, ctx->stream));
        ctx->render_launches += 1;
        if (xp.n_peers > 0) { RTE_CUDA(ctx, launch_exchange_rgb8(xp, ctx->sm_count, ctx->stream)); ctx->render_launches += 1; }
        if (last) {
            RTE_CUDA(ctx, cudaEventRecord(ctx->ev_r1, ctx->stream));
            RTE_CUDA(ctx, cudaMemcpyAsync(ctx->h_stats, ctx->d_stats, sizeof(unsigned long long) * kStatCount, cudaMemcpyDeviceToHost, ctx->stream));
            ctx->render_timed = true;
        }
        return RTE_OK;
    }
    // Reflective scenes with few primitives (rte_render32r.cu): FP32 values, interval decisions over all primitives; the tiles it
    // flags are rendered again by the exact kernel in a second launch (generic filters, no views to build).
    if (o.n_rows > 0 && fp32r_path_eligible(p)) {
        const size_t n_tiles = (size_t)((ctx->cam.width + 7) / 8) * (size_t)((o.n_rows + 3) / 4);
        int rc = reserve(ctx, ctx->s_tileflags, n_tiles);
        if (rc) return rc;
        RTE_CUDA(ctx, cudaMemsetAsync(ctx->s_tileflags.p, 0, n_tiles, ctx->stream));
        if (first) {
            RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_stats, 0, sizeof(unsigned long long) * (kStatCount + 2), ctx->stream));
            RTE_CUDA(ctx, cudaEventRecord(ctx->ev_r0, ctx->stream));
        } else {
            RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_tile_counter, 0, sizeof(unsigned long long) * 2, ctx->stream));
        }
        p.tile_flags = (uint8_t*)ctx->s_tileflags.p;
        RTE_CUDA(ctx, launch_render32r(p, ctx->sm_count, ctx->stream));
        RenderParams q = p;
        q.tile_flags = nullptr; q.tile_mask = (const uint8_t*)ctx->s_tileflags.p;
        q.plan = SmemPlan{};
        q.plan.stage_sph = 1; q.plan.stage_tri = 1;
        q.plan.sph_bytes = (uint32_t)ctx->sc.ns_pad * 16u; q.plan.tri_bytes = (uint32_t)ctx->sc.nt_pad * 48u;
        q.plan.total = (size_t)q.plan.sph_bytes + q.plan.tri_bytes;
        q.force_general = 1;
        RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_tile_counter, 0, sizeof(unsigned int), ctx->stream));
        RTE_CUDA(ctx, launch_render(q, ctx->sm_count, ctx->stream));
        ctx->render_launches += 2;
        if (last) {
            RTE_CUDA(ctx, cudaEventRecord(ctx->ev_r1, ctx->stream));
            RTE_CUDA(ctx, cudaMemcpyAsync(ctx->h_stats, ctx->d_stats, sizeof(unsigned long long) * kStatCount, cudaMemcpyDeviceToHost, ctx->stream));
            ctx->render_timed = true;
        }
        return RTE_OK;
    }
    if (o.n_rows > 0 && exact_grid_wanted(p)) {  // the exact path on the direction grids (any scene size): needs the grids, not the views
        int rc = ensure_grid(ctx, o.bias, p.grid);
        if (rc) return rc;
        p.plan.use_views = 0; p.plan.use_tri_views = 0;
    }
    if (p.pass_mode != 1 && o.n_rows > 0 && two_pass_wanted(p)) {  // nearest hits between the two launches of the fast path (16 B per pixel)
        int rc = reserve(ctx, ctx->s_hits, (size_t)o.n_rows * (size_t)ctx->cam.width * sizeof(double2));
        if (rc) return rc;
        p.hits = (double2*)ctx->s_hits.p;
    }
    if ((p.plan.use_views || p.plan.use_tri_views) && (ctx->views_dirty || ctx->views_bias != o.bias)) {
        RTE_CUDA(ctx, launch_build_views(ctx->sc, ctx->cam, o.bias, ctx->stream));
        ctx->render_launches += (ctx->sc.ns > 0 ? 1 : 0) + ((ctx->sc.nt > 0 && ctx->sc.tri_views) ? 1 : 0);
        ctx->views_dirty = false;
        ctx->views_bias = o.bias;
    }
    // RGB8 (band buffer and / or registered peer frames) = the exact FP64 operator over the FP64 HDR of the render kernel, as
    // a second, HBM-bound pass (rte_tonemap.cu: tonemap64_kernel) — the render kernels themselves only write HDR
    Tonemap64Params tp{};
    tp.n_peers = p.tonemap_op == RTE_TM_NONE ? 0 : ctx->n_peers;
    const bool want_rgb8 = p.tonemap_op != RTE_TM_NONE && (p.out.rgb8 || tp.n_peers > 0) && o.n_rows > 0;
    const bool fence = ctx->n_fence > 0 && tp.n_peers > 0;  // this render delivers a frame to the peers: it also signals them
    if (want_rgb8) {
        const int64_t n_px = (int64_t)o.n_rows * ctx->cam.width;
        if (!p.out.hdr64) {
            int rc = reserve(ctx, ctx->s_tm64, (size_t)n_px * 24);
            if (rc) return rc;
            p.out.hdr64 = (double*)ctx->s_tm64.p;
        }
        if (p.out.rgb8 && ((uintptr_t)p.out.rgb8 & 3u) != 0 && (ctx->cam.width & 3) == 0)
            return fail(ctx, RTE_ERR_INVALID, "frame.rgb8 must be 4-byte aligned");
        tp.hdr64 = p.out.hdr64; tp.n_pixels = n_px; tp.op = p.tonemap_op; tp.max_white = o.max_white; tp.jodie_a = o.jodie_a;
        tp.rgb8 = p.out.rgb8;
        for (int i = 0; i < tp.n_peers; ++i) tp.peer_rgb8[i] = ctx->peer_rgb8[i];
        tp.W = ctx->cam.width; tp.row0 = o.row0; tp.stripe_rows = o.stripe_rows; tp.stripe_period = o.stripe_period;
    }
    p.out.rgb8 = nullptr;
    p.tonemap_op = RTE_TM_NONE;
    if (p.out.depth_norm && !p.out.depth64 && o.n_rows > 0) {  // frame.depth_norm is a pass over the FP64 depth
        int rc = reserve(ctx, ctx->s_dn64, (size_t)o.n_rows * (size_t)ctx->cam.width * 8);
        if (rc) return rc;
        p.out.depth64 = (double*)ctx->s_dn64.p;
    }
    // (d_tile_counter sits right behind d_stats: one memset clears both)
    if (first) {
        RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_stats, 0, sizeof(unsigned long long) * (kStatCount + 2), ctx->stream));
        RTE_CUDA(ctx, cudaEventRecord(ctx->ev_r0, ctx->stream));
    } else {
        RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_tile_counter, 0, sizeof(unsigned int), ctx->stream));
    }
    if (o.n_rows > 0) RTE_CUDA(ctx, launch_render(p, ctx->sm_count, ctx->stream));
    if (o.n_rows > 0) ctx->render_launches += (p.hits && two_pass_wanted(p)) ? 2 : 1;
    if (p.out.depth_norm && o.n_rows > 0) {
        RTE_CUDA(ctx, launch_depth_norm(p.out.depth64, (int64_t)o.n_rows * ctx->cam.width, ctx->cam_host.near_plane, ctx->cam_host.far_plane, p.out.depth_norm,
                                        ctx->sm_count, ctx->stream));
        ctx->render_launches += 1;
    }
    if (fence) {
        ctx->fence_seq += 1;
        for (int i = 0; i < ctx->n_fence; ++i) tp.fence_flags[i] = ctx->fence_flags[i];
        tp.n_fence = ctx->n_fence; tp.fence_rank = ctx->fence_rank; tp.fence_seq = ctx->fence_seq; tp.fence_counter = ctx->d_fence_counter;
        if (!want_rgb8) { tp.n_pixels = 0; tp.rgb8 = nullptr; tp.n_peers = 0; tp.W = ctx->cam.width > 0 ? ctx->cam.width : 1; }  // a rank without rows still signals
    }
    if (want_rgb8 || fence) RTE_CUDA(ctx, launch_tonemap64(tp, ctx->sm_count, ctx->stream));
    if (want_rgb8 || fence) ctx->render_launches += 1;
    if (last) {
        RTE_CUDA(ctx, cudaEventRecord(ctx->ev_r1, ctx->stream));
        RTE_CUDA(ctx, cudaMemcpyAsync(ctx->h_stats, ctx->d_stats, sizeof(unsigned long long) * kStatCount, cudaMemcpyDeviceToHost, ctx->stream));
        ctx->render_timed = true;
    }
    return RTE_OK;
}

extern "C" int rte_render(rte_ctx* ctx, const rte_render_opts* opts, const rte_frame* frame) {
    RTE_ENTER(ctx);
    if (!ctx->has_scene || !ctx->has_camera) return fail(ctx, RTE_ERR_INVALID, "rte_render: upload a scene and set a camera first");
    if (!frame) return fail(ctx, RTE_ERR_INVALID, "rte_render: frame is NULL");
    Defaults o;
    int rc = resolve_opts(ctx, opts, o);
    if (rc) return rc;
    ctx->grid_call_px = (size_t)o.n_rows * (size_t)ctx->cam.width;
    return do_render(ctx, o, frame);
}

extern "C" int rte_render_host(rte_ctx* ctx, const rte_render_opts* opts, const rte_frame* hf) {
    RTE_ENTER(ctx);
    if (!ctx->has_scene || !ctx->has_camera) return fail(ctx, RTE_ERR_INVALID, "rte_render_host: upload a scene and set a camera first");
    if (!hf) return fail(ctx, RTE_ERR_INVALID, "rte_render_host: frame is NULL");
    Defaults o;
    int rc = resolve_opts(ctx, opts, o);
    if (rc) return rc;
    const size_t W = (size_t)ctx->cam.width;
    const size_t n = (size_t)o.n_rows * W;
    ctx->grid_call_px = n;
    rte_frame df{};
    struct Item { void* host; DevBuf* buf; size_t px_bytes; void** dev; };
    Item items[] = {
        {hf->hdr, &ctx->s_hdr, 16, (void**)&df.hdr},          {hf->hdr64, &ctx->s_hdr64, 24, (void**)&df.hdr64},
        {hf->rgb8, &ctx->s_rgb8, 3, (void**)&df.rgb8},         {hf->prim_id, &ctx->s_prim, 4, (void**)&df.prim_id},
        {hf->depth, &ctx->s_depth, 4, (void**)&df.depth},      {hf->depth64, &ctx->s_depth64, 8, (void**)&df.depth64},
        {hf->normal, &ctx->s_normal, 16, (void**)&df.normal},  {hf->albedo, &ctx->s_albedo, 16, (void**)&df.albedo},
        {hf->light_mask, &ctx->s_lmask, 4, (void**)&df.light_mask},  {hf->depth_norm, &ctx->s_dnorm, 4, (void**)&df.depth_norm}};
    for (Item& it : items) {
        if (!it.host) continue;
        if ((rc = reserve(ctx, *it.buf, n * it.px_bytes))) return rc;
        *it.dev = it.buf->p;
    }
    // Row bands: while the outputs of one band travel to the host on the copy stream, the next band is rendered on the context's
    // stream (schedule below).  Bands are whole stripes when the rows are striped; registered peer frames (multi-GPU output exchange) keep one launch.
    // With a pageable destination cudaMemcpyAsync blocks the host until band 0 has been copied, before band 1 is enqueued:
    // the split would only add a launch, so it is used for pinned destinations only.
    bool pinned = true;
    for (Item& it : items) {
        if (!it.host) continue;
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, it.host) != cudaSuccess) { (void)cudaGetLastError(); pinned = false; }
        else if (at.type != cudaMemoryTypeHost) pinned = false;
    }
    // Band schedule in eighths of the rows.  A band's copy should take about as long as the next band's render, and every launch
    // of a persistent render kernel costs ~45 us before and after it is full:
    //   exact kernels (C2 at 4K: 2.8 ms of render against 0.46 ms of copy): 7/8 + 1/8 (measured 1/4 3.19 ms, 1/6 3.13, 1/8 3.07, 1/12 3.11);
    //   FP32 path (0.86 ms of render against the same copy): 5/8 + 2/8 + 1/8 — the model t_render = 0.86 f + 0.045, t_copy = 0.457 f
    //   gives 1.25 ms for 7/8 + 1/8, 1.09 for the best two bands and 1.05 for these three; measured 1.33 -> see DESIGN.md §8.
    const int unit = o.stripe_rows > 0 ? o.stripe_rows : 4;
    int band_rows[kHostBandsMax] = {o.n_rows, 0, 0}, n_bands = 1;
    if (o.n_rows >= kHostBandRows && ctx->n_peers == 0 && pinned) {
        RenderParams q{};
        q.sc = ctx->sc; q.cam = ctx->cam; q.out = df; q.bias = o.bias; q.tonemap_op = o.tonemap_op;
        q.exact_shadows = (o.flags & RTE_FLAG_EXACT_SHADOWS) ? 1 : 0; q.no_filter = (o.flags & RTE_FLAG_NO_FILTER) ? 1 : 0;
        q.force_general = (o.flags & RTE_FLAG_GENERAL_KERNEL) ? 1 : 0; q.no_cull = (o.flags & RTE_FLAG_NO_CULL) ? 1 : 0;
        q.pass_mode = (o.flags & RTE_FLAG_SINGLE_PASS) ? 1 : (o.flags & RTE_FLAG_TWO_PASS) ? 2 : 0;
        q.exact_fp64 = (o.flags & RTE_FLAG_EXACT_FP64) ? 1 : 0;
        const bool fast = fp32_path_eligible(q);
        const int eighths[2][kHostBandsMax] = {{7, 1, 0}, {5, 2, 1}};
        int left = o.n_rows;
        n_bands = 0;
        for (int k = 0; k < kHostBandsMax && left > 0; ++k) {
            const int e = eighths[fast ? 1 : 0][k];
            if (e == 0) break;
            const bool final_band = k + 1 == kHostBandsMax || eighths[fast ? 1 : 0][k + 1] == 0;
            int r = final_band ? left : (int)(((long long)o.n_rows * e / 8 + unit - 1) / unit * unit);
            if (r > left) r = left;
            band_rows[n_bands++] = r;
            left -= r;
        }
        if (left > 0) band_rows[n_bands - 1] += left;
    }
    RTE_CUDA(ctx, cudaEventRecord(ctx->ev_copied, ctx->stream));
    RTE_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_copied, 0));  // the copy stream starts behind earlier work of the context
    for (int b = 0, y0 = 0; b < n_bands; y0 += band_rows[b], ++b) {
        Defaults ob = o;
        ob.n_rows = band_rows[b];
        if (o.stripe_rows > 0) ob.row0 = o.row0 + (y0 / o.stripe_rows) * o.stripe_period;  // y0 is a multiple of stripe_rows
        else { ob.row0 = o.row0 + y0; ob.row1 = ob.row0 + ob.n_rows; }
        const size_t off = (size_t)y0 * W;
        rte_frame bf{};
        bf.hdr = df.hdr ? df.hdr + off * 4 : nullptr;           bf.hdr64 = df.hdr64 ? df.hdr64 + off * 3 : nullptr;
        bf.rgb8 = df.rgb8 ? df.rgb8 + off * 3 : nullptr;        bf.prim_id = df.prim_id ? df.prim_id + off : nullptr;
        bf.depth = df.depth ? df.depth + off : nullptr;         bf.depth64 = df.depth64 ? df.depth64 + off : nullptr;
        bf.normal = df.normal ? df.normal + off * 4 : nullptr;  bf.albedo = df.albedo ? df.albedo + off * 4 : nullptr;
        bf.light_mask = df.light_mask ? df.light_mask + off : nullptr;  bf.depth_norm = df.depth_norm ? df.depth_norm + off : nullptr;
        const bool last = b + 1 == n_bands;
        if ((rc = do_render(ctx, ob, &bf, b == 0, last))) return rc;
        RTE_CUDA(ctx, cudaEventRecord(ctx->ev_band[b], ctx->stream));
        RTE_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_band[b], 0));
        for (Item& it : items) {
            const size_t bytes = (size_t)ob.n_rows * W * it.px_bytes;
            if (it.host && bytes)
                RTE_CUDA(ctx, cudaMemcpyAsync((char*)it.host + off * it.px_bytes, (char*)it.buf->p + off * it.px_bytes, bytes, cudaMemcpyDeviceToHost, ctx->copy_stream));
        }
    }
    RTE_CUDA(ctx, cudaEventRecord(ctx->ev_copied, ctx->copy_stream));
    RTE_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copied, 0));  // later work of the context must not overwrite the staging buffers
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (int rc = check_grid_overflow(ctx)) {  // the reused grid buffer was too small for this scene: it is sized anew, the frame rendered again
        static thread_local int depth = 0;
        if (depth == 0) { ++depth; rc = rte_render_host(ctx, opts, hf); --depth; }
        return rc;
    }
    return RTE_OK;
}

extern "C" int rte_last_render_launches(const rte_ctx* ctx) { return ctx ? ctx->render_launches : -1; }

extern "C" int rte_get_stats(rte_ctx* ctx, rte_stats* out) {
    RTE_ENTER(ctx);
    if (!out) return fail(ctx, RTE_ERR_INVALID, "rte_get_stats: out is NULL");
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (int rc = check_grid_overflow(ctx)) return rc;
    if (int rc = check_fence_timeout(ctx)) return rc;
    rte_stats& s = ctx->stats;
    if (ctx->render_timed) {
        const unsigned long long* h = ctx->h_stats;
        s.primary_rays = h[kStatPrimary]; s.shadow_rays = h[kStatShadow]; s.secondary_rays = h[kStatSecondary];
        s.sphere_tests = h[kStatSphTests]; s.plane_tests = h[kStatPlTests]; s.triangle_tests = h[kStatTriTests];
        s.exact_marches = h[kStatMarches]; s.fp64_confirms = h[kStatConfirms];
        s.sphere_tests_executed = h[kStatSphExec];
        float ms = 0.f;
        RTE_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_r0, ctx->ev_r1));
        s.kernel_ms = ms;
    }
    if (ctx->tonemap_timed) {
        float ms = 0.f;
        RTE_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_t0, ctx->ev_t1));
        s.tonemap_ms = ms;
    }
    *out = s;
    return RTE_OK;
}

// ====================================================================================== tonemap
static int check_tonemap_args(rte_ctx* ctx, const float* hdr, int64_t n, uint32_t mask, uint8_t* const rgb8[RTE_TM_COUNT]) {
    if (n < 0) return fail(ctx, RTE_ERR_INVALID, "n_pixels is negative");
    if (mask == 0 || mask >= (1u << RTE_TM_COUNT)) return fail(ctx, RTE_ERR_INVALID, "op_mask 0x%x selects no / unknown operators", mask);
    if (n > 0 && !hdr) return fail(ctx, RTE_ERR_INVALID, "hdr is NULL");
    if (!rgb8) return fail(ctx, RTE_ERR_INVALID, "rgb8 is NULL");
    for (int op = 0; op < RTE_TM_COUNT; ++op)
        if ((mask & (1u << op)) && n > 0 && !rgb8[op]) return fail(ctx, RTE_ERR_INVALID, "rgb8[%d] is NULL but its bit is set in op_mask", op);
    return RTE_OK;
}

static int do_tonemap(rte_ctx* ctx, const float* hdr, int64_t n, uint32_t mask, double max_white, double jodie_a, uint8_t* const rgb8[RTE_TM_COUNT],
                      const float* max_white_dev = nullptr) {
    TonemapParams p{};
    p.hdr = (const float4*)hdr; p.n_pixels = n; p.op_mask = mask; p.max_white_dev = max_white_dev;
    p.max_white = (float)(max_white > 0 ? max_white : 5.0);
    p.jodie_a = (float)(jodie_a > 0 ? jodie_a : 0.18);
    if (((uintptr_t)hdr & 15u) != 0) return fail(ctx, RTE_ERR_INVALID, "hdr must be 16-byte aligned");
    for (int op = 0; op < RTE_TM_COUNT; ++op) {
        p.rgb8[op] = (mask & (1u << op)) ? rgb8[op] : nullptr;
        if (p.rgb8[op] && ((uintptr_t)p.rgb8[op] & 3u) != 0) return fail(ctx, RTE_ERR_INVALID, "rgb8[%d] must be 4-byte aligned", op);
    }
    RTE_CUDA(ctx, cudaEventRecord(ctx->ev_t0, ctx->stream));
    RTE_CUDA(ctx, launch_tonemap(p, ctx->sm_count, ctx->stream));
    RTE_CUDA(ctx, cudaEventRecord(ctx->ev_t1, ctx->stream));
    ctx->tonemap_timed = true;
    return RTE_OK;
}

extern "C" int rte_tonemap(rte_ctx* ctx, const float* hdr, int64_t n_pixels, uint32_t op_mask, double max_white, double jodie_a,
                           uint8_t* const rgb8[RTE_TM_COUNT]) {
    RTE_ENTER(ctx);
    int rc = check_tonemap_args(ctx, hdr, n_pixels, op_mask, rgb8);
    if (rc) return rc;
    return do_tonemap(ctx, hdr, n_pixels, op_mask, max_white, jodie_a, rgb8);
}

extern "C" int rte_tonemap_auto_white(rte_ctx* ctx, const float* hdr, int64_t n_pixels, uint32_t op_mask, const float* max_white_dev, double jodie_a,
                                      uint8_t* const rgb8[RTE_TM_COUNT]) {
    RTE_ENTER(ctx);
    if (!max_white_dev) return fail(ctx, RTE_ERR_INVALID, "rte_tonemap_auto_white: max_white_dev is NULL");
    int rc = check_tonemap_args(ctx, hdr, n_pixels, op_mask, rgb8);
    if (rc) return rc;
    return do_tonemap(ctx, hdr, n_pixels, op_mask, 5.0, jodie_a, rgb8, max_white_dev);
}

extern "C" int rte_tonemap_host(rte_ctx* ctx, const float* hdr, int64_t n, uint32_t mask, double max_white, double jodie_a,
                                uint8_t* const rgb8[RTE_TM_COUNT]) {
    RTE_ENTER(ctx);
    int rc = check_tonemap_args(ctx, hdr, n, mask, rgb8);
    if (rc) return rc;
    if (n == 0) return RTE_OK;
    if ((rc = reserve(ctx, ctx->s_tm_in, (size_t)n * 16))) return rc;
    uint8_t* dev[RTE_TM_COUNT] = {};
    for (int op = 0; op < RTE_TM_COUNT; ++op) {
        if (!(mask & (1u << op))) continue;
        if ((rc = reserve(ctx, ctx->s_tm_out[op], (size_t)n * 3 + 16))) return rc;
        dev[op] = (uint8_t*)ctx->s_tm_out[op].p;
    }
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_tm_in.p, hdr, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = do_tonemap(ctx, (const float*)ctx->s_tm_in.p, n, mask, max_white, jodie_a, dev))) return rc;
    for (int op = 0; op < RTE_TM_COUNT; ++op)
        if (dev[op]) RTE_CUDA(ctx, cudaMemcpyAsync(rgb8[op], dev[op], (size_t)n * 3, cudaMemcpyDeviceToHost, ctx->stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTE_OK;
}

// tonemap() / tonemapAll() over the FP64 HDR the reference's functions take (std::vector<Vec3>): the exact FP64 operators
// (tm64), byte-identical to the reference's output.
extern "C" int rte_tonemap_exact_host(rte_ctx* ctx, const double* hdr64, int64_t n, uint32_t mask, double max_white, double jodie_a,
                                      uint8_t* const rgb8[RTE_TM_COUNT]) {
    RTE_ENTER(ctx);
    if (n < 0) return fail(ctx, RTE_ERR_INVALID, "n_pixels is negative");
    if (mask == 0 || mask >= (1u << RTE_TM_COUNT)) return fail(ctx, RTE_ERR_INVALID, "op_mask 0x%x selects no / unknown operators", mask);
    if ((n > 0 && !hdr64) || !rgb8) return fail(ctx, RTE_ERR_INVALID, "rte_tonemap_exact_host: NULL argument");
    for (int op = 0; op < RTE_TM_COUNT; ++op)
        if ((mask & (1u << op)) && n > 0 && !rgb8[op]) return fail(ctx, RTE_ERR_INVALID, "rgb8[%d] is NULL but its bit is set in op_mask", op);
    if (n == 0) return RTE_OK;
    int rc;
    if ((rc = reserve(ctx, ctx->s_hdr64, (size_t)n * 24))) return rc;
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_hdr64.p, hdr64, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
    const double mw = max_white > 0 ? max_white : 5.0, ja = jodie_a > 0 ? jodie_a : 0.18;
    for (int op = 0; op < RTE_TM_COUNT; ++op) {
        if (!(mask & (1u << op))) continue;
        if ((rc = reserve(ctx, ctx->s_tm_out[op], (size_t)n * 3 + 16))) return rc;
        RTE_CUDA(ctx, launch_tonemap64_values((const double*)ctx->s_hdr64.p, n, op, mw, ja, nullptr, (uint8_t*)ctx->s_tm_out[op].p, ctx->sm_count, ctx->stream));
        RTE_CUDA(ctx, cudaMemcpyAsync(rgb8[op], ctx->s_tm_out[op].p, (size_t)n * 3, cudaMemcpyDeviceToHost, ctx->stream));
    }
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTE_OK;
}

// The operator functions themselves (simple ... aces_approx, RE/RaytracingEngine.cpp:89-163) on FP64 values: Vec3 -> Vec3.
extern "C" int rte_tonemap_values_host(rte_ctx* ctx, int op, double max_white, double jodie_a, const double* in3, int64_t n, double* out3) {
    RTE_ENTER(ctx);
    if (op < 0 || op >= RTE_TM_COUNT) return fail(ctx, RTE_ERR_INVALID, "tonemap op %d is not an RTE_TM_* value", op);
    if (n < 0 || (n > 0 && (!in3 || !out3))) return fail(ctx, RTE_ERR_INVALID, "rte_tonemap_values_host: bad arguments");
    if (n == 0) return RTE_OK;
    int rc;
    if ((rc = reserve(ctx, ctx->s_hdr64, (size_t)n * 24)) || (rc = reserve(ctx, ctx->s_tm64, (size_t)n * 24))) return rc;
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_hdr64.p, in3, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
    RTE_CUDA(ctx, launch_tonemap64_values((const double*)ctx->s_hdr64.p, n, op, max_white > 0 ? max_white : 5.0, jodie_a > 0 ? jodie_a : 0.18,
                                          (double*)ctx->s_tm64.p, nullptr, ctx->sm_count, ctx->stream));
    RTE_CUDA(ctx, cudaMemcpyAsync(out3, ctx->s_tm64.p, (size_t)n * 24, cudaMemcpyDeviceToHost, ctx->stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTE_OK;
}

extern "C" int rte_max_luminance(rte_ctx* ctx, const float* hdr, int64_t n_pixels, float* out_max) {
    RTE_ENTER(ctx);
    if (!out_max || (n_pixels > 0 && !hdr) || n_pixels < 0) return fail(ctx, RTE_ERR_INVALID, "rte_max_luminance: bad arguments");
    RTE_CUDA(ctx, launch_max_luminance((const float4*)hdr, n_pixels, ctx->d_maxlum, ctx->sm_count, ctx->stream));
    unsigned int* h = (unsigned int*)&ctx->h_stats[kStatCount];
    RTE_CUDA(ctx, cudaMemcpyAsync(h, ctx->d_maxlum, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    std::memcpy(out_max, h, sizeof(float));
    return RTE_OK;
}

extern "C" int rte_max_luminance_device(rte_ctx* ctx, const float* hdr, int64_t n_pixels, float* out_max_dev) {
    RTE_ENTER(ctx);
    if (!out_max_dev || (n_pixels > 0 && !hdr) || n_pixels < 0) return fail(ctx, RTE_ERR_INVALID, "rte_max_luminance_device: bad arguments");
    if (((uintptr_t)out_max_dev & 3u) != 0) return fail(ctx, RTE_ERR_INVALID, "rte_max_luminance_device: out_max_dev must be 4-byte aligned");
    RTE_CUDA(ctx, launch_max_luminance((const float4*)hdr, n_pixels, (unsigned int*)out_max_dev, ctx->sm_count, ctx->stream));
    return RTE_OK;
}

// No counterpart in the reference (see include/rte.h): max luminance with no
// allocation of any kind in the call — the reduction's scratch is a static __device__ array of rte_tonemap.cu.
extern "C" int rte_max_luminance_noalloc(rte_ctx* ctx, const float* hdr, int64_t n_pixels, float* out_max_dev) {
    RTE_ENTER(ctx);
    if (!out_max_dev || (n_pixels > 0 && !hdr) || n_pixels < 0) return fail(ctx, RTE_ERR_INVALID, "rte_max_luminance_noalloc: bad arguments");
    if (((uintptr_t)out_max_dev & 3u) != 0 || ((uintptr_t)hdr & 15u) != 0)
        return fail(ctx, RTE_ERR_INVALID, "rte_max_luminance_noalloc: hdr must be 16-byte and out_max_dev 4-byte aligned");
    if (n_pixels > (int64_t)RTE_NOALLOC_MAX_PIXELS)
        return fail(ctx, RTE_ERR_LIMIT, "rte_max_luminance_noalloc: %lld pixels exceed RTE_NOALLOC_MAX_PIXELS = %lld (the static scratch array holds one partial per 4096 pixels); use rte_max_luminance_device",
                    (long long)n_pixels, (long long)RTE_NOALLOC_MAX_PIXELS);
    const cudaError_t e = launch_max_luminance_noalloc((const float4*)hdr, n_pixels, out_max_dev, ctx->device, ctx->stream);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, RTE_ERR_CUDA, "rte_max_luminance_noalloc: %s (%s)", cudaGetErrorString(e), cudaGetErrorName(e)); }
    return RTE_OK;
}

// ====================================================================================== FP32 peak
static int measure_peak(rte_ctx* ctx, double target_ms, double* out_tflops, double* out_ms, int packed);
extern "C" int rte_measure_fp32_peak(rte_ctx* ctx, double target_ms, double* out_tflops, double* out_ms) {
    return measure_peak(ctx, target_ms, out_tflops, out_ms, 0);
}
extern "C" int rte_measure_fp32x2_peak(rte_ctx* ctx, double target_ms, double* out_tflops, double* out_ms) {
    return measure_peak(ctx, target_ms, out_tflops, out_ms, 1);
}
static int measure_peak(rte_ctx* ctx, double target_ms, double* out_tflops, double* out_ms, int packed) {
    RTE_ENTER(ctx);
    if (!out_tflops) return fail(ctx, RTE_ERR_INVALID, "rte_measure_fp32_peak: out_tflops is NULL");
    if (!(target_ms > 0)) target_ms = 20.0;
    int rc = reserve(ctx, ctx->s_q_out, (size_t)ctx->sm_count * 8 * 256 * sizeof(float));
    if (rc) return rc;
    double flop = 0.0, best_tf = 0.0, best_ms = 0.0;
    int outer = 256;
    for (int pass = 0; pass < 4; ++pass) {  // pass 0 calibrates `outer`, the best of the other three is reported
        RTE_CUDA(ctx, cudaEventRecord(ctx->ev_t0, ctx->stream));
        RTE_CUDA(ctx, launch_ffma_peak((float*)ctx->s_q_out.p, outer, ctx->sm_count, ctx->stream, &flop, packed));
        RTE_CUDA(ctx, cudaEventRecord(ctx->ev_t1, ctx->stream));
        RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        float ms = 0.f;
        RTE_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_t0, ctx->ev_t1));
        if (pass == 0) {
            const double scale = target_ms / (ms > 1e-3f ? ms : 1e-3f);
            outer = (int)std::fmin(1.0e6, std::fmax(64.0, outer * scale));
            continue;
        }
        const double tf = flop / (ms * 1e-3) / 1e12;
        if (tf > best_tf) { best_tf = tf; best_ms = ms; }
    }
    ctx->tonemap_timed = false;  // ev_t0/ev_t1 no longer bracket a tonemap launch
    *out_tflops = best_tf;
    if (out_ms) *out_ms = best_ms;
    return RTE_OK;
}

// ====================================================================================== ray queries
extern "C" int rte_intersect_closest(rte_ctx* ctx, int64_t n, const double* origins, const double* directions, rte_hit* hits) {
    RTE_ENTER(ctx);
    if (!ctx->has_scene) return fail(ctx, RTE_ERR_INVALID, "rte_intersect_closest: upload a scene first");
    if (n < 0 || (n > 0 && (!origins || !directions || !hits))) return fail(ctx, RTE_ERR_INVALID, "rte_intersect_closest: bad arguments");
    if (n == 0) return RTE_OK;
    int rc;
    if ((rc = reserve(ctx, ctx->s_q_o, (size_t)n * 24)) || (rc = reserve(ctx, ctx->s_q_d, (size_t)n * 24)) ||
        (rc = reserve(ctx, ctx->s_q_out, (size_t)n * sizeof(rte_hit))))
        return rc;
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_q_o.p, origins, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_q_d.p, directions, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
    RTE_CUDA(ctx, launch_intersect_closest(ctx->sc, n, (const double*)ctx->s_q_o.p, (const double*)ctx->s_q_d.p, (rte_hit*)ctx->s_q_out.p,
                                           ctx->sm_count, ctx->stream));
    RTE_CUDA(ctx, cudaMemcpyAsync(hits, ctx->s_q_out.p, (size_t)n * sizeof(rte_hit), cudaMemcpyDeviceToHost, ctx->stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTE_OK;
}

extern "C" int rte_intersect_any_before(rte_ctx* ctx, int64_t n, const double* origins, const double* directions, const double* max_dist,
                                        uint8_t* out) {
    RTE_ENTER(ctx);
    if (!ctx->has_scene) return fail(ctx, RTE_ERR_INVALID, "rte_intersect_any_before: upload a scene first");
    if (n < 0 || (n > 0 && (!origins || !directions || !max_dist || !out))) return fail(ctx, RTE_ERR_INVALID, "rte_intersect_any_before: bad arguments");
    if (n == 0) return RTE_OK;
    int rc;
    if ((rc = reserve(ctx, ctx->s_q_o, (size_t)n * 24)) || (rc = reserve(ctx, ctx->s_q_d, (size_t)n * 24)) ||
        (rc = reserve(ctx, ctx->s_q_m, (size_t)n * 8)) || (rc = reserve(ctx, ctx->s_q_out, (size_t)n)))
        return rc;
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_q_o.p, origins, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_q_d.p, directions, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
    RTE_CUDA(ctx, cudaMemcpyAsync(ctx->s_q_m.p, max_dist, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    RTE_CUDA(ctx, launch_intersect_any_before(ctx->sc, n, (const double*)ctx->s_q_o.p, (const double*)ctx->s_q_d.p, (const double*)ctx->s_q_m.p,
                                              (uint8_t*)ctx->s_q_out.p, ctx->sm_count, ctx->stream));
    RTE_CUDA(ctx, cudaMemcpyAsync(out, ctx->s_q_out.p, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTE_OK;
}

// ====================================================================================== memory helpers
// ====================================================================================== peer frames
extern "C" int rte_set_rgb8_peers(rte_ctx* ctx, int n_frames, uint8_t* const* frames) {
    RTE_ENTER(ctx);
    if (n_frames < 0 || n_frames > kMaxPeers) return fail(ctx, RTE_ERR_LIMIT, "rte_set_rgb8_peers: %d frames (at most %d)", n_frames, kMaxPeers);
    if (n_frames > 0 && !frames) return fail(ctx, RTE_ERR_INVALID, "rte_set_rgb8_peers: frames is NULL");
    for (int i = 0; i < n_frames; ++i)
        if (!frames[i] || ((uintptr_t)frames[i] & 3u) != 0) return fail(ctx, RTE_ERR_INVALID, "rte_set_rgb8_peers: frame %d is NULL or not 4-byte aligned", i);
    for (int i = 0; i < n_frames; ++i) ctx->peer_rgb8[i] = frames[i];
    ctx->n_peers = n_frames;
    return RTE_OK;
}

extern "C" int rte_set_fence_peers(rte_ctx* ctx, int rank, int n_ranks, uint32_t* const* flags) {
    RTE_ENTER(ctx);
    if (n_ranks < 0 || n_ranks > kMaxPeers) return fail(ctx, RTE_ERR_LIMIT, "rte_set_fence_peers: %d ranks (at most %d)", n_ranks, kMaxPeers);
    if (n_ranks > 0 && (!flags || rank < 0 || rank >= n_ranks)) return fail(ctx, RTE_ERR_INVALID, "rte_set_fence_peers: flags is NULL or rank %d outside [0,%d)", rank, n_ranks);
    for (int i = 0; i < n_ranks; ++i)
        if (!flags[i] || ((uintptr_t)flags[i] & 3u) != 0) return fail(ctx, RTE_ERR_INVALID, "rte_set_fence_peers: flags[%d] is NULL or misaligned", i);
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < n_ranks; ++i) ctx->fence_flags[i] = flags[i];
    ctx->n_fence = n_ranks; ctx->fence_rank = rank; ctx->fence_seq = 0;
    if (n_ranks > 0) {
        RTE_CUDA(ctx, cudaMemsetAsync(flags[rank], 0, sizeof(uint32_t) * kMaxPeers, ctx->stream));
        RTE_CUDA(ctx, cudaMemsetAsync(ctx->d_fence_counter, 0, sizeof(unsigned int), ctx->stream));
        RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return RTE_OK;
}

extern "C" int rte_frame_fence(rte_ctx* ctx, double timeout_ms) {
    RTE_ENTER(ctx);
    if (ctx->n_fence <= 0) return fail(ctx, RTE_ERR_INVALID, "rte_frame_fence: no fence peers registered (rte_set_fence_peers)");
    const double ms = timeout_ms > 0 ? timeout_ms : 2000.0;
    unsigned int* dev_word = nullptr;
    RTE_CUDA(ctx, cudaHostGetDevicePointer((void**)&dev_word, ctx->h_fence_timeout, 0));
    RTE_CUDA(ctx, launch_fence_wait(ctx->fence_flags[ctx->fence_rank], ctx->n_fence, ctx->fence_seq, (unsigned long long)(ms * 1e6), dev_word, ctx->stream));
    return RTE_OK;
}

static_assert(sizeof(cudaIpcMemHandle_t) == RTE_IPC_HANDLE_BYTES, "rte.h: RTE_IPC_HANDLE_BYTES");
extern "C" int rte_ipc_export(rte_ctx* ctx, const void* dev, uint8_t handle[RTE_IPC_HANDLE_BYTES]) {
    RTE_ENTER(ctx);
    if (!dev || !handle) return fail(ctx, RTE_ERR_INVALID, "rte_ipc_export: NULL argument");
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void*>(dev));
    if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, RTE_ERR_CUDA, "cudaIpcGetMemHandle: %s (the pointer must come from rte_malloc)", cudaGetErrorString(e)); }
    std::memcpy(handle, &h, sizeof(h));
    return RTE_OK;
}
extern "C" int rte_ipc_open(rte_ctx* ctx, const uint8_t handle[RTE_IPC_HANDLE_BYTES], void** out_dev) {
    RTE_ENTER(ctx);
    if (!handle || !out_dev) return fail(ctx, RTE_ERR_INVALID, "rte_ipc_open: NULL argument");
    *out_dev = nullptr;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    cudaError_t e = cudaIpcOpenMemHandle(out_dev, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, RTE_ERR_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e)); }
    return RTE_OK;
}
extern "C" int rte_ipc_close(rte_ctx* ctx, void* dev) {
    RTE_ENTER(ctx);
    if (!dev) return RTE_OK;
    for (int i = 0; i < ctx->n_peers; ++i)
        if (ctx->peer_rgb8[i] == dev) return fail(ctx, RTE_ERR_INVALID, "rte_ipc_close: the mapping is still registered with rte_set_rgb8_peers");
    cudaError_t e = cudaIpcCloseMemHandle(dev);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, RTE_ERR_CUDA, "cudaIpcCloseMemHandle: %s", cudaGetErrorString(e)); }
    return RTE_OK;
}

extern "C" int rte_malloc(rte_ctx* ctx, size_t bytes, void** out_dev) {
    RTE_ENTER(ctx);
    if (!out_dev) return fail(ctx, RTE_ERR_INVALID, "rte_malloc: out is NULL");
    *out_dev = nullptr;
    cudaError_t e = cudaMalloc(out_dev, bytes ? bytes : 16);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, RTE_ERR_CUDA, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e)); }
    return RTE_OK;
}
extern "C" int rte_free(rte_ctx* ctx, void* dev) {
    RTE_ENTER(ctx);
    if (dev) RTE_CUDA(ctx, cudaFree(dev));
    return RTE_OK;
}
extern "C" int rte_memcpy_d2h(rte_ctx* ctx, void* host, const void* dev, size_t bytes) {
    RTE_ENTER(ctx);
    if (bytes && (!host || !dev)) return fail(ctx, RTE_ERR_INVALID, "rte_memcpy_d2h: NULL pointer");
    if (bytes) RTE_CUDA(ctx, cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTE_OK;
}
extern "C" int rte_memcpy_h2d(rte_ctx* ctx, void* dev, const void* host, size_t bytes) {
    RTE_ENTER(ctx);
    if (bytes && (!host || !dev)) return fail(ctx, RTE_ERR_INVALID, "rte_memcpy_h2d: NULL pointer");
    if (bytes) RTE_CUDA(ctx, cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    RTE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTE_OK;
}
extern "C" int rte_malloc_host(size_t bytes, void** out_host) {
    if (!out_host) return fail(nullptr, RTE_ERR_INVALID, "rte_malloc_host: out is NULL");
    *out_host = nullptr;
    cudaError_t e = cudaMallocHost(out_host, bytes ? bytes : 16);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(nullptr, e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? RTE_ERR_NO_DEVICE : RTE_ERR_CUDA,
                    "cudaMallocHost(%zu): %s", bytes, cudaGetErrorString(e));
    }
    return RTE_OK;
}
extern "C" int rte_free_host(void* host) {
    if (host && cudaFreeHost(host) != cudaSuccess) { cudaGetLastError(); return fail(nullptr, RTE_ERR_CUDA, "cudaFreeHost failed"); }
    return RTE_OK;
}

// ====================================================================================== PPM
// writePPM, RE/Image.cpp:11-31: "P6\n<W> <H>\n255\n" + W*H RGB triplets; the reference streams one char at
// a time, this is one fwrite from the caller's (pinned) buffer.  Error texts are the reference's.
extern "C" int rte_write_ppm(const char* filename, const uint8_t* rgb8, uint64_t width, uint64_t height) {
    if (!filename || (!rgb8 && width != 0 && height != 0)) return fail(nullptr, RTE_ERR_INVALID, "rte_write_ppm: NULL argument");
    FILE* f = std::fopen(filename, "wb");
    if (!f) return fail(nullptr, RTE_ERR_IO, "Could not open file for writing");
    const size_t n = (size_t)width * (size_t)height * 3;
    bool ok = std::fprintf(f, "P6\n%llu %llu\n255\n", (unsigned long long)width, (unsigned long long)height) > 0;
    ok = ok && (n == 0 || std::fwrite(rgb8, 1, n, f) == n);
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) return fail(nullptr, RTE_ERR_IO, "Error occurred while writing to file");
    return RTE_OK;
}

// Row N4: the reference shells out to `ffmpeg` to turn every .ppm into a .png and to `del` to remove the
// .ppm (RE/RaytracingEngine.cpp:317-327, Windows only).  Here the PNG is written directly: 8-bit RGB, filter 0 on every
// scanline, zlib stream of stored (uncompressed) deflate blocks — self-contained, no zlib, no subprocess.
namespace {
struct PngOut {
    FILE* f;
    bool ok = true;
    uint32_t crc = 0;
    static uint32_t crc_step(uint32_t c, const uint8_t* p, size_t n) {
        static uint32_t table[256];
        static bool init = false;
        if (!init) {
            for (uint32_t i = 0; i < 256; ++i) {
                uint32_t v = i;
                for (int k = 0; k < 8; ++k) v = (v & 1u) ? 0xEDB88320u ^ (v >> 1) : v >> 1;
                table[i] = v;
            }
            init = true;
        }
        for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
        return c;
    }
    void raw(const void* p, size_t n) { ok = ok && (n == 0 || std::fwrite(p, 1, n, f) == n); }
    void be32(uint32_t v) { const uint8_t b[4] = {(uint8_t)(v >> 24), (uint8_t)(v >> 16), (uint8_t)(v >> 8), (uint8_t)v}; raw(b, 4); }
    void chunk_begin(const char* type, uint32_t len) { be32(len); crc = 0xFFFFFFFFu; data(type, 4); }
    void data(const void* p, size_t n) { crc = crc_step(crc, (const uint8_t*)p, n); raw(p, n); }
    void chunk_end() { be32(crc ^ 0xFFFFFFFFu); }
};
}  // namespace

extern "C" int rte_write_png(const char* filename, const uint8_t* rgb8, uint64_t width, uint64_t height) {
    if (!filename || !rgb8) return fail(nullptr, RTE_ERR_INVALID, "rte_write_png: NULL argument");
    if (width == 0 || height == 0 || width > 0x7FFFFFFFull / 4 || height > 0x7FFFFFFFull)
        return fail(nullptr, RTE_ERR_INVALID, "rte_write_png: %llux%llu is not a PNG size", (unsigned long long)width, (unsigned long long)height);
    const uint64_t row = 1 + 3 * width, total = row * height;       // filter byte + RGB per scanline
    const uint64_t n_blocks = (total + 65534) / 65535;
    const uint64_t idat = 2 + total + 5 * n_blocks + 4;              // zlib header, stored blocks, adler32
    if (idat > 0x7FFFFFFFull) return fail(nullptr, RTE_ERR_LIMIT, "rte_write_png: image too large for one IDAT chunk");
    FILE* f = std::fopen(filename, "wb");
    if (!f) return fail(nullptr, RTE_ERR_IO, "Could not open file for writing");
    PngOut o{f};
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    o.raw(sig, 8);
    o.chunk_begin("IHDR", 13);
    {
        const uint8_t h[13] = {(uint8_t)(width >> 24), (uint8_t)(width >> 16), (uint8_t)(width >> 8), (uint8_t)width,
                               (uint8_t)(height >> 24), (uint8_t)(height >> 16), (uint8_t)(height >> 8), (uint8_t)height,
                               8, 2, 0, 0, 0};  // 8 bits, colour type 2 (RGB), deflate, adaptive filtering, no interlace
        o.data(h, 13);
    }
    o.chunk_end();
    o.chunk_begin("IDAT", (uint32_t)idat);
    {
        const uint8_t zh[2] = {0x78, 0x01};
        o.data(zh, 2);
        std::vector<uint8_t> line((size_t)row);
        uint32_t a = 1, b = 0;  // adler32 of the uncompressed stream
        uint64_t done = 0;      // bytes of the stream emitted so far
        uint32_t left_in_block = 0;
        for (uint64_t y = 0; y < height; ++y) {
            line[0] = 0;
            std::memcpy(line.data() + 1, rgb8 + (size_t)(3 * width * y), (size_t)(3 * width));
            for (size_t i = 0; i < line.size(); ++i) { a += line[i]; if (a >= 65521u) a -= 65521u; b += a; if (b >= 65521u) b -= 65521u; }
            size_t off = 0;
            while (off < line.size()) {
                if (left_in_block == 0) {
                    const uint64_t remain = total - done;
                    const uint32_t len = (uint32_t)(remain < 65535 ? remain : 65535);
                    const uint8_t bh[5] = {(uint8_t)(remain <= 65535 ? 1 : 0), (uint8_t)len, (uint8_t)(len >> 8), (uint8_t)~len, (uint8_t)(~len >> 8)};
                    o.data(bh, 5);
                    left_in_block = len;
                }
                const size_t n = std::min<size_t>(line.size() - off, left_in_block);
                o.data(line.data() + off, n);
                off += n; done += n; left_in_block -= (uint32_t)n;
            }
        }
        const uint32_t adler = (b << 16) | a;
        const uint8_t ad[4] = {(uint8_t)(adler >> 24), (uint8_t)(adler >> 16), (uint8_t)(adler >> 8), (uint8_t)adler};
        o.data(ad, 4);
    }
    o.chunk_end();
    o.chunk_begin("IEND", 0);
    o.chunk_end();
    const bool ok = (std::fclose(f) == 0) && o.ok;
    if (!ok) return fail(nullptr, RTE_ERR_IO, "Error occurred while writing to file");
    return RTE_OK;
}
