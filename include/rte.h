/*
 * rte.h — C ABI of the B200-native render path (librte_b200.so).
 *
 * The reference (Sorax5/RaytracingEngine, "RE/" = RaytracingEngine/) has no FFI: its de-facto
 * boundary is the public C++ API of `Scene` plus the free functions of RaytracingEngine.cpp and
 * Image.cpp.  Each entry point below names the reference interface it replaces (file:line).
 * The C++20 host mirror (include/rte/engine.hpp) and the Python binding
 * (raytracingengine_b200/engine.py, struct mirrors in _abi.py) are thin layers over exactly these symbols.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++/torch types cross this boundary;
 *   - every function returns 0 on success, a negative rte_status otherwise; the message is
 *     available from rte_last_error() (the C++ wrapper rethrows it as std::runtime_error, which
 *     is the reference's error convention, RE/RaytracingEngine.cpp:38, RE/Image.cpp:15,27);
 *   - one rte_ctx per GPU, externally synchronised; all work is enqueued on the ctx stream;
 *   - image layout: row-major, idx = y*W + x, row 0 = top (RE/Scene.h:214-216,321-324);
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef RTE_H_
#define RTE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTE_VERSION 1

typedef enum rte_status {
    RTE_OK = 0,
    RTE_ERR_INVALID = -1,   /* bad argument / inconsistent description */
    RTE_ERR_CUDA = -2,      /* CUDA runtime error (sticky for the ctx) */
    RTE_ERR_NO_DEVICE = -3, /* no usable CUDA device: there is no CPU path */
    RTE_ERR_IO = -4,        /* file open / write failure (RE/Image.cpp:14-16,26-28) */
    RTE_ERR_LIMIT = -5,     /* a documented capacity is exceeded (2^30 primitives per class, 4096 lights, RTE_MAX_AA, RTE_MAX_DEPTH, ...) */
    RTE_ERR_TIMEOUT = -6    /* rte_frame_fence: a peer did not deliver its frame in time */
} rte_status;

/* HitType, RE/Shape.h:21-26 */
enum { RTE_HIT_NONE = 0, RTE_HIT_SPHERE = 1, RTE_HIT_PLANE = 2, RTE_HIT_TRIANGLE = 3 };

/* Tonemap operators in tonemapAll order, RE/RaytracingEngine.cpp:205-211,303-311 */
enum {
    RTE_TM_SIMPLE = 0,
    RTE_TM_REINHARD_SIMPLE = 1,
    RTE_TM_REINHARD_EXTENDED = 2,
    RTE_TM_REINHARD_EXTENDED_LUMINANCE = 3,
    RTE_TM_REINHARD_JODIE = 4,
    RTE_TM_UNCHARTED2 = 5,
    RTE_TM_ACES = 6,
    RTE_TM_COUNT = 7,
    RTE_TM_NONE = -1
};

/* ------------------------------------------------------------------------------------------
 * Scene description (host side, FP64 exactly like the reference's structs).
 * ---------------------------------------------------------------------------------------- */

/* Material, RE/Shape.h:13-19 (same field order, 56 bytes) */
typedef struct rte_material {
    double color[3];
    double shininess;        /* default 128 */
    double specular;         /* default 0 */
    double transparency;     /* default 0 */
    double refractive_index; /* default 1 */
} rte_material;

/* Sphere(r, pos, mat), RE/Shape.h:59-70 */
typedef struct rte_sphere {
    double center[3];
    double radius;
    int32_t material; /* index into rte_scene_desc.materials */
    int32_t reserved;
} rte_sphere;

/* Plane(pos, norm, mat), RE/Shape.h:135-147 — `normal` is given un-normalised, the library
 * normalises it in FP64 exactly like the constructor does (RE/Shape.h:142). */
typedef struct rte_plane {
    double point[3];
    double normal[3];
    int32_t material;
    int32_t reserved;
} rte_plane;

/* Triangle(v0,v1,v2,mat,transform), RE/Shape.h:188-200 — only transform.position is ever
 * applied by the reference (tv0..tv2); rotation/scale are ignored there and absent here. */
typedef struct rte_triangle {
    double v0[3], v1[3], v2[3];
    double translate[3];
    int32_t material;
    int32_t reserved;
} rte_triangle;

/* Model(indices, transform, material, vertexPositions), RE/Shape.h:248-256.
 * Vertex positions and indices of all models live in two shared arrays of the scene. */
typedef struct rte_model {
    int64_t first_vertex; /* offset (in vertices) into model_vertices */
    int64_t vertex_count;
    int64_t first_index;  /* offset into model_indices; indices are relative to first_vertex */
    int64_t index_count;  /* multiple of 3 */
    double translate[3];
    int32_t material;
    int32_t reserved;
} rte_model;

/* Light{position,color,intensity}, RE/Light.h:6-12 */
typedef struct rte_light {
    double position[3];
    double color[3];
    double intensity;
} rte_light;

/* Camera(position, focal, width, height, near, far), RE/Math.h:85-97 */
typedef struct rte_camera {
    double position[3];
    double focal;
    uint64_t width, height;
    double near_plane, far_plane;          /* only frame.depth_norm reads them (HitInfo::normalizedDistance) */
    int32_t anti_aliasing; /* Camera::antiAliasingAmount, RE/Math.h:94: samples per pixel, 1..RTE_MAX_AA */
    int32_t jitter_seed;   /* seed of the jitter stream below (upstream seeds from std::random_device, RE/Math.h:109) */
} rte_camera;
/* GeneratePixelAt (RE/Scene.h:283-304): the pixel is the mean of anti_aliasing samples; sample 0 is unjittered, sample
 * s >= 1 adds (jx, jy) in [0,1)^2 to the screen point (RE/Math.h:103-116, where invAA = 1/double(bool) = 1).  Upstream
 * draws the jitter from a thread-local mt19937 seeded by random_device — not reproducible — so this library defines
 * a counter-based stream instead (statistically equivalent, identical on every run / GPU count / tile order):
 *   key = ((uint64)(uint32)jitter_seed << 32) ^ ((uint64)(y*W + x) * 8192 + 2*s + axis),   axis 0 = x, 1 = y
 *   z = key + 0x9E3779B97F4A7C15; z = (z ^ z>>30) * 0xBF58476D1CE4E5B9; z = (z ^ z>>27) * 0x94D049BB133111EB; z ^= z>>31
 *   jitter = (z >> 11) * 2^-53 */
#define RTE_MAX_AA 4096

/* The vectors a reference `Scene` owns (RE/Scene.h:17-23), flattened. */
typedef struct rte_scene_desc {
    const rte_material* materials; int64_t n_materials;
    const rte_sphere* spheres;     int64_t n_spheres;
    const rte_plane* planes;       int64_t n_planes;
    const rte_triangle* triangles; int64_t n_triangles;
    const rte_model* models;       int64_t n_models;
    const double* model_vertices;  int64_t n_model_vertices; /* xyz triples */
    const int32_t* model_indices;  int64_t n_model_indices;
    const rte_light* lights;       int64_t n_lights;
} rte_scene_desc;

/* ------------------------------------------------------------------------------------------
 * Render options / outputs
 * ---------------------------------------------------------------------------------------- */

/* TraceRay's recursion is evaluated on a per-pixel stack of this many frames (the reference's maxRecursion is 10). */
#define RTE_MAX_DEPTH 24
typedef struct rte_render_opts {
    int32_t row0, row1;     /* render rows [row0,row1) of the camera image (row band of one GPU);
                               row1 <= 0 means camera.height */
    int32_t max_depth;      /* Scene::maxRecursion, RE/Scene.h:24; <= 0 means 10; at most RTE_MAX_DEPTH */
    int32_t tonemap_op;     /* RTE_TM_* fused into the render pass (RGB8 output), or RTE_TM_NONE */
    double bias;            /* RE/Scene.h:291; <= 0 means 1e-3 */
    double max_white;       /* reinhardExtended*(…, max_white); <= 0 means 5.0 (RE/RaytracingEngine.cpp:193,195) */
    double jodie_a;         /* reinhardJodie(…, a); <= 0 means 0.18 (RE/RaytracingEngine.cpp:150) */
    int32_t spp_seed;       /* counter-based jitter seed when camera.anti_aliasing > 1 */
    int32_t flags;          /* RTE_FLAG_* (validation switches; 0 = production path) */
    /* Interleaved row stripes (multi-GPU load balance): with stripe_rows > 0 the call
     * renders rows row0 + k*stripe_period + [0, stripe_rows) for k = 0,1,... that lie below row1, packed
     * densely in that order into the output buffers.  stripe_rows must be a multiple of 4 and
     * <= stripe_period.  0 = the contiguous band [row0,row1). */
    int32_t stripe_rows, stripe_period;
} rte_render_opts;

/* Validation switches for rte_render_opts.flags.  Both must leave every output bit-identical:
 *   EXACT_SHADOWS  every shadow ray takes the literal RE/Scene.h:35-77 march instead of the any-hit query;
 *   NO_FILTER      the FP32 filters accept everything, i.e. every primitive is tested in FP64;
 *   GENERAL_KERNEL never take the lean kernel specialised for diffuse scenes without AOVs;
 *   NO_CULL        the lean kernel scans every view in full instead of the per-tile culled copy;
 *   SINGLE_PASS /  the lean kernel traces primary rays and shades in one launch / in two launches (primary hits, then
 *   TWO_PASS       shading); without either flag large frames of heavy scenes take two launches, everything else one.
 *   EXACT_GRID     the exact path of a diffuse scene walks the per-origin direction grids (one pixel per thread, no shared-memory
 *                  staging) even where the view-based kernels apply; it is what scenes too large for those kernels take anyway.
 * And one switch that selects the numerics of the PRODUCTION path:
 *   EXACT_FP64     never take the FP32 production path.  Without it, a call that asks for hdr / rgb8 / prim_id / light_mask
 *                  only, on a diffuse scene (spheres, planes, triangles, models), is shaded in FP32 (hit index and shadow visibility bit-exact — every
 *                  FP32 decision is certain within its error bound or the pixel is re-rendered in FP64 —, HDR to ~1e-6 relative,
 *                  RGB8 within +-1 LSB of the reference: the bars BASELINE.json states).  The same holds for scenes with
 *                  specular (reflecting) or transparent (refracting) materials and at most 32 spheres + triangles — main()'s
 *                  scene —: their chains / trees of secondary rays end once the path weight is below 1e-7, so
 *                  rte_stats counts fewer secondary / shadow rays than the reference traces.  Every FP64 output (hdr64,
 *                  depth64), the other AOVs and every flag above select the exact FP64 path, whose results are the
 *                  reference's own bits. */
enum { RTE_FLAG_EXACT_SHADOWS = 1, RTE_FLAG_NO_FILTER = 2, RTE_FLAG_GENERAL_KERNEL = 4, RTE_FLAG_NO_CULL = 8, RTE_FLAG_SINGLE_PASS = 16, RTE_FLAG_TWO_PASS = 32,
       RTE_FLAG_EXACT_FP64 = 64, RTE_FLAG_EXACT_GRID = 128 };

/* Output buffers of one render call.  Every pointer is optional (NULL = not wanted) and is a
 * DEVICE pointer for rte_render() and a HOST pointer for rte_render_host().  Each buffer holds
 * (row1-row0)*W elements in band-local row-major order. */
typedef struct rte_frame {
    float* hdr;          /* float4 per pixel: linear RGB, w = 0 (RenderImage() result, RE/Scene.h:311) */
    double* hdr64;       /* 3 doubles per pixel, the same values before rounding to FP32 */
    uint8_t* rgb8;       /* 3 bytes per pixel, opts.tonemap_op applied + toColor (RE/RaytracingEngine.cpp:113-121) */
    /* AOVs of the primary hit (the remnants of the reference's depth/normal/colour maps,
     * RE/Scene.h:278-281, RE/Shape.h:40-42; miss values documented in DESIGN.md) */
    uint32_t* prim_id;   /* (type << 30) | index, 0xFFFFFFFF on miss (HitInfo.type/.index, RE/Shape.h:29-31) */
    float* depth;        /* HitInfo.distance, +inf on miss */
    double* depth64;     /* the same in FP64 */
    float* normal;       /* float4: HitInfo.normal (geometric, un-flipped), 0 on miss */
    float* albedo;       /* float4: HitInfo.material.color, sky colour on miss */
    uint32_t* light_mask;/* bit j: light j contributes at the primary hit (transmittance > bias);
                            bit 16+j: a shadow ray was cast to light j (guards RE/Scene.h:89-101) */
    float* depth_norm;   /* HitInfo::normalizedDistance(camera), RE/Shape.h:40-42: (distance - near_plane) / (far_plane -
                            near_plane), evaluated in FP64 like upstream and rounded to FP32; +inf on miss (the depth map
                            main() used to write, RE/Scene.h:278-281) */
} rte_frame;

/* Counters accumulated by the render kernels of the last rte_render*() call. */
typedef struct rte_stats {
    uint64_t primary_rays;
    uint64_t shadow_rays;      /* pixel-light pairs passing the guards at RE/Scene.h:89-101 */
    uint64_t secondary_rays;   /* reflection rays that were intersected (depth >= 1) */
    uint64_t sphere_tests;     /* algorithmic test counts: closest-hit rays test */
    uint64_t plane_tests;      /*   every primitive, shadow rays count canonical-order tests up  */
    uint64_t triangle_tests;   /*   to and including the first occluder                           */
    uint64_t exact_marches;    /* shadow rays resolved by the exact RE/Scene.h:35-77 march */
    uint64_t fp64_confirms;    /* FP64 re-evaluations triggered by the FP32 filter; FP32 production path: pixel samples re-rendered in FP64 */
    double kernel_ms;          /* device time of the render kernel (CUDA events on the ctx stream) */
    double tonemap_ms;         /* device time of the last tonemap kernel */
    uint64_t sphere_tests_executed; /* ray-sphere filter evaluations the kernel really ran (any-hit rays stop
                                       when their whole warp is decided; padding entries included) */
} rte_stats;

typedef struct rte_ctx rte_ctx;

/* Hit record of one ray query: HitInfo, RE/Shape.h:28-34 (material by index). */
typedef struct rte_hit {
    double distance;
    double normal[3];
    double point[3];
    uint32_t prim_id; /* as rte_frame.prim_id */
    int32_t material;
} rte_hit;

/* ------------------------------------------------------------------------------------------
 * Entry points
 * ---------------------------------------------------------------------------------------- */

/* Library/ABI version and availability (no device needed). */
int rte_version(void);
int rte_device_count(void);
/* Message of the last failure on this ctx (or, with ctx == NULL, of the calling thread). */
const char* rte_last_error(const rte_ctx* ctx);

/* One context per GPU.  Fails with RTE_ERR_NO_DEVICE when no sm_100 device is visible. */
int rte_create(int device, rte_ctx** out);
void rte_destroy(rte_ctx* ctx);
/* Use an externally owned cudaStream_t (e.g. torch's current stream); NULL = ctx-private stream. */
int rte_set_stream(rte_ctx* ctx, void* cuda_stream);
int rte_synchronize(rte_ctx* ctx);

/* Replaces Scene::Add{Sphere,Plane,Light,Triangle,Model} (RE/Scene.h:208-212) + the implicit
 * "scene is ready" of RenderImage: validates, flattens models into triangles in canonical
 * order, packs FP32 SoA filter arrays + FP64 exact arrays and copies them to the device.
 * The description is copied; the caller keeps ownership of its arrays. */
int rte_upload_scene(rte_ctx* ctx, const rte_scene_desc* scene);
/* explicit Scene(const Camera&), RE/Scene.h:200 */
int rte_set_camera(rte_ctx* ctx, const rte_camera* camera);

/* Scene::RenderImage(), RE/Scene.h:311-328 (+ fused tonemap()/toColor, RE/RaytracingEngine.cpp:165-174).
 * `frame` holds device pointers. Asynchronous on the ctx stream. */
int rte_render(rte_ctx* ctx, const rte_render_opts* opts, const rte_frame* frame);
/* Same with host buffers: device staging + D2H copies inside the call, synchronous.  This is the
 * call the end-to-end number is measured through. */
int rte_render_host(rte_ctx* ctx, const rte_render_opts* opts, const rte_frame* host_frame);
int rte_get_stats(rte_ctx* ctx, rte_stats* out);
/* Kernel launches of this library that the last rte_render / rte_render_host call enqueued (view build, render
 * launches, RGB8 pass), for callers that account for them (bench.py gpu_launches); negative on error. */
int rte_last_render_launches(const rte_ctx* ctx);

/* tonemap() / tonemapAll(), RE/RaytracingEngine.cpp:165-214: `hdr` = float4 per pixel (device),
 * rgb8[op] = 3 bytes per pixel (device) for every op whose bit is set in op_mask. */
int rte_tonemap(rte_ctx* ctx, const float* hdr, int64_t n_pixels, uint32_t op_mask,
                double max_white, double jodie_a, uint8_t* const rgb8[RTE_TM_COUNT]);
int rte_tonemap_host(rte_ctx* ctx, const float* hdr, int64_t n_pixels, uint32_t op_mask,
                     double max_white, double jodie_a, uint8_t* const rgb8[RTE_TM_COUNT]);
/* tonemap() / tonemapAll() over the FP64 HDR the reference's functions take (std::vector<Vec3>, RE/RaytracingEngine.cpp:165,176):
 * hdr64 = 3 doubles per pixel (HOST), rgb8[op] = 3 bytes per pixel (HOST).  The exact FP64 operators in the reference's
 * evaluation order: the bytes are the reference's own (rte_tonemap / rte_tonemap_host are the fast FP32 pass, +-1 LSB). */
int rte_tonemap_exact_host(rte_ctx* ctx, const double* hdr64, int64_t n_pixels, uint32_t op_mask,
                           double max_white, double jodie_a, uint8_t* const rgb8[RTE_TM_COUNT]);
/* The operator functions themselves — simple, reinhardSimple, reinhardExtended(max_white), reinhardExtendedLuminance(max_white),
 * reinhardJodie(a), uncharted2, aces_approx (RE/RaytracingEngine.cpp:89-163) — on n FP64 colours: in3 / out3 = xyz triples (HOST). */
int rte_tonemap_values_host(rte_ctx* ctx, int op, double max_white, double jodie_a, const double* in3, int64_t n, double* out3);
/* Max of luminance() (RE/RaytracingEngine.cpp:100-104) over a device HDR buffer; the per-rank
 * value that an auto-max-white run all-reduces (MAX) across GPUs. Result on the host. */
int rte_max_luminance(rte_ctx* ctx, const float* hdr, int64_t n_pixels, float* out_max);
/* Stream-ordered variants for the multi-GPU auto-max-white pipeline: the per-rank max goes to a
 * device float (no host round trip), the caller all-reduces it (MAX) across ranks on the same stream, and the
 * tonemap pass reads max_white from that device float.  *max_white_dev <= 0 falls back to the reference's 5.0. */
int rte_max_luminance_device(rte_ctx* ctx, const float* hdr, int64_t n_pixels, float* out_max_dev);
int rte_tonemap_auto_white(rte_ctx* ctx, const float* hdr, int64_t n_pixels, uint32_t op_mask,
                           const float* max_white_dev, double jodie_a, uint8_t* const rgb8[RTE_TM_COUNT]);

/* Allocation-free variant of rte_max_luminance_device.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  The call allocates nothing — no cudaMalloc / cudaFree, no allocating library routine: the
 * reduction's scratch (one partial maximum per 4096 pixels + a counter) is a file-scope static __device__ array of
 * 64.25 KB that every GPU loading the library holds until exit, so at most RTE_NOALLOC_MAX_PIXELS pixels per call (more:
 * RTE_ERR_LIMIT).  hdr = float4 per pixel (device, 16-byte aligned), *out_max_dev = device float, stream-ordered on the ctx
 * stream; same value as rte_max_luminance_device (NaN luminances ignored, 0 for n_pixels == 0).  All calls on one GPU share
 * the array (it keeps the last call's contents; the call clears what must start at zero): calls that OVERLAP on the GPU —
 * two contexts or streams of one device — race; calls in stream order do not. */
#define RTE_NOALLOC_MAX_PIXELS 67108864
int rte_max_luminance_noalloc(rte_ctx* ctx, const float* hdr, int64_t n_pixels, float* out_max_dev);

/* Measurement aid (no reference counterpart): runs a register-resident FFMA kernel on every SM of the
 * ctx device for about `target_ms` and reports the achieved FP32 rate — the measured FP32-FMA roofline
 * denominator bench.py quotes the render kernel against (MEASURED_PEAKS.json has none). */
int rte_measure_fp32_peak(rte_ctx* ctx, double target_ms, double* out_tflops, double* out_ms);
/* The same with packed fma.rn.f32x2 instructions (two FMAs per lane and instruction, sm_100+). */
int rte_measure_fp32x2_peak(rte_ctx* ctx, double target_ms, double* out_tflops, double* out_ms);

/* Scene::IntersectClosest(ray), RE/Scene.h:218-257, for a batch of rays (host arrays:
 * origins/directions = xyz triples).  hits[i].prim_id == 0xFFFFFFFF means std::nullopt. */
int rte_intersect_closest(rte_ctx* ctx, int64_t n_rays, const double* origins,
                          const double* directions, rte_hit* hits);
/* Scene::IntersectAnyBefore(ray, maxDist), RE/Scene.h:259-276: out[i] = 1 if any primitive
 * reports 0 < t < max_dist[i]. */
int rte_intersect_any_before(rte_ctx* ctx, int64_t n_rays, const double* origins,
                             const double* directions, const double* max_dist, uint8_t* out);

/* Device memory helpers for callers without their own allocator. */
/* Multi-GPU output without a separate all-gather (no reference counterpart — the reference is one
 * process writing one std::vector): every rank registers the full-frame RGB8 buffers (W*H*3 bytes, 4-byte aligned)
 * of all ranks of the NVSwitch domain, its own included; each render with opts.tonemap_op != RTE_TM_NONE then stores
 * the RGB8 of every pixel it renders at 3*(y*W + x) of EVERY registered frame (peer stores over NVLink, issued by the
 * render kernel while it computes), on top of frame.rgb8 (which may be NULL).  The caller fences the frame with
 * rte_frame_fence() (below) or any cross-rank barrier on the same stream.  n_frames = 0 unregisters.  At most RTE_MAX_PEERS frames. */
#define RTE_MAX_PEERS 8
int rte_set_rgb8_peers(rte_ctx* ctx, int n_frames, uint8_t* const* frames);
/* Frame fence for the above without a collective library.  Every rank owns an array of RTE_MAX_PEERS uint32 flags
 * (rte_malloc; mapped into the peers like the frames) and registers the arrays of all ranks, flags[r] = rank r's array,
 * flags[rank] its own (zeroed by this call; follow it with one cross-rank barrier before the first render).  From then on
 * the kernel that performs a render's peer stores ends by writing the render's sequence number to flags[r][rank] of every
 * rank r (system-scope release store over NVLink, after its stores), and rte_frame_fence() enqueues a device-side wait on
 * the context's stream until every rank has delivered the frame of the last render call: work enqueued behind it sees the
 * complete frame.  A peer that does not deliver within timeout_ms (<= 0: 2000) makes the NEXT rte_synchronize /
 * rte_get_stats of this context fail with RTE_ERR_TIMEOUT.  Ranks must issue the same sequence of render calls.
 * n_ranks = 0 unregisters. */
int rte_set_fence_peers(rte_ctx* ctx, int rank, int n_ranks, uint32_t* const* flags);
int rte_frame_fence(rte_ctx* ctx, double timeout_ms);
/* CUDA IPC plumbing for the above when ranks are processes: export a buffer obtained from rte_malloc as an opaque
 * 64-byte handle, map a peer's handle into this process (enables peer access), unmap it. */
#define RTE_IPC_HANDLE_BYTES 64
int rte_ipc_export(rte_ctx* ctx, const void* dev, uint8_t handle[RTE_IPC_HANDLE_BYTES]);
int rte_ipc_open(rte_ctx* ctx, const uint8_t handle[RTE_IPC_HANDLE_BYTES], void** out_dev);
int rte_ipc_close(rte_ctx* ctx, void* dev);

int rte_malloc(rte_ctx* ctx, size_t bytes, void** out_dev);
int rte_free(rte_ctx* ctx, void* dev);
int rte_memcpy_d2h(rte_ctx* ctx, void* host, const void* dev, size_t bytes);
int rte_memcpy_h2d(rte_ctx* ctx, void* dev, const void* host, size_t bytes);
/* Pinned host memory (used for the RGB8 staging buffer of writePPM). */
int rte_malloc_host(size_t bytes, void** out_host);
int rte_free_host(void* host);

/* writePPM(filename, pixels, width, height), RE/Image.cpp:11-31: "P6\n<W> <H>\n255\n" + RGB
 * bytes, one fwrite from the (pinned) host buffer. Needs no device. */
int rte_write_ppm(const char* filename, const uint8_t* rgb8, uint64_t width, uint64_t height);
/* What main() gets from `ffmpeg -i x.ppm x.png` + `del x.ppm` (RE/RaytracingEngine.cpp:317-327): the same pixels as an
 * 8-bit RGB PNG (stored deflate blocks; no external tool).  Needs no GPU. */
int rte_write_png(const char* filename, const uint8_t* rgb8, uint64_t width, uint64_t height);

#ifdef __cplusplus
}
#endif
#endif /* RTE_H_ */
