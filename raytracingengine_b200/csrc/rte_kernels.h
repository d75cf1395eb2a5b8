// rte_kernels.h — host-visible launch interface of the CUDA kernels (internal to librte_b200.so).
#pragma once
#include <cuda_runtime.h>

#include "rte_device.cuh"
#include "rte_grid.cuh"

namespace rte {

enum StatSlot {
    kStatPrimary = 0, kStatShadow, kStatSecondary, kStatSphTests, kStatPlTests, kStatTriTests,
    kStatMarches, kStatConfirms, kStatSphExec, kStatCount
};

// what the render kernel stages into dynamic shared memory (rte_render.cu: plan_smem)
struct SmemPlan {
    uint32_t view_bytes, sph_bytes, tri_bytes;
    int32_t use_views;   // sphere phases of primary / shadow rays go through the per-origin views (rte_views.cuh)
    int32_t stage_sph;   // generic sph32 array is in shared memory (else read from global by the rare generic queries)
    int32_t stream_views;  // views do not fit: view_bytes = 2 buffers of one view each, refilled by TMA while the other is scanned
    uint32_t one_view_bytes;
    uint32_t tview_bytes;  // triangle views (bounding spheres of the triangles), resident
    int32_t use_tri_views; // triangle phases of primary / shadow rays go through them
    int32_t stage_tri;     // generic tri32 array is in shared memory (else read from global by the rare generic queries)
    size_t total;
};
SmemPlan plan_smem(const DevScene& sc, size_t budget);

constexpr int kMaxPeers = 8;  // GPUs of one NVSwitch domain whose frame buffers a render can store into

struct RenderParams {
    DevScene sc;
    SmemPlan plan;
    DevCamera cam;
    int32_t row0, row1;      // band [row0,row1)
    int32_t n_rows;          // rows this launch renders (== row1-row0 unless striped)
    int32_t stripe_rows, stripe_period;  // interleaved stripes (0 = contiguous), see rte_render_opts
    int32_t max_depth;       // Scene::maxRecursion, RE/Scene.h:24
    int32_t tonemap_op;      // RTE_TM_* or RTE_TM_NONE
    int32_t exact_shadows;   // RTE_FLAG_EXACT_SHADOWS: every shadow ray takes the literal march
    int32_t no_filter;       // RTE_FLAG_NO_FILTER: FP32 filters accept everything
    int32_t force_general;   // RTE_FLAG_GENERAL_KERNEL: never take the lean fast kernel (validation)
    double bias, max_white, jodie_a;
    rte_frame out;           // device pointers, any may be NULL
    int32_t cull;            // fast kernel: per-warp cull scratch follows the staged arrays in dynamic shared memory (rte_views.cuh)
    int32_t no_cull;         // RTE_FLAG_NO_CULL: never use it (validation / A-B)
    int32_t cta_cull;        // streamed views: the CTA-level cull scratch follows the per-warp scratch (rte_views.cuh)
    int32_t pass_mode;       // fast kernel: 0 = auto, 1 = RTE_FLAG_SINGLE_PASS, 2 = RTE_FLAG_TWO_PASS (primary-ray launch + shading launch)
    double2* hits;           // two-pass fast path: {t, (kind << 30 | index) as the low word of the second double} per band pixel
    int32_t exact_grid;      // RTE_FLAG_EXACT_GRID: the exact path walks the direction grids (render_exact_grid_kernel) even where the views fit
    int32_t exact_fp64;      // RTE_FLAG_EXACT_FP64: never take the FP32 production path (rte_render32.cu)
    DevGrid grid;            // FP32 path: per-origin direction grids (rte_grid.cuh), built per camera / bias
    uint8_t* tile_flags;               // reflective FP32 path (rte_render32r.cu): one byte per 8 x 4 tile, set when the tile needs the exact kernel
    const uint8_t* tile_mask;          // exact kernel render_kernel<1, *>: render only the tiles whose byte is set (and write RGB8 itself)
    unsigned int* tile_counter;        // zeroed before launch
    unsigned long long* stats;         // kStatCount counters, zeroed before launch
};

// dynamic shared memory the render / query kernels need for this scene (FP32 filter arrays)
size_t scene_smem_bytes(const DevScene& sc);
// largest dynamic shared memory size the device allows per block (opt-in)
cudaError_t launch_render(const RenderParams& p, int sm_count, cudaStream_t stream);
// FP32 production path (rte_render32.cu) and the exact fix-up of its flagged pixels (rte_render.cu)
bool fp32_path_eligible(const RenderParams& p);
bool exact_grid_wanted(const RenderParams& p);  // exact path: render_exact_grid_kernel (needs p.grid)
// direction grids: count + scan (cell_start[n_cells] = number of entries), then — with `entries` of at least that capacity — fill
// (`aggregates`: grid_scan_blocks(n_cells) 64-bit words of scratch; `epoch`: a number that differs from build to build)
int grid_scan_blocks(int n_cells);
cudaError_t launch_grid_count_scan(const DevScene& sc, const DevCamera& cam, double bias, int G, uint32_t* count, uint32_t* cell_start,
                                   unsigned long long* aggregates, uint32_t epoch, uint32_t* total_host, cudaStream_t stream);
cudaError_t launch_grid_fill(const DevScene& sc, const DevCamera& cam, double bias, int G, uint32_t* count, uint32_t* cell_start, uint32_t* entries,
                             uint32_t capacity, cudaStream_t stream);
cudaError_t launch_render32(const RenderParams& p, int sm_count, cudaStream_t stream);
// reflective scenes with few primitives (rte_render32r.cu); tiles it flags are rendered by launch_render with p.tile_mask
bool fp32r_path_eligible(const RenderParams& p);
cudaError_t launch_render32r(const RenderParams& p, int sm_count, cudaStream_t stream);
// would launch_render split this call into a primary-ray launch and a shading launch (it then needs p.hits)?
bool two_pass_wanted(const RenderParams& p);
// (re)builds sc.views for this camera / bias (rte_views.cuh); a no-op for scenes without spheres
cudaError_t launch_build_views(const DevScene& sc, const DevCamera& cam, double bias, cudaStream_t stream);

// Scene::IntersectClosest / IntersectAnyBefore for a batch of rays (device arrays)
cudaError_t launch_intersect_closest(const DevScene& sc, int64_t n, const double* origins, const double* dirs,
                                     rte_hit* hits, int sm_count, cudaStream_t stream);
cudaError_t launch_intersect_any_before(const DevScene& sc, int64_t n, const double* origins, const double* dirs,
                                        const double* max_dist, uint8_t* out, int sm_count, cudaStream_t stream);

// The exact (FP64) operator + quantiser of a render call: hdr64 (3 doubles per band pixel, written by the render kernel) ->
// RGB8 in the band buffer and in every registered peer frame (rte_set_rgb8_peers: this GPU's and its peers' full frames,
// mapped over NVLink; pixel (x, y) goes to 3*(y*W + x) — the all-gather of the row stripes, fused into this pass).
struct Tonemap64Params {
    const double* hdr64;
    int64_t n_pixels;        // band pixels = n_rows * W
    int32_t op;              // RTE_TM_*
    double max_white, jodie_a;
    uint8_t* rgb8;           // band-local, may be NULL
    uint8_t* peer_rgb8[kMaxPeers];
    int32_t n_peers;
    int32_t W, row0, stripe_rows, stripe_period;  // band row -> image row (rte_render_opts)
    // frame fence (rte_set_fence_peers): when this launch's peer stores are done, flags[r][rank] = seq on every rank r
    uint32_t* fence_flags[kMaxPeers];
    int32_t n_fence, fence_rank;
    uint32_t fence_seq;
    unsigned int* fence_counter;  // CTAs of this launch that are done (self-resetting)
};
// waits (device side, on `stream`) until flags[r] >= seq for every r < n; after timeout_ns writes 1 to *timed_out (mapped host word)
cudaError_t launch_fence_wait(const uint32_t* flags, int n, uint32_t seq, unsigned long long timeout_ns, unsigned int* timed_out, cudaStream_t stream);
cudaError_t launch_tonemap64(const Tonemap64Params& p, int sm_count, cudaStream_t stream);
// FP32 path, multi-GPU: copies the band's RGB8 (p.rgb8, p.n_pixels, band row -> image row like above) into every p.peer_rgb8 frame and signals the fence
cudaError_t launch_exchange_rgb8(const Tonemap64Params& p, int sm_count, cudaStream_t stream);

// tonemapAll / tonemap over a float4 HDR buffer (stand-alone HBM-bound pass)
struct TonemapParams {
    const float4* hdr;
    int64_t n_pixels;
    uint32_t op_mask;
    float max_white, jodie_a;
    uint8_t* rgb8[RTE_TM_COUNT];
    const float* max_white_dev;  // non-NULL: max_white is read from device memory (auto max-white after the all-reduce)
};
cudaError_t launch_tonemap(const TonemapParams& p, int sm_count, cudaStream_t stream);
// frame.depth_norm: (depth64 - near) / (far - near) per pixel, FP64 like HitInfo::normalizedDistance (RE/Shape.h:40-42), rounded to FP32
cudaError_t launch_depth_norm(const double* depth64, int64_t n_pixels, double near_plane, double far_plane, float* out, int sm_count, cudaStream_t stream);
// max luminance (RE/RaytracingEngine.cpp:100-104) into *out (device float, bit pattern max as uint)
cudaError_t launch_max_luminance(const float4* hdr, int64_t n_pixels, unsigned int* out_bits, int sm_count,
                                 cudaStream_t stream);

// tm64 over 3 doubles per pixel: the operator's FP64 result (out3) and / or its toColor bytes (rgb8)
cudaError_t launch_tonemap64_values(const double* in3, int64_t n, int op, double max_white, double jodie_a, double* out3, uint8_t* rgb8, int sm_count,
                                    cudaStream_t stream);
// rte_max_luminance_noalloc: the same maximum into *out_dev (device float) with its scratch in a file-scope device array of
// rte_tonemap.cu (no cudaMalloc); `device` = the current device; n_pixels <= RTE_NOALLOC_MAX_PIXELS
cudaError_t launch_max_luminance_noalloc(const float4* hdr, int64_t n_pixels, float* out_dev, int device, cudaStream_t stream);

// register-resident FFMA chains on every SM (rte_peak.cu): *flop_out = FP32 flop the launch executes
cudaError_t launch_ffma_peak(float* scratch, int outer, int sm_count, cudaStream_t stream, double* flop_out, int packed);

}  // namespace rte
