// rte_tonemap.cu — stand-alone tonemap / quantise pass over a float4 HDR buffer (sm_100a), HBM-bound.
//
// Replaces tonemap() / tonemapAll() + toColor (RE/RaytracingEngine.cpp:113-121,165-214; "RE/" =
// /root/reference/RaytracingEngine/).  Algorithmic traffic: 12 B read + 3 B written per
// pixel and operator; all seven operators in one pass = 33 B/px.  The implementation reads the float4
// (16 B/px) once and writes every selected operator from registers:
//   * one thread = 4 consecutive pixels: 4 x LDG.128 in, 3 x STG.32 per operator out (12 B, aligned);
//   * grid = a multiple of the SM count, grid-stride loop, streaming loads (ld.global.cs) because
//     nothing is re-read.
#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "rte_kernels.h"
#include "rte_tonemap.cuh"

namespace rte {

constexpr int kTmThreads = 256;

__device__ __forceinline__ float4 ldg_stream(const float4* p) { return __ldcs(p); }

// one operator over the 4 prepared pixels of a thread -> 12 bytes = 3 aligned words: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
template <int kOp>
__device__ __forceinline__ void store_op(const tm32::Shared (&s)[4], const tm32::Consts& K, uint8_t* rgb8, int64_t q) {
    uint32_t b[12];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const tm32::f3 y = tm32::apply_op<kOp>(s[j], K);
        b[3 * j] = tm32::qbyte(y.x); b[3 * j + 1] = tm32::qbyte(y.y); b[3 * j + 2] = tm32::qbyte(y.z);
    }
    uint32_t* dst = reinterpret_cast<uint32_t*>(rgb8 + 12 * q);
    __stcs(dst + 0, tm32::pack4(b[0], b[1], b[2], b[3]));
    __stcs(dst + 1, tm32::pack4(b[4], b[5], b[6], b[7]));
    __stcs(dst + 2, tm32::pack4(b[8], b[9], b[10], b[11]));
}
template <int kOp>
__device__ __forceinline__ void store_op_tail(const tm32::Shared& s, const tm32::Consts& K, uint8_t* rgb8, int64_t i) {
    const tm32::f3 y = tm32::apply_op<kOp>(s, K);
    uint8_t* d = rgb8 + 3 * i;
    d[0] = (uint8_t)(tm32::qbyte(y.x) & 255u); d[1] = (uint8_t)(tm32::qbyte(y.y) & 255u); d[2] = (uint8_t)(tm32::qbyte(y.z) & 255u);
}

template <uint32_t kMask>
__global__ void __launch_bounds__(kTmThreads, 4) tonemap_kernel(const __grid_constant__ TonemapParams P, const tm32::Consts K0) {
    tm32::Consts K = K0;
    if (P.max_white_dev) {  // auto max-white: the (all-reduced) max luminance of the frame, produced on this stream
        const float w = __ldg(P.max_white_dev);
        K.inv_wsq = w > 0.0f ? 1.0f / (w * w) : K0.inv_wsq;
    }
    const int64_t n_quads = P.n_pixels >> 2;
    const uint32_t mask = kMask ? kMask : P.op_mask;
#define RTE_TM_EACH(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6)
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_quads; q += (int64_t)gridDim.x * blockDim.x) {
        const float4* src = P.hdr + 4 * q;
        float4 px[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) px[j] = ldg_stream(src + j);
        tm32::Shared s[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) s[j] = tm32::prepare<kMask>(tm32::mk3(px[j].x, px[j].y, px[j].z), mask);
#define RTE_TM_DO(op) if (mask & (1u << op)) store_op<op>(s, K, P.rgb8[op], q);
        RTE_TM_EACH(RTE_TM_DO)
#undef RTE_TM_DO
    }
    // tail (n_pixels % 4) — bytes
    const int64_t tail0 = n_quads << 2;
    const int64_t i = tail0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.n_pixels) {
        const float4 v = P.hdr[i];
        const tm32::Shared s = tm32::prepare<kMask>(tm32::mk3(v.x, v.y, v.z), mask);
#define RTE_TM_DO(op) if (mask & (1u << op)) store_op_tail<op>(s, K, P.rgb8[op], i);
        RTE_TM_EACH(RTE_TM_DO)
#undef RTE_TM_DO
    }
#undef RTE_TM_EACH
}

// Developer probe (environment RTE_TM_NULL=1, tools/tm_probe.py): the memory pattern of the 7-operator pass WITHOUT its
// arithmetic — 4 x LDG.128 in, 7 x 12 bytes out per thread, as 3 x STG.32 (kWide = false, the production pattern) or
// transposed through shared memory into full-sector STG.128 (kWide = true).  Measured on B200 (4K, back to back, two rotating
// inputs): 67.8 / 63.7 us against 62.1 us for tonemap_kernel<127> itself — the pass is bound by its access pattern (133 MB
// read + 174 MB written through eight streams = 4.9 TB/s of actual DRAM traffic), not by the operators' arithmetic: a packed
// FP32x2 version of the operators (-33 % instructions, 55 us under ncu's cold-cache replay) ran 68.8 us back to back.
template <bool kWide>
__global__ void __launch_bounds__(kTmThreads, 4) tonemap_null_kernel(const __grid_constant__ TonemapParams P) {
    __shared__ uint32_t s_w[kTmThreads / 32][RTE_TM_COUNT][96];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t n_quads = P.n_pixels >> 2;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_quads; q += (int64_t)gridDim.x * blockDim.x) {
        const float4* src = P.hdr + 4 * q;
        float4 px[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) px[j] = ldg_stream(src + j);
        const uint32_t a = __float_as_uint(px[0].x) ^ __float_as_uint(px[1].y), b = __float_as_uint(px[2].z) ^ __float_as_uint(px[3].x),
                       c = __float_as_uint(px[0].y) ^ __float_as_uint(px[1].z) ^ __float_as_uint(px[2].x) ^ __float_as_uint(px[3].y) ^
                           __float_as_uint(px[0].z) ^ __float_as_uint(px[1].x) ^ __float_as_uint(px[2].y) ^ __float_as_uint(px[3].z);
        if (!kWide) {
#pragma unroll
            for (int op = 0; op < RTE_TM_COUNT; ++op) {
                uint32_t* dst = reinterpret_cast<uint32_t*>(P.rgb8[op] + 12 * q);
                __stcs(dst + 0, a + op); __stcs(dst + 1, b + op); __stcs(dst + 2, c + op);
            }
        } else {  // (whole warps only: n_quads a multiple of 32 in the probe)
#pragma unroll
            for (int op = 0; op < RTE_TM_COUNT; ++op) { s_w[warp][op][3 * lane] = a + op; s_w[warp][op][3 * lane + 1] = b + op; s_w[warp][op][3 * lane + 2] = c + op; }
            __syncwarp();
            if (lane < 24) {
#pragma unroll
                for (int op = 0; op < RTE_TM_COUNT; ++op)
                    reinterpret_cast<uint4*>(P.rgb8[op] + 12 * (q - lane))[lane] = reinterpret_cast<const uint4*>(s_w[warp][op])[lane];
            }
            __syncwarp();
        }
    }
}

static float u2p_host(float x) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
}

cudaError_t launch_tonemap(const TonemapParams& p, int sm_count, cudaStream_t stream) {
    if (p.n_pixels <= 0) return cudaSuccess;
    tm32::Consts k;
    k.inv_wsq = 1.0f / (p.max_white * p.max_white);
    k.jodie_a = p.jodie_a;
    k.white_scale = 1.0f / u2p_host(11.2f);
    const int64_t n_quads = (p.n_pixels + 3) / 4;
    int64_t blocks = (n_quads + kTmThreads - 1) / kTmThreads;
    constexpr uint32_t kAll = (1u << RTE_TM_COUNT) - 1u;
    if (const char* probe = std::getenv("RTE_TM_NULL")) {  // developer probe, see tonemap_null_kernel
        const unsigned nb = (unsigned)std::min<int64_t>(blocks, (int64_t)sm_count * 4);
        if (probe[0] == '2') tonemap_null_kernel<true><<<nb, kTmThreads, 0, stream>>>(p);
        else tonemap_null_kernel<false><<<nb, kTmThreads, 0, stream>>>(p);
        return cudaGetLastError();
    }
    const void* fn = p.op_mask == kAll ? (const void*)tonemap_kernel<kAll>
                   : p.op_mask == (1u << RTE_TM_ACES) ? (const void*)tonemap_kernel<(1u << RTE_TM_ACES)> : (const void*)tonemap_kernel<0u>;
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kTmThreads, 0);
    if (e != cudaSuccess) return e;
    const int64_t cap = (int64_t)sm_count * (per_sm > 0 ? per_sm : 1);  // one resident wave, grid-stride loop
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (p.op_mask == kAll) tonemap_kernel<kAll><<<(unsigned)blocks, kTmThreads, 0, stream>>>(p, k);
    else if (p.op_mask == (1u << RTE_TM_ACES)) tonemap_kernel<(1u << RTE_TM_ACES)><<<(unsigned)blocks, kTmThreads, 0, stream>>>(p, k);
    else tonemap_kernel<0u><<<(unsigned)blocks, kTmThreads, 0, stream>>>(p, k);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------ exact pass of a render call
// One thread = 4 consecutive band pixels: 12 doubles in (streaming loads), tm64 (the reference's FP64 evaluation order),
// 12 bytes = 3 aligned words out — to the band buffer and to every peer frame.  HBM-bound: 96 B read + 12 B written per
// thread; kept out of the render kernel because that kernel is bound by instruction fetch.
__global__ void __launch_bounds__(kTmThreads) tonemap64_kernel(const __grid_constant__ Tonemap64Params P) {
    const int64_t n = P.n_pixels;
    const int W = P.W;
    const bool words = (W & 3) == 0;  // a quad never straddles two rows, rows are word aligned
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; 4 * q < n; q += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p0 = 4 * q;
        const int cnt = (int)(n - p0 < 4 ? n - p0 : 4);
        uint32_t px[4] = {0, 0, 0, 0};
        for (int j = 0; j < cnt; ++j) {
            const double* h = P.hdr64 + 3 * (p0 + j);
            const d3 v = mk(__ldcs(h), __ldcs(h + 1), __ldcs(h + 2));
            const uchar3 u = tm64::to_color(tm64::apply(P.op, v, P.max_white, P.jodie_a));
            px[j] = (uint32_t)u.x | ((uint32_t)u.y << 8) | ((uint32_t)u.z << 16);
        }
        const uint32_t w0 = px[0] | (px[1] << 24), w1 = (px[1] >> 8) | (px[2] << 16), w2 = (px[2] >> 16) | (px[3] << 8);
        // image position of the quad's first pixel (peer frames hold the whole image)
        const int64_t band_y = p0 / W;
        const int x = (int)(p0 - band_y * W);
        const int sr = P.stripe_rows, sp = P.stripe_period;
        const int64_t y = P.row0 + (sr > 0 ? (band_y / sr) * sp + band_y % sr : band_y);
        const int64_t ip0 = y * W + x;
        if (words && cnt == 4) {
            if (P.rgb8) { uint32_t* d = reinterpret_cast<uint32_t*>(P.rgb8 + 3 * p0); __stcs(d, w0); __stcs(d + 1, w1); __stcs(d + 2, w2); }
            for (int i = 0; i < P.n_peers; ++i) {
                uint32_t* d = reinterpret_cast<uint32_t*>(P.peer_rgb8[i] + 3 * ip0);
                __stcs(d, w0); __stcs(d + 1, w1); __stcs(d + 2, w2);
            }
        } else {
            for (int j = 0; j < cnt; ++j) {
                const int64_t pj = p0 + j, by = pj / W;
                const int64_t ipj = (P.row0 + (sr > 0 ? (by / sr) * sp + by % sr : by)) * W + (pj - by * W);
                const uint8_t b0 = (uint8_t)(px[j] & 255u), b1 = (uint8_t)((px[j] >> 8) & 255u), b2 = (uint8_t)(px[j] >> 16);
                if (P.rgb8) { uint8_t* d = P.rgb8 + 3 * pj; d[0] = b0; d[1] = b1; d[2] = b2; }
                for (int i = 0; i < P.n_peers; ++i) { uint8_t* d = P.peer_rgb8[i] + 3 * ipj; d[0] = b0; d[1] = b1; d[2] = b2; }
            }
        }
    }
    // Frame fence over NVLink, fused into the kernel that does the peer stores: every thread makes its stores visible
    // system-wide, the last CTA of the launch to get here tells every rank "rank `fence_rank` has delivered frame `seq`".
    if (P.n_fence > 0) {
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned int done = atomicAdd(P.fence_counter, 1u);
            if (done == gridDim.x - 1) {
                *P.fence_counter = 0u;  // for the next launch (stream order)
                __threadfence_system();
                for (int r = 0; r < P.n_fence; ++r)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(P.fence_flags[r] + P.fence_rank), "r"(P.fence_seq) : "memory");
            }
        }
    }
}

// One warp: lane r polls flags[r] (this rank's own memory, written by rank r over NVLink) until it reaches `seq`.
__global__ void fence_wait_kernel(const uint32_t* flags, int n, uint32_t seq, unsigned long long timeout_ns, unsigned int* timed_out) {
    const int r = (int)threadIdx.x;
    if (r >= n) return;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (true) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + r) : "memory");
        if ((int32_t)(v - seq) >= 0) return;  // (sequence numbers wrap)
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) { *timed_out = 1u; return; }
        __nanosleep(200);
    }
}

// ------------------------------------------------------------------------------ multi-GPU output exchange of the FP32 path
// The FP32 render kernel writes its RGB8 to the band buffer of its own GPU; this launch, right behind it on the stream, copies
// the band's rows into every registered full frame at their image rows (this GPU's and its peers', mapped over NVLink) with
// 16-byte stores — the all-gather of the row stripes as plain wide peer stores — and signals the frame fence like
// tonemap64_kernel does.  (Measured on 2 x B200: storing from the render kernel itself, 24 bytes per tile row and frame, costs the
// kernel ~60 us; counting finished tiles to send whole segments needs a __threadfence per tile and costs more.)
__global__ void __launch_bounds__(kTmThreads) exchange_rgb8_kernel(const __grid_constant__ Tonemap64Params P) {
    const int64_t row_bytes = (int64_t)P.W * 3;
    const int64_t n_rows = P.W > 0 ? P.n_pixels / P.W : 0;
    const int sr = P.stripe_rows, sp = P.stripe_period;
    bool v16 = ((row_bytes | (int64_t)(uintptr_t)P.rgb8) & 15) == 0;
    for (int i = 0; i < P.n_peers; ++i) v16 = v16 && ((uintptr_t)P.peer_rgb8[i] & 15) == 0;
    if (v16) {
        // four 16-byte pieces per thread and round: the four loads are in flight together, then 4 x n_peers stores (one piece
        // per round left the NVLink stores of a thread waiting behind its single load: 515 GB/s out per GPU at N = 8)
        constexpr int kPieces = 4;
        const int64_t per_row = row_bytes >> 4, total = per_row * n_rows, stride = (int64_t)gridDim.x * blockDim.x;
        for (int64_t k0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k0 < total; k0 += stride * kPieces) {
            uint4 v[kPieces];
            int64_t off[kPieces];
#pragma unroll
            for (int j = 0; j < kPieces; ++j) {
                const int64_t k = k0 + j * stride;
                off[j] = -1;
                if (k < total) {
                    const int64_t br = k / per_row, c = k - br * per_row;
                    const int64_t iy = P.row0 + (sr > 0 ? (br / sr) * sp + br % sr : br);
                    v[j] = __ldcs(reinterpret_cast<const uint4*>(P.rgb8 + br * row_bytes) + c);
                    off[j] = iy * row_bytes + (c << 4);
                }
            }
            for (int i = 0; i < P.n_peers; ++i) {
#pragma unroll
                for (int j = 0; j < kPieces; ++j)
                    if (off[j] >= 0) __stcs(reinterpret_cast<uint4*>(P.peer_rgb8[i] + off[j]), v[j]);
            }
        }
    } else {
        const int64_t total = row_bytes * n_rows;
        for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
            const int64_t br = k / row_bytes, c = k - br * row_bytes;
            const int64_t iy = P.row0 + (sr > 0 ? (br / sr) * sp + br % sr : br);
            const uint8_t v = P.rgb8[br * row_bytes + c];
            for (int i = 0; i < P.n_peers; ++i) P.peer_rgb8[i][iy * row_bytes + c] = v;
        }
    }
    if (P.n_fence > 0) {
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned int done = atomicAdd(P.fence_counter, 1u);
            if (done == gridDim.x - 1) {
                *P.fence_counter = 0u;  // for the next launch (stream order)
                __threadfence_system();
                for (int r = 0; r < P.n_fence; ++r)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(P.fence_flags[r] + P.fence_rank), "r"(P.fence_seq) : "memory");
            }
        }
    }
}
cudaError_t launch_exchange_rgb8(const Tonemap64Params& p, int sm_count, cudaStream_t stream) {
    const int64_t units = (p.n_pixels * 3 + 15) / 16;
    int64_t blocks = (units + kTmThreads - 1) / kTmThreads;
    const int64_t cap = (int64_t)sm_count * 4;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    exchange_rgb8_kernel<<<(unsigned)blocks, kTmThreads, 0, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_fence_wait(const uint32_t* flags, int n, uint32_t seq, unsigned long long timeout_ns, unsigned int* timed_out, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    fence_wait_kernel<<<1, 32, 0, stream>>>(flags, n, seq, timeout_ns, timed_out);
    return cudaGetLastError();
}

cudaError_t launch_tonemap64(const Tonemap64Params& p, int sm_count, cudaStream_t stream) {
    if (p.n_pixels <= 0 || (!p.rgb8 && p.n_peers == 0)) {
        if (p.n_fence <= 0) return cudaSuccess;
        // nothing to store (a rank without rows), but the peers wait for this rank's flag: an empty launch signals
    }
    const int64_t quads = (p.n_pixels + 3) / 4;
    int64_t blocks = (quads + kTmThreads - 1) / kTmThreads;
    const int64_t cap = (int64_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    tonemap64_kernel<<<(unsigned)blocks, kTmThreads, 0, stream>>>(p);
    return cudaGetLastError();
}

// HitInfo::normalizedDistance over the frame's FP64 depth (RE/Shape.h:40-42); a miss (+inf) stays +inf for every finite near < far
__global__ void __launch_bounds__(kTmThreads) depth_norm_kernel(const double* __restrict__ depth64, int64_t n, double near_plane, double far_plane, float* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = (float)div64(depth64[i] - near_plane, far_plane - near_plane);
}
cudaError_t launch_depth_norm(const double* depth64, int64_t n, double near_plane, double far_plane, float* out, int sm_count, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    int64_t blocks = (n + kTmThreads - 1) / kTmThreads;
    if (blocks > (int64_t)sm_count * 8) blocks = (int64_t)sm_count * 8;
    depth_norm_kernel<<<(unsigned)blocks, kTmThreads, 0, stream>>>(depth64, n, near_plane, far_plane, out);
    return cudaGetLastError();
}

// max over pixels of luminance() (RE/RaytracingEngine.cpp:100-104); non-negative floats order like their bits
__global__ void __launch_bounds__(kTmThreads) max_luminance_kernel(const float4* __restrict__ hdr, int64_t n, unsigned int* out_bits) {
    float m = 0.0f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = __ldcs(hdr + i);
        m = fmaxf(m, tm32::luminance(tm32::mk3(v.x, v.y, v.z)));  // fmaxf drops NaN
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
    __shared__ float s_m[kTmThreads / 32];
    if ((threadIdx.x & 31) == 0) s_m[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = threadIdx.x < kTmThreads / 32 ? s_m[threadIdx.x] : 0.0f;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
        if (threadIdx.x == 0) atomicMax(out_bits, __float_as_uint(fmaxf(m, 0.0f)));
    }
}

// ------------------------------------------------------------------------------ exact operators over FP64 values
// rte_tonemap_exact_host / rte_tonemap_values_host: tm64 (the reference's FP64 evaluation order, RE/RaytracingEngine.cpp:70-163)
// over 3 doubles per pixel; `out3` receives the operator's Vec3 result, `rgb8` its toColor bytes (either may be NULL).
__global__ void __launch_bounds__(kTmThreads) tonemap64_values_kernel(const double* __restrict__ in3, int64_t n, int op, double max_white, double jodie_a,
                                                                      double* __restrict__ out3, uint8_t* __restrict__ rgb8) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const d3 v = mk(in3[3 * i], in3[3 * i + 1], in3[3 * i + 2]);
        const d3 m = tm64::apply(op, v, max_white, jodie_a);
        if (out3) { out3[3 * i] = m.x; out3[3 * i + 1] = m.y; out3[3 * i + 2] = m.z; }
        if (rgb8) { const uchar3 u = tm64::to_color(m); rgb8[3 * i] = u.x; rgb8[3 * i + 1] = u.y; rgb8[3 * i + 2] = u.z; }
    }
}
cudaError_t launch_tonemap64_values(const double* in3, int64_t n, int op, double max_white, double jodie_a, double* out3, uint8_t* rgb8, int sm_count,
                                    cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    int64_t blocks = (n + kTmThreads - 1) / kTmThreads;
    if (blocks > (int64_t)sm_count * 8) blocks = (int64_t)sm_count * 8;
    tonemap64_values_kernel<<<(unsigned)blocks, kTmThreads, 0, stream>>>(in3, n, op, max_white, jodie_a, out3, rgb8);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------ allocation-free variant
// rte_max_luminance_noalloc (include/rte.h; no counterpart in the reference):
// the same reduction in two stages — every block reduces one fixed chunk of kNaPixelsPerBlock pixels to a partial maximum,
// the last block to finish (device-scope ticket) reduces the partials — with the partials and the ticket counter in a
// file-scope device array instead of memory from cudaMalloc.  Every GPU that loads the library holds the array (64.25 KB)
// until exit; every call on a device shares it, so calls that overlap on the GPU race.
constexpr int kNaPixelsPerBlock = kTmThreads * 16;
constexpr int kNaMaxBlocks = 16384;  // x 4096 pixels = RTE_NOALLOC_MAX_PIXELS
static_assert((long long)kNaMaxBlocks * kNaPixelsPerBlock == RTE_NOALLOC_MAX_PIXELS, "rte.h: RTE_NOALLOC_MAX_PIXELS");
// [0, kNaMaxBlocks): per-block partial maxima (float bits); [kNaMaxBlocks]: blocks done.  In host code this name is only a
// stand-in for the symbol: kernels get the address cudaGetSymbolAddress returns (noalloc_scratch_ptr below).
static __device__ __align__(256) unsigned int g_noalloc_scratch[kNaMaxBlocks + 64];

__device__ __forceinline__ float block_max(float m, float* s_m) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
    if ((threadIdx.x & 31) == 0) s_m[threadIdx.x >> 5] = m;
    __syncthreads();
    m = threadIdx.x < kTmThreads / 32 ? s_m[threadIdx.x] : 0.0f;
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
    }
    return m;  // valid in thread 0
}

__global__ void __launch_bounds__(kTmThreads) max_luminance_noalloc_kernel(const float4* __restrict__ hdr, int64_t n, unsigned int* scratch,
                                                                           float* __restrict__ out) {
    __shared__ float s_m[kTmThreads / 32];
    __shared__ bool s_last;
    const int64_t base = (int64_t)blockIdx.x * kNaPixelsPerBlock;
    const int64_t end = base + kNaPixelsPerBlock < n ? base + kNaPixelsPerBlock : n;
    float m = 0.0f;
    for (int64_t i = base + threadIdx.x; i < end; i += kTmThreads) {
        const float4 v = __ldcs(hdr + i);
        m = fmaxf(m, tm32::luminance(tm32::mk3(v.x, v.y, v.z)));  // fmaxf drops NaN, like launch_max_luminance
    }
    m = block_max(m, s_m);
    if (threadIdx.x == 0) {
        scratch[blockIdx.x] = __float_as_uint(fmaxf(m, 0.0f));
        __threadfence();
        s_last = atomicAdd(scratch + kNaMaxBlocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    m = 0.0f;
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += kTmThreads) m = fmaxf(m, __uint_as_float(__ldcg(scratch + b)));
    __syncthreads();  // (s_m is reused)
    m = block_max(m, s_m);
    if (threadIdx.x == 0) {
        *out = m;
        scratch[kNaMaxBlocks] = 0u;  // the next call (stream order) starts from zero; the launcher clears it as well
    }
}

// The device address of g_noalloc_scratch on the CURRENT device, looked up once per device (cudaGetSymbolAddress is not an
// allocation: the array is part of the module image).
static cudaError_t noalloc_scratch_ptr(int device, unsigned int** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static unsigned int* ptrs[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[device]) {
        unsigned int* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        ptrs[device] = p;
    }
    *out = ptrs[device];
    return cudaSuccess;
}

cudaError_t launch_max_luminance_noalloc(const float4* hdr, int64_t n_pixels, float* out_dev, int device, cudaStream_t stream) {
    if (n_pixels <= 0) return cudaMemsetAsync(out_dev, 0, sizeof(float), stream);
    if (n_pixels > (int64_t)RTE_NOALLOC_MAX_PIXELS) return cudaErrorInvalidValue;  // (the entry point rejects it with a message first)
    unsigned int* scratch = nullptr;
    cudaError_t e = noalloc_scratch_ptr(device, &scratch);
    if (e != cudaSuccess) return e;
    // the array keeps the last call's contents: the ticket counter must start at zero (the partials are overwritten before they are read)
    if ((e = cudaMemsetAsync(scratch + kNaMaxBlocks, 0, sizeof(unsigned int), stream)) != cudaSuccess) return e;
    const unsigned blocks = (unsigned)((n_pixels + kNaPixelsPerBlock - 1) / kNaPixelsPerBlock);
    max_luminance_noalloc_kernel<<<blocks, kTmThreads, 0, stream>>>(hdr, n_pixels, scratch, out_dev);
    return cudaGetLastError();
}

cudaError_t launch_max_luminance(const float4* hdr, int64_t n_pixels, unsigned int* out_bits, int sm_count, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(out_bits, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if (n_pixels <= 0) return cudaSuccess;
    int64_t blocks = (n_pixels + kTmThreads - 1) / kTmThreads;
    const int64_t cap = (int64_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    max_luminance_kernel<<<(unsigned)blocks, kTmThreads, 0, stream>>>(hdr, n_pixels, out_bits);
    return cudaGetLastError();
}

}  // namespace rte
