// rte_tonemap.cuh — the seven tonemap operators + 8-bit quantiser.
//
// Two flavours:
//   * tm64::  FP64, the reference's exact evaluation order (float-typed constants widened where the
//             reference does) — used once per pixel by the fused render epilogue;
//   * tm32::  FP32 with shared reciprocals — used by the stand-alone HBM-bound tonemap pass.
// Reference: RE/RaytracingEngine.cpp:70-163 ("RE/" = /root/reference/RaytracingEngine/).
#pragma once
#include "rte_device.cuh"

namespace rte {
namespace tm64 {

// ClampVec3, RE/RaytracingEngine.cpp:70-76: std::min(maxVal, std::max(minVal, v)); NaN -> minVal
__device__ __forceinline__ double clamp_mm(double v, double lo, double hi) { return std_min(hi, std_max(lo, v)); }
__device__ __forceinline__ d3 clamp3(d3 v, double lo, double hi) { return mk(clamp_mm(v.x, lo, hi), clamp_mm(v.y, lo, hi), clamp_mm(v.z, lo, hi)); }

// uncharted2_tonemap_partial, :78-87 — C*B, D*E, D*F, E/F are float expressions in the reference
__device__ __forceinline__ d3 uncharted2_partial(d3 x) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    const float CB = __fmul_rn(C, B), DE = __fmul_rn(D, E), DF = __fmul_rn(D, F), EF = __fdiv_rn(E, F);
    const d3 num = adds(x * adds(x * (double)A, (double)CB), (double)DE);
    const d3 den = adds(x * adds(x * (double)A, (double)B), (double)DF);
    return subs(num / den, (double)EF);
}
// aces_approx, :89-98
__device__ __forceinline__ d3 aces(d3 v) {
    v = v * (double)0.6f;
    const float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    const d3 num = v * adds(v * (double)a, (double)b);
    const d3 den = adds(v * adds(v * (double)c, (double)d), (double)e);
    return clamp3(num / den, 0.0, 1.0);
}
// luminance / change_luminance, :100-110
__device__ __forceinline__ double luminance(d3 c) { return dot(c, mk(0.2126, 0.7152, 0.0722)); }
__device__ __forceinline__ d3 change_luminance(d3 c, double l_out) { const double l_in = luminance(c); return c * (l_out / l_in); }
// simple, :123-131
__device__ __forceinline__ d3 simple(d3 c) { return mk(std_min(1.0, std_max(0.0, c.x)), std_min(1.0, std_max(0.0, c.y)), std_min(1.0, std_max(0.0, c.z))); }
// reinhardSimple, :133-135
__device__ __forceinline__ d3 reinhard(d3 c) { return c / adds(c, 1.0); }
// reinhardExtended, :137-141
__device__ __forceinline__ d3 reinhard_ext(d3 c, double max_white) {
    const double wsq = max_white * max_white;
    const d3 numerator = c * adds(c / mk(wsq, wsq, wsq), 1.0);
    return numerator / adds(c, 1.0);
}
// reinhardExtendedLuminance, :143-148
__device__ __forceinline__ d3 reinhard_ext_lum(d3 c, double max_white) {
    const double L_old = luminance(c);
    const double numerator = L_old * (1.0 + (L_old / (max_white * max_white)));
    const double l_new = numerator / (1.0 + L_old);
    return change_luminance(c, l_new);
}
// reinhardJodie, :150-154
__device__ __forceinline__ d3 jodie(d3 c, double a) {
    const double L = luminance(c);
    const double L_mapped = (a / log(2.0 + pow((L / 0.85), 1.7))) * log(1.0 + L);
    return change_luminance(c, L_mapped);
}
// uncharted2, :156-163
__device__ __forceinline__ d3 uncharted2(d3 c) {
    const d3 curr = uncharted2_partial(c * 2.0);
    const d3 whiteScale = mk(1.0, 1.0, 1.0) / uncharted2_partial(mk(11.2, 11.2, 11.2));
    return curr * whiteScale;
}
__device__ __forceinline__ d3 apply(int op, d3 c, double max_white, double a) {
    switch (op) {
        case RTE_TM_SIMPLE: return simple(c);
        case RTE_TM_REINHARD_SIMPLE: return reinhard(c);
        case RTE_TM_REINHARD_EXTENDED: return reinhard_ext(c, max_white);
        case RTE_TM_REINHARD_EXTENDED_LUMINANCE: return reinhard_ext_lum(c, max_white);
        case RTE_TM_REINHARD_JODIE: return jodie(c, a);
        case RTE_TM_UNCHARTED2: return uncharted2(c);
        default: return aces(c);
    }
}
// toColor, :113-121: clamp to [0,1], * 255.0, truncate
__device__ __forceinline__ uchar3 to_color(d3 p) {
    const d3 c = clamp3(p, 0.0, 1.0);
    return make_uchar3((unsigned char)(c.x * 255.0), (unsigned char)(c.y * 255.0), (unsigned char)(c.z * 255.0));
}

}  // namespace tm64

// FP32 flavour for the stand-alone HBM-bound pass (tonemapAll over a float4 HDR buffer).  Results may
// differ from the FP64 reference by one quantisation step when c*255 lands within ~1e-4 of an integer
// (tolerance of the pass: +-1 LSB, tests/test_gpu_tonemap.py); divisions are rcp.approx based and
// log/pow are lg2/ex2.approx based, i.e. 2 instructions each instead of the ~10/~50 of the IEEE forms.
namespace tm32 {

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ float sat(float v) { return fminf(1.0f, fmaxf(0.0f, v)); }  // NaN -> 0 like ClampVec3
__device__ __forceinline__ float qdiv(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ float luminance(f3 c) { return fmaf(c.z, 0.0722f, fmaf(c.y, 0.7152f, c.x * 0.2126f)); }

__device__ __forceinline__ float u2p(float x) {  // uncharted2_tonemap_partial, RE/RaytracingEngine.cpp:78-87
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    const float num = fmaf(x, fmaf(x, A, C * B), D * E), den = fmaf(x, fmaf(x, A, B), D * F);
    return qdiv(num, den) - E / F;
}
__device__ __forceinline__ float aces1(float x) {  // aces_approx, :89-98
    const float v = x * 0.6f;
    return sat(qdiv(v * fmaf(v, 2.51f, 0.03f), fmaf(v, fmaf(v, 2.43f, 0.59f), 0.14f)));
}

struct Consts {
    float inv_wsq;      // 1 / max_white^2
    float jodie_a;
    float white_scale;  // 1 / uncharted2_partial(11.2)
};

// all operators of `mask` for one pixel; out[op] = quantised RGB packed as r | g<<8 | b<<16
template <uint32_t kMask>
__device__ __forceinline__ void apply_all(f3 c, const Consts& k, uint32_t mask, uint32_t out[RTE_TM_COUNT]) {
    auto pack = [](float r, float g, float b) -> uint32_t {
        return __float2uint_rz(sat(r) * 255.0f) | (__float2uint_rz(sat(g) * 255.0f) << 8) | (__float2uint_rz(sat(b) * 255.0f) << 16);
    };
    const uint32_t m = kMask ? kMask : mask;
    const f3 r1 = mk3(__frcp_rn(c.x + 1.0f), __frcp_rn(c.y + 1.0f), __frcp_rn(c.z + 1.0f));  // shared by ops 1 and 2
    if (m & (1u << RTE_TM_SIMPLE)) out[RTE_TM_SIMPLE] = pack(c.x, c.y, c.z);
    if (m & (1u << RTE_TM_REINHARD_SIMPLE)) out[RTE_TM_REINHARD_SIMPLE] = pack(c.x * r1.x, c.y * r1.y, c.z * r1.z);
    if (m & (1u << RTE_TM_REINHARD_EXTENDED))
        out[RTE_TM_REINHARD_EXTENDED] = pack(c.x * fmaf(c.x, k.inv_wsq, 1.0f) * r1.x, c.y * fmaf(c.y, k.inv_wsq, 1.0f) * r1.y,
                                             c.z * fmaf(c.z, k.inv_wsq, 1.0f) * r1.z);
    if (m & ((1u << RTE_TM_REINHARD_EXTENDED_LUMINANCE) | (1u << RTE_TM_REINHARD_JODIE))) {
        const float L = luminance(c);
        const float invL = __frcp_rn(L);  // L == 0 -> inf, 0 * inf = NaN -> 0 after sat, like 0/0 in the reference
        if (m & (1u << RTE_TM_REINHARD_EXTENDED_LUMINANCE)) {
            const float s = qdiv(L * fmaf(L, k.inv_wsq, 1.0f), 1.0f + L) * invL;
            out[RTE_TM_REINHARD_EXTENDED_LUMINANCE] = pack(c.x * s, c.y * s, c.z * s);
        }
        if (m & (1u << RTE_TM_REINHARD_JODIE)) {  // (a / ln(2 + (L/0.85)^1.7)) * ln(1 + L), :150-154
            const float p = exp2f(1.7f * __log2f(L * (1.0f / 0.85f)));
            const float s = qdiv(k.jodie_a * __log2f(1.0f + L), __log2f(2.0f + p)) * invL;  // ln ratios == log2 ratios
            out[RTE_TM_REINHARD_JODIE] = pack(c.x * s, c.y * s, c.z * s);
        }
    }
    if (m & (1u << RTE_TM_UNCHARTED2))
        out[RTE_TM_UNCHARTED2] = pack(u2p(c.x * 2.0f) * k.white_scale, u2p(c.y * 2.0f) * k.white_scale, u2p(c.z * 2.0f) * k.white_scale);
    if (m & (1u << RTE_TM_ACES)) out[RTE_TM_ACES] = pack(aces1(c.x), aces1(c.y), aces1(c.z));
}

}  // namespace tm32
}  // namespace rte
