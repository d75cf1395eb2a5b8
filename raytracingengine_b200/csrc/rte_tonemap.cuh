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
// one out-of-line copy: the fused epilogue runs it once per pixel, and inlining seven operators (pow / log included)
// into the render kernels only adds to their instruction footprint
static __device__ __noinline__ d3 apply(int op, d3 c, double max_white, double a) {
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
// static_cast<uint8_t>(c * 255.0) for c in [0,1] without a double -> integer conversion instruction: adding 2^52 with
// round-towards-zero leaves floor(c * 255.0) in the low mantissa bits
__device__ __forceinline__ unsigned char trunc_byte(double c01) {
    return (unsigned char)(__double2loint(__dadd_rz(c01 * 255.0, 4503599627370496.0)) & 0xFF);
}
__device__ __forceinline__ uchar3 to_color(d3 p) {
    const d3 c = clamp3(p, 0.0, 1.0);
    return make_uchar3(trunc_byte(c.x), trunc_byte(c.y), trunc_byte(c.z));
}

}  // namespace tm64

// FP32 flavour for the stand-alone HBM-bound pass (tonemapAll over a float4 HDR buffer).  Results may differ
// from the FP64 reference by one quantisation step when c*255 lands within ~1e-4 of an integer (tolerance of
// the pass: +-1 LSB, tests/test_gpu_tonemap.py).  Built to stay off the XU pipe (16 lanes/clk/SM), which bounded
// the first version (ncu r01a: XU 77 %, DRAM 32 %):
//   * divisions are one MUFU.RCP (rcp.approx.ftz) + one FMUL; log/pow are MUFU.LG2 / MUFU.EX2;
//   * clamp to [0,1] is the .SAT modifier of the producing FMUL/FFMA (NaN -> 0 exactly like ClampVec3's
//     std::max(0, NaN), +inf -> 1);
//   * float -> byte is NOT an F2I (XU pipe): fma.rz(y, 255, 2^23) leaves floor(y*255) in the low mantissa byte,
//     and four such bytes are merged with PRMT (ALU pipe).
namespace tm32 {

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mul_sat(float a, float b) { float r; asm("mul.rn.sat.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float luminance(f3 c) { return fmaf(c.z, 0.0722f, fmaf(c.y, 0.7152f, c.x * 0.2126f)); }
// toColor's (uint8)(c * 255.0) for c already in [0,1]: the byte sits in bits 0..7 of the result
__device__ __forceinline__ uint32_t qbyte(float y01) { return __float_as_uint(__fmaf_rz(y01, 255.0f, 8388608.0f)); }

struct Consts {
    float inv_wsq;      // 1 / max_white^2
    float jodie_a;
    float white_scale;  // 1 / uncharted2_partial(11.2)
};

// Per-pixel values shared by several operators (computed once, kept while the operators are applied one after the
// other so that only one operator's 12 output bytes are live at a time: 48 registers instead of 128).
struct Shared {
    f3 c;     // the pixel
    f3 r1;    // 1 / (1 + c): Reinhard, Reinhard extended
    f3 cc;    // c limited for the rational operators (see below)
    float L;  // luminance
};
template <uint32_t kMask>
__device__ __forceinline__ Shared prepare(f3 c, uint32_t mask) {
    const uint32_t m = kMask ? kMask : mask;
    Shared s;
    s.c = c;
    if (m & ((1u << RTE_TM_REINHARD_SIMPLE) | (1u << RTE_TM_REINHARD_EXTENDED))) s.r1 = mk3(rcp(c.x + 1.0f), rcp(c.y + 1.0f), rcp(c.z + 1.0f));
    if (m & ((1u << RTE_TM_REINHARD_EXTENDED_LUMINANCE) | (1u << RTE_TM_REINHARD_JODIE))) s.L = luminance(c);
    if (m & ((1u << RTE_TM_UNCHARTED2) | (1u << RTE_TM_ACES))) {
        // The two rational operators square their argument: beyond 1e15 they sit on their asymptote (to 1e-15) and FP32
        // would overflow where the FP64 reference does not.  +inf stays special: the reference evaluates inf/inf = NaN -> 0.
        const float qn = __int_as_float(0x7fc00000), big = 1.0e15f;
        auto lim = [&](float v) { return v == INFINITY ? qn : (v > big ? big : v); };  // NaN passes through (fminf would drop it)
        s.cc = mk3(lim(c.x), lim(c.y), lim(c.z));
    }
    return s;
}

// One operator on one prepared pixel -> three channels in [0,1] (NaN already mapped to 0 by .sat)
template <int kOp>
__device__ __forceinline__ f3 apply_op(const Shared& s, const Consts& k) {
    const f3 c = s.c;
    if (kOp == RTE_TM_SIMPLE) return mk3(__saturatef(c.x), __saturatef(c.y), __saturatef(c.z));  // :123-131
    if (kOp == RTE_TM_REINHARD_SIMPLE) return mk3(mul_sat(c.x, s.r1.x), mul_sat(c.y, s.r1.y), mul_sat(c.z, s.r1.z));  // x / (1 + x), :133-135
    if (kOp == RTE_TM_REINHARD_EXTENDED)  // x (1 + x / w^2) / (1 + x), :137-141
        return mk3(mul_sat(c.x * fmaf(c.x, k.inv_wsq, 1.0f), s.r1.x), mul_sat(c.y * fmaf(c.y, k.inv_wsq, 1.0f), s.r1.y),
                   mul_sat(c.z * fmaf(c.z, k.inv_wsq, 1.0f), s.r1.z));
    if (kOp == RTE_TM_REINHARD_EXTENDED_LUMINANCE) {  // x * (Ln / L), Ln = L (1 + L / w^2) / (1 + L), :143-148
        const float L = s.L;
        float f = fmaf(L, k.inv_wsq, 1.0f) * rcp(1.0f + L);
        f = (L == 0.0f) ? __int_as_float(0x7fc00000) : f;  // the reference divides 0 by 0 here: NaN -> 0 in every channel
        return mk3(mul_sat(c.x, f), mul_sat(c.y, f), mul_sat(c.z, f));
    }
    if (kOp == RTE_TM_REINHARD_JODIE) {  // x * (Lm / L), Lm = (a / ln(2 + (L/0.85)^1.7)) * ln(1 + L), :150-154
        const float L = s.L;
        const float q = 1.7f * lg2(L * (1.0f / 0.85f));
        const float lgp = q > 120.0f ? q : lg2(2.0f + ex2(q));  // (L/0.85)^1.7 overflows FP32 long after the 2 stopped mattering
        const float f = (k.jodie_a * lg2(1.0f + L)) * rcp(lgp * L);  // ln ratios == log2 ratios
        return mk3(mul_sat(c.x, f), mul_sat(c.y, f), mul_sat(c.z, f));
    }
    if (kOp == RTE_TM_UNCHARTED2) {  // :78-87,156-163
        auto u2 = [&](float v) {
            const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
            const float x = v + v;
            const float num = fmaf(x, fmaf(x, A, C * B), D * E), den = fmaf(x, fmaf(x, A, B), D * F);
            return mul_sat(fmaf(num, rcp(den), -(E / F)), k.white_scale);
        };
        return mk3(u2(s.cc.x), u2(s.cc.y), u2(s.cc.z));
    }
    auto ac = [](float x) {  // aces_approx, :89-98
        const float v = x * 0.6f;
        return mul_sat(v * fmaf(v, 2.51f, 0.03f), rcp(fmaf(v, fmaf(v, 2.43f, 0.59f), 0.14f)));
    };
    return mk3(ac(s.cc.x), ac(s.cc.y), ac(s.cc.z));
}

// 4 pixels x RGB (bytes in bits 0..7 of twelve words) -> three packed 32-bit words
__device__ __forceinline__ uint32_t pack4(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3) {
    return __byte_perm(__byte_perm(b0, b1, 0x0040), __byte_perm(b2, b3, 0x0040), 0x5410);
}

}  // namespace tm32
}  // namespace rte
