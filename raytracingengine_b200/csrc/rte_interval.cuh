// rte_interval.cuh — FP32 interval tests of one ray against one primitive (shared by the FP32 production kernels,
// rte_render32.cu and rte_render32r.cu).  Every function classifies a primitive as a certain miss, a certain hit with
// bounds of the reported distance, or "unsure"; the callers re-render pixels with an unsure decision exactly.
//
// The ray is {origin o = o_hi + o_lo (two floats per component), unit direction L (FP32)}.  `eo` / `ed` (where present) bound
// how far this ray may be from the REFERENCE's ray — |o - o_ref| <= eo, |L - L_ref| <= ed — so that a "certain" answer is
// certain for the reference's ray too (primary rays: eo = 0, ed = 4 eps is part of the built-in margins; reflected rays
// carry the accumulated deviation, rte_render32r.cu).
#pragma once
#include "rte_device.cuh"

namespace rte {

__device__ __forceinline__ float sqrt_approx(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

struct Plane32 {
    float nx, ny, nz, d;  // n (FP64 normalised, rounded), d = n.p0 (FP64 dot, rounded)
    int mat;
};

// ------------------------------------------------------------------------------ one sphere candidate against one ray
// The ray is {origin o (FP64), unit direction L (FP32, |L| = 1 +- 4 eps, within 4 eps of the reference's direction)}.
//   g = C - o to a relative error of ~2 eps (see below),  b = g.L,  l = g - b L,  l2 = |l|^2  — no cancellation between
//   quantities of the size of the scene; every absolute error below is bounded by  dl = 20 eps (|b| + |l|) + 1e-7
//   (direction error 4 eps |b| + roundings ~5 eps |g| + parameter scale 4 eps |b|, |g| <= |b| + |l|).
// Returns the class of the candidate and, for a certain hit, bounds of the near root t0 and of the far root t1.
enum { kCandMiss = 0, kCandHit = 1, kCandUnsure = 2 };
struct Cand32 {
    int cls;
    float t0_lo, t0_hi, t1_lo, t1_hi;  // kCandHit: t0 in [t0_lo, t0_hi], t1 in [t1_lo, t1_hi]; kCandUnsure: the chord, if any, lies in [t0_lo, t1_hi]
};
__device__ __forceinline__ Cand32 classify_sphere(const float4 hi, const float4 lo, float ohx, float ohy, float ohz, float olx, float oly, float olz,
                                                  float Lx, float Ly, float Lz, float eo = 0.0f, float ed = 0.0f) {
    // g = C - o from the hi + lo halves: (hi_C - hi_o) is one correctly rounded FP32 subtraction (relative error eps of |g|), the
    // lo halves carry what the roundings of C and o to FP32 lost — the same accuracy as rounding the FP64 difference, without
    // an FP64 -> FP32 conversion per candidate (those issue on the XU pipe, like MUFU)
    const float gx = (hi.x - ohx) + (lo.x - olx), gy = (hi.y - ohy) + (lo.y - oly), gz = (hi.z - ohz) + (lo.z - olz);
    const float b = __fmaf_rn(gz, Lz, __fmaf_rn(gy, Ly, gx * Lx));
    const float lx = __fmaf_rn(-b, Lx, gx), ly = __fmaf_rn(-b, Ly, gy), lz = __fmaf_rn(-b, Lz, gz);
    const float l2 = __fmaf_rn(lz, lz, __fmaf_rn(ly, ly, lx * lx));
    const float l = sqrt_approx(l2) * 1.000001f;
    // + the deviation from the reference's ray: a point of it at parameter t <= |b| + l is within eo + ed t of ours, and b itself moves by ed |g|
    const float dl = __fmaf_rn(20.0f * kEps32 + 2.0f * ed, fabsf(b) + l, 1e-7f + 2.0f * eo);
    const float ra = hi.w;  // |r| rounded to FP32 (what the reference squares)
    const float r_hi = ra * 1.0000002f, r_lo = ra * 0.9999998f;
    Cand32 c;
    const float lm = fmaxf(l * 0.999998f - dl, 0.0f), lp = l + dl;   // lower / upper bound of the true distance line - centre
    if (lm > r_hi) { c.cls = kCandMiss; c.t0_lo = c.t0_hi = c.t1_lo = c.t1_hi = 0.0f; return c; }
    const float hc_hi = sqrt_approx(fmaxf(__fmaf_rn(r_hi, r_hi, -lm * lm), 0.0f)) * 1.000002f + dl;
    const float h_lo2 = __fmaf_rn(r_lo, r_lo, -lp * lp);
    c.t0_lo = b - hc_hi - dl;
    c.t1_hi = b + hc_hi + dl;
    if (!(r_lo > 0.0f) || !(h_lo2 > 0.0f)) { c.cls = kCandUnsure; c.t0_hi = c.t1_hi; c.t1_lo = c.t0_lo; return c; }  // grazing: hit or miss
    const float hc_lo = fmaxf(sqrt_approx(h_lo2) * 0.999998f - dl, 0.0f);
    c.cls = kCandHit;
    c.t0_hi = b - hc_lo + dl;
    c.t1_lo = b + hc_lo - dl;
    return c;
}

// Cheap rejection of a cell's candidate before the interval classification: with the hi halves only (absolute error of g up
// to d = 2 eps (|o|_1 + |g|)), the squared distance line - centre  l2 = |g|^2 - (g.L)^2  carries an error of at most
// E = 4 eps |g|^2 (cancellation, |L| = 1 +- 4 eps) + 2 |g| d + d^2; certainly a miss when l2 - E > (1.001 r)^2.
// (sqrt-free: |g| <= (|g|^2 + 1) / 2.)  Most entries of a cell are spheres the ray passes at a distance.
__device__ __forceinline__ bool far_miss(const float4 hi, float ox, float oy, float oz, float o1, float Lx, float Ly, float Lz) {
    const float gx = hi.x - ox, gy = hi.y - oy, gz = hi.z - oz;
    const float b = __fmaf_rn(gz, Lz, __fmaf_rn(gy, Ly, gx * Lx));
    const float gg = __fmaf_rn(gz, gz, __fmaf_rn(gy, gy, gx * gx));
    const float l2 = __fmaf_rn(-b, b, gg);
    const float gn = 0.5f * gg + 0.5f;                                  // >= |g|
    const float E = kEps32 * __fmaf_rn(4.0f * gn, o1 + gn, 16.0f * gg) + 1e-12f;  // >= 4 eps gg + 2 gn d + d^2 with d = 2 eps (o1 + gn)
    const float r = hi.w * 1.001f;
    return l2 - E > r * r;
}

// Sphere::Intersect's reported distance (RE/Shape.h:89-97: t0 unless t0 < 1e-6, then t1 unless t1 < 1e-6) as an interval.
// Returns kCandMiss / kCandHit ([lo, hi]) / kCandUnsure (a hit, if any, reports a distance >= lo).
__device__ __forceinline__ int reported_t(const Cand32& c, float& lo, float& hi) {
    const float kRefEps = 1e-6f;
    lo = 0.0f; hi = 0.0f;
    if (c.cls == kCandMiss || c.t1_hi < kRefEps) return kCandMiss;         // no root can reach 1e-6
    if (c.cls == kCandHit) {
        if (c.t0_lo > kRefEps) { lo = c.t0_lo; hi = c.t0_hi; return kCandHit; }
        if (c.t0_hi < kRefEps && c.t1_lo > kRefEps) { lo = c.t1_lo; hi = c.t1_hi; return kCandHit; }  // origin inside: far root
    }
    lo = fmaxf(c.t0_lo, 0.0f); hi = c.t1_hi;
    return kCandUnsure;
}

// ------------------------------------------------------------------------------ nearest hit from intervals
struct Near32 {
    int kind, idx;        // RTE_HIT_*; best certain hit so far
    float lo, hi;         // its distance interval
    float amb_lo;         // the smallest distance anything ambiguous (overlapping interval, grazing sphere) could report
};
__device__ __forceinline__ void near_hit(Near32& n, int kind, int idx, float lo, float hi) {
    if (n.kind == RTE_HIT_NONE || hi < n.lo) { n.kind = kind; n.idx = idx; n.lo = lo; n.hi = hi; return; }  // certainly closer
    if (lo > n.hi) return;                                                                                  // certainly farther
    n.amb_lo = fminf(n.amb_lo, fminf(lo, n.lo));  // overlap: canonical order / strict '<' (RE/Scene.h:223) would have to decide
    if (lo < n.lo) { n.kind = kind; n.idx = idx; n.lo = lo; n.hi = hi; }
}
__device__ __forceinline__ void near_unsure(Near32& n, float lo) {
    if (n.kind != RTE_HIT_NONE && lo > n.hi) return;  // whatever it is, it is farther than the best (which only gets closer)
    n.amb_lo = fminf(n.amb_lo, lo);
}
__device__ __forceinline__ bool near_ambiguous(const Near32& n) { return n.kind == RTE_HIT_NONE ? n.amb_lo < INFINITY : n.amb_lo <= n.hi; }

// ------------------------------------------------------------------------------ one plane against one ray
// Plane::Intersect, RE/Shape.h:149-159: denom = n.d; |denom| > 1e-6; t = ((p0 - o).n) / denom; hit iff t >= 0.
// FP32: den = n.L (error <= ed = 8 eps), num = d - n.o (error <= en = 4 eps (|d| + |o|_1) + 1e-9), so
//   t = |num| / |den|  lies in  [(|num| - en) / (|den| + ed), (|num| + en) / (|den| - ed)].
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
struct PlaneND { float ad, an, ed, en; int cls; };  // cls: kCandMiss (no hit), kCandUnsure, kCandHit (t = an / ad >= 0 for certain)
__device__ __forceinline__ PlaneND plane_nd(const Plane32& p, float ox, float oy, float oz, float o1, float Lx, float Ly, float Lz, float eo = 0.0f, float ed = 0.0f) {
    const float den = __fmaf_rn(p.nz, Lz, __fmaf_rn(p.ny, Ly, p.nx * Lx));
    const float num = p.d - __fmaf_rn(p.nz, oz, __fmaf_rn(p.ny, oy, p.nx * ox));
    PlaneND r;
    r.ed = 8.0f * kEps32 + ed; r.en = __fmaf_rn(4.0f * kEps32, fabsf(p.d) + o1, 1e-9f) + eo;  // (|n| = 1: the origin's deviation moves num by at most eo)
    r.ad = fabsf(den); r.an = fabsf(num);
    if (r.ad < 1e-6f - r.ed) r.cls = kCandMiss;                          // parallel: rejected by the reference
    else if (r.ad < 1e-6f + r.ed || r.an <= r.en) r.cls = kCandUnsure;    // |denom| ~ 1e-6, or the origin on the plane (t = 0 is a hit upstream)
    else r.cls = ((num < 0.0f) != (den < 0.0f)) ? kCandMiss : kCandHit;  // t < 0: miss
    return r;
}
// Primary rays: the distance as an interval.
__device__ __forceinline__ int plane_interval(const Plane32& p, float ox, float oy, float oz, float o1, float Lx, float Ly, float Lz, float& lo, float& hi,
                                              float eo = 0.0f, float ed = 0.0f) {
    const PlaneND r = plane_nd(p, ox, oy, oz, o1, Lx, Ly, Lz, eo, ed);
    lo = 0.0f; hi = 0.0f;
    if (r.cls != kCandHit) return r.cls;
    lo = (r.an - r.en) * rcp_approx(r.ad + r.ed) * 0.999999f;
    hi = (r.an + r.en) * rcp_approx(r.ad - r.ed) * 1.000001f;
    return hi < 1.5f * lo + 1e-6f ? kCandHit : kCandUnsure;  // (a uselessly wide interval: leave the pixel to the exact code)
}
// Shadow rays: only the position of t against (bias, maxDist) matters — no division, and no dot product per light either:
//   den * dist = n . (Lp - so) = (n . Lp - d) + (d - n . so) = A + num,
// A a constant of (plane, light) (Scene32::plA), num = d - n . so a value of (plane, pixel) that the caller computes once per pixel.
//   t = num * dist / (A + num);   errors: num <= en, A + num <= e2 = eps |A| + en + eps |A + num|, dist relative 6 eps.
//   kCandMiss: no effect (no hit, t < 0, or certainly beyond the light); kCandHit: certainly bias < t < maxDist; else kCandUnsure.
__device__ __forceinline__ int plane_shadow(float num, float en, float A, float dist, float bias_hi, float tmax_lo, float tmax_hi) {
    const float den2 = A + num;
    const float ad = fabsf(den2), an = fabsf(num);
    const float e2 = __fmaf_rn(kEps32, fabsf(A) + ad, en);
    const float par = 1e-6f * dist;                 // |denom| > 1e-6 of the reference, scaled by dist
    if (ad + e2 < par * 0.99999f) return kCandMiss;  // parallel: rejected upstream
    if (ad - e2 < par * 1.00001f || an <= en) return kCandUnsure;
    if ((num < 0.0f) != (den2 < 0.0f)) return kCandMiss;  // t < 0
    const float n_lo = (an - en) * dist * 0.999999f, n_hi = (an + en) * dist * 1.000001f, d_lo = ad - e2, d_hi = ad + e2;
    if (n_lo > tmax_hi * d_hi) return kCandMiss;
    if (n_lo > bias_hi * d_hi && n_hi < tmax_lo * d_lo) return kCandHit;
    return kCandUnsure;
}

// ------------------------------------------------------------------------------ one triangle against one ray
// Triangle::Intersect, RE/Shape.h:202-220 (Moeller-Trumbore): h = d x e2, a = e1.h, |a| >= 1e-6, s = o - a0, u = (s.h)/a in [0, 1],
// q = s x e1, v = (d.q)/a >= 0, u + v <= 1, t = (e2.q)/a > 1e-6.  Evaluated without the division: with sigma = sign(a), A = |a|,
//   U = sigma s.h,  V = sigma d.q,  T = sigma e2.q:   0 <= U <= A,  V >= 0,  U + V <= A,  T > 1e-6 A.
// Error bounds against the reference's values (edges rounded to FP32: eps |e|; s from hi + lo halves: 2 eps |s| + eo; direction
// ed; three or four roundings per product), with E1 >= |e1|, E2 >= |e2|, S >= |s|, g = 12 eps:
//   |da| <= (ed + g) E1 E2,   |dU| <= E2 (eo + S (ed + g)),   |dV| <= E1 (eo + S (ed + g)),   |dT| <= E1 E2 (eo + S g).
// (struct Tri32: rte_device.cuh)
__device__ __forceinline__ int tri_interval(const Tri32& T, float ohx, float ohy, float ohz, float olx, float oly, float olz, float Dx, float Dy, float Dz,
                                            float eo, float ed, float& lo, float& hi) {
    lo = 0.0f; hi = 0.0f;
    const float hx = Dy * T.e2[2] - Dz * T.e2[1], hy = Dz * T.e2[0] - Dx * T.e2[2], hz = Dx * T.e2[1] - Dy * T.e2[0];
    const float a = __fmaf_rn(T.e1[2], hz, __fmaf_rn(T.e1[1], hy, T.e1[0] * hx));
    const float g = 12.0f * kEps32, EE = T.E1 * T.E2;
    const float ea = (ed + g) * EE, A = fabsf(a);
    if (A + ea < 1e-6f) return kCandMiss;  // parallel: rejected upstream
    const float sx = (ohx - T.ah[0]) + (olx - T.al[0]), sy = (ohy - T.ah[1]) + (oly - T.al[1]), sz = (ohz - T.ah[2]) + (olz - T.al[2]);
    const float S = sqrt_approx(__fmaf_rn(sz, sz, __fmaf_rn(sy, sy, sx * sx))) * 1.00001f;
    const float sg = a < 0.0f ? -1.0f : 1.0f;
    const float U = sg * __fmaf_rn(sz, hz, __fmaf_rn(sy, hy, sx * hx));
    const float w = __fmaf_rn(S, ed + g, eo);
    const float eU = T.E2 * w, eV = T.E1 * w, eT = EE * __fmaf_rn(S, g, eo);
    // certainly outside one of the conditions: a miss whatever the others say (most rays leave here, before q is needed)
    if (U + eU < 0.0f || U - eU > A + ea) return kCandMiss;
    const float qx = sy * T.e1[2] - sz * T.e1[1], qy = sz * T.e1[0] - sx * T.e1[2], qz = sx * T.e1[1] - sy * T.e1[0];
    const float V = sg * __fmaf_rn(Dz, qz, __fmaf_rn(Dy, qy, Dx * qx));
    if (V + eV < 0.0f || (U + V) - (eU + eV) > A + ea) return kCandMiss;
    const float Tn = sg * __fmaf_rn(T.e2[2], qz, __fmaf_rn(T.e2[1], qy, T.e2[0] * qx));
    if (Tn + eT < 0.0f) return kCandMiss;
    const bool sure_a = A - ea > 1e-6f;
    if (sure_a && (Tn + eT) * 1.00001f < 1e-6f * (A - ea)) return kCandMiss;  // t <= 1e-6
    const bool inside = sure_a && U - eU > 0.0f && U + eU < A - ea && V - eV > 0.0f && (U + V) + (eU + eV) < A - ea;
    if (sure_a) {
        lo = fmaxf((Tn - eT) * rcp_approx(A + ea) * 0.999999f, 0.0f);
        hi = (Tn + eT) * rcp_approx(A - ea) * 1.000001f;
    } else {
        hi = INFINITY;
    }
    if (inside && lo > 1.00001e-6f && hi < 1.5f * lo + 1e-6f) return kCandHit;
    return kCandUnsure;  // a hit, if any, reports a distance >= lo
}

}  // namespace rte
