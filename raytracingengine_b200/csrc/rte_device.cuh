// rte_device.cuh — device-side scene layout and exact (FP64) primitive arithmetic.
//
// Design: the brute-force loops over primitives run in FP32 FMA as *conservative
// filters* over SoA arrays staged in shared memory; every candidate that passes a filter is
// re-evaluated here in FP64 with exactly the reference's operation order, so nearest-hit index,
// shadow visibility, depth and normals are the reference's own FP64 values.  This translation unit
// is compiled with -fmad=false: FP64 expressions below are never contracted (the x86 oracle is built
// with -ffp-contract=off), FP32 filter code asks for FMAs explicitly with __fmaf_rn.
//
// "RE/" = /root/reference/RaytracingEngine/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rte.h"

namespace rte {

// ------------------------------------------------------------------------------ device scene
struct DevPlane {   // Plane, RE/Shape.h:135-147: normal already normalised in FP64 (ctor :142)
    double n[3];
    double p0[3];
    int32_t mat;
    int32_t pad;
};
struct DevLight {   // Light, RE/Light.h:6-12; E = color * intensity (RE/Scene.h:110)
    double pos[3];
    double E[3];
};
struct DevTri64 {   // Triangle, RE/Shape.h:198-227, evaluated once on the host in FP64:
    double a0[3];   //   a0 = v0 + T
    double e1[3];   //   e1 = (v1 + T) - a0
    double e2[3];   //   e2 = (v2 + T) - a0
    double n[3];    //   normalize((v1 - v0) x (v2 - v0))  — from the untranslated vertices
};

// FP32 interval test of a triangle (rte_interval.cuh: tri_interval), built at upload next to DevTri64
struct __align__(16) Tri32 {
    float ah[3], al[3];   // a0 = ah + al
    float e1[3], e2[3];   // edges (FP64 differences, rounded)
    float n[3];           // unit normal (FP64, rounded): HitInfo.normal
    float E1, E2;         // upper bounds of |e1|, |e2|
    int mat;
    uint32_t report;      // HitInfo.index: the triangle's, or its MODEL's (RE/Shape.h:276)
    float pad;            // (80 bytes: five 16-byte loads from global memory)
};

struct DevScene {
    // FP32 filter arrays (sources of the TMA bulk copies into shared memory)
    const float4* sph32;  // {cx, cy, cz, r*r}; padded to a multiple of 32 with never-hit entries
    const float* views;   // (1+nl) per-origin views of the spheres (rte_views.cuh), rebuilt per camera / bias; may be NULL
    int32_t reflective;   // some material has specular > 0: reflection rays (generic closest-hit) will be traced
    int32_t nt_pad32;     // triangles padded to a multiple of 32 (one candidate mask per 32 bounding spheres)
    const float* tri_views;  // (1+nl) per-origin views of the triangles' bounding spheres, same format as `views`; may be NULL
    const float* views_cold;      // per view and sphere {f = C - O, ng}: read from global memory when a candidate is resolved
    const float* tri_views_cold;  // same for the triangles' bounding spheres
    const float4* tri32;  // 3 x float4 per triangle: {a0,_}, {e1,_}, {e2,_}; padded to a multiple of 2
    int32_t ns, ns_pad, nt, nt_pad;
    // FP64 exact arrays
    const double4* sph64; // {cx, cy, cz, r}
    const Tri32* tri_iv;  // FP32 path: one per (flattened) triangle
    const float4* sph_hl; // FP32 path: two float4 per sphere, {hi(C), |r|} and {lo(C), 0} with C = hi + lo to ~2^-48 (no FP64 -> FP32 conversion per candidate)
    const int32_t* sph_mat;
    const DevPlane* planes;
    const DevTri64* tri64;
    const int2* tri_info; // {material, reported index (triangle index or MODEL index, RE/Shape.h:276)}
    const rte_material* materials;  // nm entries + one default Material at index nm
    const DevLight* lights;
    int32_t np, nm, nl;
    int32_t transparent;  // some material has transparency > 0: refraction rays, transmittance marches through occluders
    // bounds used by the filter margins
    float sph_extent;  // max_i(|C_i| + r_i)
    float tri_extent;  // max |vertex|
    float tri_edge;    // max edge length
    float pad1;
};

struct DevCamera {  // Camera, RE/Math.h:85-97
    double pos[3];
    double focal;
    double half_w, half_h;  // double(width)/2.0, double(height)/2.0 (RE/Math.h:100-101)
    int32_t width, height;
    int32_t aa;             // Camera::antiAliasingAmount
    uint32_t jitter_seed;   // rte_camera.jitter_seed
};

// ------------------------------------------------------------------------------ Vec3 (FP64)
struct d3 { double x, y, z; };

__device__ __forceinline__ d3 mk(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ d3 ld3(const double* p) { return mk(p[0], p[1], p[2]); }
__device__ __forceinline__ d3 operator+(d3 a, d3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }   // RE/Math.h:14
__device__ __forceinline__ d3 operator-(d3 a, d3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }   // :16
__device__ __forceinline__ d3 operator*(d3 a, double s) { return mk(a.x * s, a.y * s, a.z * s); }      // :18
__device__ __forceinline__ d3 operator*(d3 a, d3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }   // :19
// (measured: out-of-line copies of the FP64 division / square root shrink the kernels by 12 % but lose more in calls than
// they win in instruction fetch on every config but one — they stay inline)
__device__ __forceinline__ double div64(double a, double b) { return a / b; }
__device__ __forceinline__ double sqrt64(double a) { return sqrt(a); }
__device__ __forceinline__ d3 operator/(d3 a, double s) { return mk(a.x / s, a.y / s, a.z / s); }      // :21
__device__ __forceinline__ d3 operator/(d3 a, d3 b) { return mk(div64(a.x, b.x), div64(a.y, b.y), div64(a.z, b.z)); }   // :20
__device__ __forceinline__ d3 operator-(d3 a) { return mk(-a.x, -a.y, -a.z); }                         // :23
__device__ __forceinline__ d3 adds(d3 a, double s) { return mk(a.x + s, a.y + s, a.z + s); }           // :15
__device__ __forceinline__ d3 subs(d3 a, double s) { return mk(a.x - s, a.y - s, a.z - s); }           // :17
__device__ __forceinline__ double dot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }        // :27
__device__ __forceinline__ d3 cross(d3 a, d3 b) {                                                       // :28
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ double length(d3 a) { return sqrt64(dot(a, a)); }                           // :30
// {a / |a|, |a|} — optionally as ONE out-of-line copy (RTE_UNIT_OL): a square root and three divisions are ~100 SASS
// instructions, and the render kernels normalise at five places per ray
struct d4 { double x, y, z, w; };
#ifndef RTE_UNIT_OL
#define RTE_UNIT_OL 1
#endif
#if RTE_UNIT_OL
static __device__ __noinline__ d4 unit_len(d3 a) {
#else
__device__ __forceinline__ d4 unit_len(d3 a) {
#endif
    d4 r;
    r.w = length(a);
    r.x = a.x / r.w; r.y = a.y / r.w; r.z = a.z / r.w;
    return r;
}
__device__ __forceinline__ d3 normalize(d3 a) {                                                         // :31-37
    const d4 u = unit_len(a);
    if (u.w <= 1e-12) return mk(0.0, 0.0, 0.0);
    return mk(u.x, u.y, u.z);
}
__device__ __forceinline__ d3 reflect(d3 i, d3 n) { return i - (n * 2.0) * dot(i, n); }                // :39-41
__device__ __forceinline__ double std_max(double a, double b) { return (a < b) ? b : a; }              // std::max
__device__ __forceinline__ double std_min(double a, double b) { return (b < a) ? b : a; }              // std::min
__device__ __forceinline__ double std_clamp(double v, double lo, double hi) { return (v < lo) ? lo : (hi < v) ? hi : v; }
// Vec3::refract, RE/Math.h:43-52 (both arguments are normalised again, like upstream)
__device__ __forceinline__ d3 refract(d3 self, d3 normal, double eta) {
    const d3 I = normalize(self), N = normalize(normal);
    const double cosi = std_clamp(dot(I, N), -1.0, 1.0);
    const double k = 1.0 - eta * eta * (1.0 - cosi * cosi);
    if (k < 0.0) return mk(0.0, 0.0, 0.0);
    return I * eta - N * (eta * cosi + sqrt64(k));
}

// the jitter stream of rte.h (rte_camera): sample s >= 1 of pixel y*W + x, axis 0 / 1
__device__ __forceinline__ double jitter01(uint32_t seed, int64_t pixel, int s, int axis) {
    unsigned long long z = (((unsigned long long)seed) << 32) ^ ((unsigned long long)pixel * 8192ull + (unsigned long long)(2 * s + axis));
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * 0x1.0p-53;
}
// Camera::getRay(px, py, aa), RE/Math.h:99-121; sample 0 is the unjittered ray (jitter = 0.0, RE/Scene.h:292)
__device__ __forceinline__ d3 camera_dir(const DevCamera& c, int px, int py, int sample = 0) {
    double sx = (double)px - c.half_w;
    double sy = c.half_h - (double)py;
    double jx = 0.0, jy = 0.0;
    if (sample > 0) {
        const double invAA = 1.0 / 1.0;  // 1.0 / static_cast<double>(bool aa), RE/Math.h:106
        const int64_t pixel = (int64_t)py * c.width + px;
        jx = jitter01(c.jitter_seed, pixel, sample, 0) * invAA;
        jy = jitter01(c.jitter_seed, pixel, sample, 1) * invAA;
    }
    sx += jx;
    sy += jy;
    const d3 screen = mk(sx, sy, c.pos[2] + c.focal);
    return normalize(screen - ld3(c.pos));
}

// ------------------------------------------------------------------------------ exact primitive tests
// Sphere::Intersect, RE/Shape.h:72-98
__device__ __forceinline__ bool sphere_exact(const double4 s, d3 o, d3 d, double& t_out) {
    const d3 oc = o - mk(s.x, s.y, s.z);
    const double a = dot(d, d);
    const double b = 2.0 * dot(oc, d);
    const double c = dot(oc, oc) - s.w * s.w;
    const double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return false;
    const double sq = sqrt64(disc);
    double t0 = div64(-b - sq, 2.0 * a);
    double t1 = div64(-b + sq, 2.0 * a);
    if (t0 > t1) { const double tmp = t0; t0 = t1; t1 = tmp; }
    double t = t0;
    if (t < 1e-6) { t = t1; if (t < 1e-6) return false; }
    t_out = t;
    return true;
}
// Plane::Intersect, RE/Shape.h:149-159
#ifndef RTE_PLANE_OL
#define RTE_PLANE_OL 0
#endif
#if RTE_PLANE_OL
static __device__ __noinline__ bool plane_exact(const DevPlane& p, d3 o, d3 d, double& t_out) {
#else
__device__ __forceinline__ bool plane_exact(const DevPlane& p, d3 o, d3 d, double& t_out) {
#endif
    const d3 n = ld3(p.n);
    const double denom = dot(n, d);
    if (fabs(denom) > 1e-6) {
        const d3 p0l0 = ld3(p.p0) - o;
        const double t = div64(dot(p0l0, n), denom);
        if (t >= 0.0) { t_out = t; return true; }
    }
    return false;
}
// Triangle::Intersect, RE/Shape.h:202-220 (a0/e1/e2 are the host-evaluated :204-206)
__device__ __forceinline__ bool tri_exact(const DevTri64& tr, d3 o, d3 d, double& t_out) {
    const d3 e1 = ld3(tr.e1), e2 = ld3(tr.e2);
    const d3 h = cross(d, e2);
    const double a = dot(e1, h);
    if (a > -1e-6 && a < 1e-6) return false;
    const double f = div64(1.0, a);
    const d3 s = o - ld3(tr.a0);
    const double u = f * dot(s, h);
    if (u < 0.0 || u > 1.0) return false;
    const d3 q = cross(s, e1);
    const double v = f * dot(d, q);
    if (v < 0.0 || u + v > 1.0) return false;
    const double t = f * dot(e2, q);
    if (t > 1e-6) { t_out = t; return true; }
    return false;
}

constexpr uint32_t kMiss = 0xFFFFFFFFu;
constexpr float kEps32 = 5.9604645e-8f;  // 2^-24

}  // namespace rte
