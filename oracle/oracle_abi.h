/*
 * oracle_abi.h — TEST INFRASTRUCTURE ONLY.
 *
 * C ABI shared by the two CPU checkers of the render path:
 *   oracle/_ref/libref_oracle.so  — the reference's own headers (/root/reference/RaytracingEngine),
 *                                   compiled unmodified behind oracle/ref_shim.cpp;
 *   oracle/liboracle_port.so      — oracle/oracle_port.c, a plain-C FP64 restatement of the same
 *                                   algorithm (every function cites the reference file:line).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load these libraries.  The product (librte_b200.so) never links or calls them.
 *
 * Both libraries export the same symbols with the prefix below so that a test can load both
 * and compare them bit for bit.
 */
#ifndef ORACLE_ABI_H_
#define ORACLE_ABI_H_

#include "../include/rte.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Which pixels to evaluate: rows row0, row0+row_step, ... < row1 ; all columns.
 * Output arrays are indexed by (row_slot * W + x) with row_slot = (y-row0)/row_step. */
typedef struct orc_opts {
    int32_t row0, row1, row_step;
    int32_t threads;      /* OpenMP threads; <= 0 = omp_get_max_threads() */
    int32_t max_depth;    /* port only (the reference hard-codes 10, RE/Scene.h:24); <= 0 = 10 */
    int32_t want_aovs;    /* fill the AOV / visibility outputs (slower) */
    double bias;          /* port only (reference: constexpr 1e-3, RE/Scene.h:291); <= 0 = 1e-3 */
} orc_opts;

typedef struct orc_outputs {
    double* hdr;          /* 3 doubles per pixel: GeneratePixelAt(x,y), RE/Scene.h:283-304 */
    uint32_t* prim_id;    /* primary hit (type<<30)|index, 0xFFFFFFFF = miss */
    double* depth;        /* HitInfo.distance, +inf on miss */
    double* normal;       /* 3 doubles, HitInfo.normal (un-flipped), 0 on miss */
    double* albedo;       /* 3 doubles, HitInfo.material.color; sky colour on miss */
    uint32_t* light_mask; /* bit j: transmittance > bias for light j at the primary hit;
                             bit 16+j: shadow ray cast (guards RE/Scene.h:89-101) */
    uint32_t* tie_mask;   /* bit j: the march of RE/Scene.h:35-77 took a `t <= bias` step for light j
                             (the "documented ties within bias");
                             bit 31: second-best primary hit within 1e-5 relative of the best */
    /* counters over the evaluated pixels */
    uint64_t primary_rays, shadow_rays, secondary_rays;
    uint64_t sphere_tests, plane_tests, triangle_tests; /* SURVEY.md §8(d) algorithmic counts */
    uint64_t closest_queries;  /* IntersectClosest calls actually made by the reference algorithm */
    double render_seconds;     /* wall clock of the pixel loop only (RE/RaytracingEngine.cpp:292-296) */
    int32_t threads_used;
    int32_t reserved;
} orc_outputs;

#ifndef ORC_PREFIX
#error "define ORC_PREFIX (ref_ or port_) before including oracle_abi.h"
#endif
#define ORC_CAT2(a, b) a##b
#define ORC_CAT(a, b) ORC_CAT2(a, b)
#define ORC_FN(name) ORC_CAT(ORC_PREFIX, name)

/* "reference" or "port" */
const char* ORC_FN(kind)(void);

/* Scene::RenderImage / GeneratePixelAt over the selected rows.  camera.anti_aliasing > 1: the port uses the counter-based
 * jitter stream of include/rte.h (bit-reproducible); the reference build uses upstream's own mt19937/random_device
 * jitter (RE/Math.h:103-116), so its multi-sample images are a STATISTICAL reference only. */
int ORC_FN(render)(const rte_scene_desc* scene, const rte_camera* camera, const orc_opts* opts,
                   orc_outputs* out);

/* Scene::IntersectClosest for a batch of rays (RE/Scene.h:218-257). */
int ORC_FN(intersect_closest)(const rte_scene_desc* scene, int64_t n, const double* origins,
                              const double* directions, rte_hit* hits);
/* Scene::IntersectAnyBefore (RE/Scene.h:259-276). */
int ORC_FN(intersect_any_before)(const rte_scene_desc* scene, int64_t n, const double* origins,
                                 const double* directions, const double* max_dist, uint8_t* out);

/* Camera::getRay(px,py,false), RE/Math.h:99-121: dir[3] out. */
void ORC_FN(camera_ray)(const rte_camera* camera, uint64_t px, uint64_t py, double* origin3,
                        double* dir3);

/* tonemapAll(), RE/RaytracingEngine.cpp:176-214 with the call-site constants (max_white 5.0,
 * a 0.18): hdr = 3 doubles per pixel, rgb8[op] = 3 bytes per pixel (NULL = skip).
 * hdr_out[op] (optional, 3 doubles per pixel) receives the operator result before toColor. */
int ORC_FN(tonemap_all)(const double* hdr, int64_t n_pixels, uint8_t* const rgb8[RTE_TM_COUNT],
                        double* const hdr_out[RTE_TM_COUNT]);
/* Parameterised variant (port only honours max_white / a different from the call-site values). */
int ORC_FN(tonemap_one)(int op, double max_white, double jodie_a, const double* hdr,
                        int64_t n_pixels, uint8_t* rgb8);

/* writePPM, RE/Image.cpp:11-31 */
int ORC_FN(write_ppm)(const char* filename, const uint8_t* rgb8, uint64_t width, uint64_t height);

#ifdef __cplusplus
}
#endif
#endif
