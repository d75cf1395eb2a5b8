/*
 * oracle_port.c — TEST INFRASTRUCTURE ONLY (see oracle_abi.h).
 *
 * Plain-C FP64 restatement of the reference's per-pixel render path.  Every function cites the
 * reference file:line it follows ("RE/" = /root/reference/RaytracingEngine/).  Floating-point
 * operations are written in the reference's evaluation order and the file is compiled with
 * -ffp-contract=off, so results are bit-identical to oracle/_ref/libref_oracle.so (the
 * reference's own headers); tests/test_oracle.py pins that, plus the known-answer vectors of
 * SURVEY.md §8(c).  Parity status: PINNED against the reference built in this container.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_PREFIX port_
#include "oracle_abi.h"

/* ---------------------------------------------------------------- Vec3, RE/Math.h:9-71 ------ */
typedef struct { double x, y, z; } v3;

static inline v3 V(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static inline v3 vp(const double* p) { return V(p[0], p[1], p[2]); }
static inline v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }          /* :14 */
static inline v3 adds(v3 a, double s) { return V(a.x + s, a.y + s, a.z + s); }            /* :15 */
static inline v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }          /* :16 */
static inline v3 subs(v3 a, double s) { return V(a.x - s, a.y - s, a.z - s); }            /* :17 */
static inline v3 mul(v3 a, double s) { return V(a.x * s, a.y * s, a.z * s); }             /* :18 */
static inline v3 mulv(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }         /* :19 */
static inline v3 divv(v3 a, v3 b) { return V(a.x / b.x, a.y / b.y, a.z / b.z); }         /* :20 */
static inline v3 divs(v3 a, double s) { return V(a.x / s, a.y / s, a.z / s); }            /* :21 */
static inline v3 neg(v3 a) { return V(-a.x, -a.y, -a.z); }                                /* :23 */
static inline double dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }        /* :27 */
static inline v3 cross(v3 a, v3 b) {                                                       /* :28 */
    return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline double length(v3 a) { return sqrt(dot(a, a)); }                              /* :30 */
static inline v3 normalize(v3 a) {                                                         /* :31-37 */
    double len = length(a);
    if (len <= 1e-12) return V(0, 0, 0);
    return divs(a, len);
}
static inline v3 reflect(v3 i, v3 n) { return sub(i, mul(mul(n, 2.0), dot(i, n))); }      /* :39-41 */
static inline double clampd(double v, double lo, double hi) { return (v < lo) ? lo : (hi < v) ? hi : v; } /* std::clamp */
static inline double maxd(double a, double b) { return (a < b) ? b : a; }                  /* std::max(a,b) */
static inline double mind(double a, double b) { return (b < a) ? b : a; }                  /* std::min(a,b) */
static v3 refract(v3 self, v3 normal, double eta) {                                        /* :43-52 */
    v3 I = normalize(self), N = normalize(normal);
    double cosi = clampd(dot(I, N), -1.0, 1.0);
    double k = 1.0 - eta * eta * (1.0 - cosi * cosi);
    if (k < 0.0) return V(0, 0, 0);
    return sub(mul(I, eta), mul(N, eta * cosi + sqrt(k)));
}

typedef struct { v3 o, d; } ray_t;                                                         /* Rayon, :78-83 */
static inline v3 point_at(ray_t r, double t) { return add(r.o, mul(r.d, t)); }            /* :82 */

/* the counter-based jitter stream include/rte.h defines for rte_camera (upstream: mt19937 seeded by random_device) */
static double jitter01(uint32_t seed, uint64_t pixel, int s, int axis) {
    uint64_t z = (((uint64_t)seed) << 32) ^ (pixel * 8192ull + (uint64_t)(2 * s + axis));
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * 0x1.0p-53;
}
/* Camera::getRay(px,py,aa), RE/Math.h:99-121: sample 0 is unjittered (jitter = 0.0); sample >= 1 adds [0,1)^2 * invAA */
static ray_t camera_ray_sample(const rte_camera* c, uint64_t px, uint64_t py, int sample) {
    double sx = ((double)px) - (double)c->width / 2.0;
    double sy = (double)c->height / 2.0 - ((double)py);
    double jx = 0.0, jy = 0.0;
    if (sample > 0) {
        const double invAA = 1.0 / (double)1;  /* 1.0 / static_cast<double>(bool aa), RE/Math.h:106 */
        const uint64_t pixel = py * c->width + px;
        jx = jitter01((uint32_t)c->jitter_seed, pixel, sample, 0) * invAA;
        jy = jitter01((uint32_t)c->jitter_seed, pixel, sample, 1) * invAA;
    }
    sx += jx;
    sy += jy;
    v3 pos = vp(c->position);
    v3 screen = V(sx, sy, pos.z + c->focal);
    ray_t r = {pos, normalize(sub(screen, pos))};
    return r;
}
static ray_t camera_ray(const rte_camera* c, uint64_t px, uint64_t py) { return camera_ray_sample(c, px, py, 0); }

/* ---------------------------------------------------------------- flattened scene ----------- */
typedef struct { v3 c; double r; int mat; } sph_t;
typedef struct { v3 p0, n; int mat; } pln_t;
typedef struct { v3 v0, v1, v2, T; int mat; uint32_t report_index; } tri_t;
typedef struct {
    int64_t ns, np, nt, nl, nm;
    sph_t* s; pln_t* p; tri_t* t;     /* t: standalone triangles then model triangles, canonical order */
    const rte_light* l;
    const rte_material* m;
    rte_material default_mat;
} scene_t;

static const rte_material* mat_of(const scene_t* sc, int idx) {
    return (idx >= 0 && idx < sc->nm) ? &sc->m[idx] : &sc->default_mat;
}

static scene_t* scene_build(const rte_scene_desc* d) {
    scene_t* sc = (scene_t*)calloc(1, sizeof(scene_t));
    sc->default_mat.shininess = 128.0;  /* Material defaults, RE/Shape.h:13-19 */
    sc->default_mat.refractive_index = 1.0;
    sc->ns = d->n_spheres; sc->np = d->n_planes; sc->nl = d->n_lights; sc->nm = d->n_materials;
    sc->l = d->lights; sc->m = d->materials;
    sc->s = (sph_t*)malloc(sizeof(sph_t) * (size_t)(sc->ns + 1));
    for (int64_t i = 0; i < sc->ns; ++i) {
        sc->s[i].c = vp(d->spheres[i].center); sc->s[i].r = d->spheres[i].radius; sc->s[i].mat = d->spheres[i].material;
    }
    sc->p = (pln_t*)malloc(sizeof(pln_t) * (size_t)(sc->np + 1));
    for (int64_t i = 0; i < sc->np; ++i) {
        sc->p[i].p0 = vp(d->planes[i].point);
        sc->p[i].n = normalize(vp(d->planes[i].normal));  /* Plane ctor, RE/Shape.h:142 */
        sc->p[i].mat = d->planes[i].material;
    }
    int64_t nt = d->n_triangles;
    for (int64_t i = 0; i < d->n_models; ++i) nt += d->models[i].index_count / 3;
    sc->nt = nt;
    sc->t = (tri_t*)malloc(sizeof(tri_t) * (size_t)(nt + 1));
    int64_t k = 0;
    for (int64_t i = 0; i < d->n_triangles; ++i, ++k) {
        const rte_triangle* t = &d->triangles[i];
        sc->t[k].v0 = vp(t->v0); sc->t[k].v1 = vp(t->v1); sc->t[k].v2 = vp(t->v2); sc->t[k].T = vp(t->translate);
        sc->t[k].mat = t->material; sc->t[k].report_index = (uint32_t)i;
    }
    /* Model::GetHitInfoAt walks the index triples in order and reports the MODEL index, RE/Shape.h:269-283 */
    for (int64_t m = 0; m < d->n_models; ++m) {
        const rte_model* mo = &d->models[m];
        const double* vb = d->model_vertices + 3 * mo->first_vertex;
        const int32_t* ib = d->model_indices + mo->first_index;
        for (int64_t i = 0; i + 2 < mo->index_count; i += 3, ++k) {
            sc->t[k].v0 = vp(vb + 3 * ib[i]); sc->t[k].v1 = vp(vb + 3 * ib[i + 1]); sc->t[k].v2 = vp(vb + 3 * ib[i + 2]);
            sc->t[k].T = vp(mo->translate); sc->t[k].mat = mo->material; sc->t[k].report_index = (uint32_t)m;
        }
    }
    return sc;
}
static void scene_free(scene_t* sc) { free(sc->s); free(sc->p); free(sc->t); free(sc); }

/* ---------------------------------------------------------------- primitive tests ----------- */
/* Sphere::Intersect, RE/Shape.h:72-98 */
static int sphere_intersect(const sph_t* s, ray_t ray, double* t_out) {
    v3 oc = sub(ray.o, s->c);
    double a = dot(ray.d, ray.d);
    double b = 2.0 * dot(oc, ray.d);
    double c = dot(oc, oc) - s->r * s->r;
    double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return 0;
    double sq = sqrt(disc);
    double t0 = (-b - sq) / (2.0 * a);
    double t1 = (-b + sq) / (2.0 * a);
    if (t0 > t1) { double tmp = t0; t0 = t1; t1 = tmp; }
    const double eps = 1e-6;
    double t = t0;
    if (t < eps) { t = t1; if (t < eps) return 0; }
    *t_out = t;
    return 1;
}
/* Plane::Intersect, RE/Shape.h:149-159 */
static int plane_intersect(const pln_t* p, ray_t ray, double* t_out) {
    double denom = dot(p->n, ray.d);
    if (fabs(denom) > 1e-6) {
        v3 p0l0 = sub(p->p0, ray.o);
        double t = dot(p0l0, p->n) / denom;
        if (t >= 0.0) { *t_out = t; return 1; }
    }
    return 0;
}
/* Triangle::Intersect, RE/Shape.h:202-220 (tv0..tv2 :198-200) */
static int tri_intersect(const tri_t* tr, ray_t ray, double* t_out) {
    const double EPSILON = 1e-6;
    v3 a0 = add(tr->v0, tr->T);
    v3 e1 = sub(add(tr->v1, tr->T), a0);
    v3 e2 = sub(add(tr->v2, tr->T), a0);
    v3 h = cross(ray.d, e2);
    double a = dot(e1, h);
    if (a > -EPSILON && a < EPSILON) return 0;
    double f = 1.0 / a;
    v3 s = sub(ray.o, a0);
    double u = f * dot(s, h);
    if (u < 0.0 || u > 1.0) return 0;
    v3 q = cross(s, e1);
    double v = f * dot(ray.d, q);
    if (v < 0.0 || u + v > 1.0) return 0;
    double t = f * dot(e2, q);
    if (t > EPSILON) { *t_out = t; return 1; }
    return 0;
}

/* HitInfo, RE/Shape.h:28-34 */
typedef struct { int type; double t; uint32_t index; int mat; v3 n, p; } hit_t;

/* Scene::IntersectClosest, RE/Scene.h:218-257: spheres, planes, triangles, models; strict `<` */
static int closest(const scene_t* sc, ray_t ray, hit_t* h) {
    int found = 0; double best = 0.0; int btype = 0; int64_t bi = -1; double t;
    for (int64_t i = 0; i < sc->ns; ++i)
        if (sphere_intersect(&sc->s[i], ray, &t) && (!found || t < best)) { found = 1; best = t; btype = RTE_HIT_SPHERE; bi = i; }
    for (int64_t i = 0; i < sc->np; ++i)
        if (plane_intersect(&sc->p[i], ray, &t) && (!found || t < best)) { found = 1; best = t; btype = RTE_HIT_PLANE; bi = i; }
    for (int64_t i = 0; i < sc->nt; ++i)
        if (tri_intersect(&sc->t[i], ray, &t) && (!found || t < best)) { found = 1; best = t; btype = RTE_HIT_TRIANGLE; bi = i; }
    if (!found) return 0;
    h->type = btype; h->t = best; h->p = point_at(ray, best);
    if (btype == RTE_HIT_SPHERE) {          /* GetHitInfoAt, RE/Shape.h:104-121 */
        h->index = (uint32_t)bi; h->mat = sc->s[bi].mat; h->n = normalize(sub(h->p, sc->s[bi].c));
    } else if (btype == RTE_HIT_PLANE) {    /* RE/Shape.h:165-180 */
        h->index = (uint32_t)bi; h->mat = sc->p[bi].mat; h->n = sc->p[bi].n;
    } else {                                /* RE/Shape.h:222-243: normal from the untranslated vertices */
        const tri_t* tr = &sc->t[bi];
        h->index = tr->report_index; h->mat = tr->mat;
        h->n = normalize(cross(sub(tr->v1, tr->v0), sub(tr->v2, tr->v0)));
    }
    return 1;
}

/* ---------------------------------------------------------------- counters ------------------ */
typedef struct {
    uint64_t shadow, secondary, sph, pl, tri, queries;
    int count_tests;   /* also enumerate the canonical-order shadow tests */
} cnt_t;

static void count_closest_tests(const scene_t* sc, cnt_t* c) {
    c->sph += (uint64_t)sc->ns; c->pl += (uint64_t)sc->np; c->tri += (uint64_t)sc->nt;
}
/* shadow ray: primitives in canonical order up to and including the first one whose own
 * Intersect() reports t in (bias, maxDist) */
static void count_shadow_tests(const scene_t* sc, ray_t r, double maxDist, double bias, cnt_t* c) {
    double t;
    for (int64_t i = 0; i < sc->ns; ++i) { c->sph++; if (sphere_intersect(&sc->s[i], r, &t) && t > bias && t < maxDist) return; }
    for (int64_t i = 0; i < sc->np; ++i) { c->pl++; if (plane_intersect(&sc->p[i], r, &t) && t > bias && t < maxDist) return; }
    for (int64_t i = 0; i < sc->nt; ++i) { c->tri++; if (tri_intersect(&sc->t[i], r, &t) && t > bias && t < maxDist) return; }
}

/* ---------------------------------------------------------------- shading ------------------- */
/* Scene::computeTransmittance, RE/Scene.h:35-77 */
static double transmittance(const scene_t* sc, ray_t ray, double maxDist, double bias, int* took_step, cnt_t* c) {
    double T = 1.0, traveled = 0.0;
    ray_t r = ray;
    int safety = 64;
    while (safety-- > 0 && T > 1e-4 && traveled < maxDist) {
        hit_t hit;
        c->queries++;
        if (!closest(sc, r, &hit)) break;
        const double t = hit.t;
        if (t <= 0.0) {
            r.o = add(r.o, mul(r.d, bias));
            traveled += bias;
            if (took_step) *took_step = 1;
            continue;
        }
        if (t <= bias) {
            r.o = add(point_at(r, t), mul(r.d, bias));
            traveled += t + bias;
            if (took_step) *took_step = 1;
            continue;
        }
        if (traveled + t >= maxDist) break;
        const double tr = clampd(mat_of(sc, hit.mat)->transparency, 0.0, 1.0);
        T *= tr;
        const v3 newOrigin = add(point_at(r, t), mul(r.d, bias));
        traveled += t + bias;
        r.o = newOrigin;
    }
    return clampd(T, 0.0, 1.0);
}

/* Scene::directLightning, RE/Scene.h:79-129.  light_mask/tie_mask (optional) as in oracle_abi.h */
static v3 direct_lighting(const scene_t* sc, const hit_t* hit, v3 viewDir, v3 normalIn, double bias, cnt_t* c,
                          uint32_t* light_mask, uint32_t* tie_mask) {
    const rte_material* material = mat_of(sc, hit->mat);
    v3 normal = normalize(normalIn);
    v3 diffuseAcc = V(0, 0, 0), specAcc = V(0, 0, 0);
    for (int64_t j = 0; j < sc->nl; ++j) {
        const rte_light* light = &sc->l[j];
        v3 vecToLight = sub(vp(light->position), hit->p);
        const double d = length(vecToLight);
        if (d <= 0.0) continue;
        v3 L = divs(vecToLight, d);
        const double ndl = maxd(0.0, dot(normal, L));
        if (ndl <= 0.0) continue;
        if (d <= bias) continue;
        ray_t shadowRay = {add(hit->p, mul(normal, bias)), L};
        c->shadow++;
        if (light_mask && j < 16) *light_mask |= 1u << (16 + j);
        if (c->count_tests) count_shadow_tests(sc, shadowRay, d - bias, bias, c);
        int took = 0;
        const double T = transmittance(sc, shadowRay, d - bias, bias, &took, c);
        if (tie_mask && took && j < 16) *tie_mask |= 1u << j;
        if (T <= bias) continue;
        if (light_mask && j < 16) *light_mask |= 1u << j;
        v3 emitted = mul(vp(light->color), light->intensity);
        v3 contribution = mul(mul(emitted, 1.0 / (d * d)), ndl);
        diffuseAcc = add(diffuseAcc, mul(contribution, T));
        if (material->transparency <= 0.0 && material->specular > 0.0) {
            v3 H = normalize(add(L, viewDir));
            double ndh = maxd(0.0, dot(normal, H));
            if (ndh > 0.0) {
                const double specFactor = pow(ndh, material->shininess);
                specAcc = add(specAcc, mul(mul(mul(emitted, 1.0 / (d * d)), specFactor), T));
            }
        }
    }
    v3 diffuse = mulv(vp(material->color), diffuseAcc);
    v3 specular = mul(specAcc, material->specular);
    return add(diffuse, specular);
}

/* Scene::backgroundColor, RE/Scene.h:30-33 */
static v3 background(ray_t ray) {
    double t = 0.5 * (normalize(ray.d).y + 1.0);
    return add(mul(V(1.0, 1.0, 1.0), 1.0 - t), mul(V(0.5, 0.7, 1.0), t));
}
/* Scene::fresnel, RE/Scene.h:26-28 */
static double fresnel(double cosTheta, double F0) { return F0 + (1.0 - F0) * pow(1.0 - cosTheta, 5.0); }

typedef struct {  /* what the AOV pass wants to know about the depth-0 hit */
    int have; hit_t hit; uint32_t light_mask, tie_mask;
} primary_t;

/* Scene::TraceRay, RE/Scene.h:131-198 */
static v3 trace(const scene_t* sc, ray_t ray, int depth, int max_depth, double bias, cnt_t* c, primary_t* prim) {
    if (depth >= max_depth) return background(ray);
    hit_t hit;
    c->queries++;
    if (depth > 0) c->secondary++;
    if (c->count_tests) count_closest_tests(sc, c);
    if (!closest(sc, ray, &hit)) return background(ray);
    const rte_material* material = mat_of(sc, hit.mat);
    const v3 incoming = normalize(ray.d);
    const int frontFace = dot(hit.n, incoming) < 0.0;
    const v3 normal = frontFace ? hit.n : neg(hit.n);
    const v3 viewDir = neg(incoming);
    const double cosTheta = maxd(0.0, dot(normal, viewDir));
    const double etaI = 1.0, etaT = material->refractive_index;
    const double f0 = pow((etaT - etaI) / (etaT + etaI), 2.0);
    double fresnelAmount = fresnel(cosTheta, f0);
    double transparency = clampd(material->transparency, 0.0, 1.0);

    uint32_t lm = 0, tm = 0;
    v3 localLight = direct_lighting(sc, &hit, viewDir, normal, bias, c, prim ? &lm : NULL, prim ? &tm : NULL);
    if (prim) { prim->have = 1; prim->hit = hit; prim->light_mask = lm; prim->tie_mask = tm; }
    v3 finalLight = V(0, 0, 0);
    if (transparency < 1.0) finalLight = add(finalLight, mul(localLight, 1.0 - transparency));
    if (transparency > 0.0) {
        const double eta = frontFace ? (etaI / etaT) : (etaT / etaI);
        v3 refractDir = refract(incoming, normal, eta);
        if (length(refractDir) > bias) {
            refractDir = normalize(refractDir);
            ray_t rr = {add(hit.p, mul(refractDir, bias * 1e2)), refractDir};
            v3 tc = trace(sc, rr, depth + 1, max_depth, bias, c, NULL);
            finalLight = add(finalLight, mul(tc, transparency * (1.0 - fresnelAmount)));
        } else {
            fresnelAmount = 1.0;
        }
    }
    const double reflectiveness = (transparency > 0.0) ? fresnelAmount : material->specular;
    if (reflectiveness > bias) {
        v3 reflectDir = normalize(reflect(incoming, normal));
        ray_t rr = {add(hit.p, mul(reflectDir, bias)), reflectDir};
        v3 rc = trace(sc, rr, depth + 1, max_depth, bias, c, NULL);
        finalLight = add(finalLight, mul(rc, reflectiveness));
    }
    return finalLight;
}

/* ---------------------------------------------------------------- frame loop ---------------- */
static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }

const char* port_kind(void) { return "port"; }

void port_camera_ray(const rte_camera* camera, uint64_t px, uint64_t py, double* o3, double* d3) {
    ray_t r = camera_ray(camera, px, py);
    o3[0] = r.o.x; o3[1] = r.o.y; o3[2] = r.o.z; d3[0] = r.d.x; d3[1] = r.d.y; d3[2] = r.d.z;
}

/* near-tie flag: second-best primary candidate within 1e-5 relative of the best */
static int primary_near_tie(const scene_t* sc, ray_t ray, double best) {
    double second = INFINITY, t; int seen = 0;
    for (int64_t i = 0; i < sc->ns + sc->np + sc->nt; ++i) {
        int ok = i < sc->ns ? sphere_intersect(&sc->s[i], ray, &t)
               : i < sc->ns + sc->np ? plane_intersect(&sc->p[i - sc->ns], ray, &t)
                                     : tri_intersect(&sc->t[i - sc->ns - sc->np], ray, &t);
        if (!ok) continue;
        if (t == best && !seen) seen = 1; else if (t < second) second = t;
    }
    return second - best <= 1e-5 * fabs(best);
}

/* Scene::RenderImage / GeneratePixelAt, RE/Scene.h:283-328 (camera.anti_aliasing samples, jitter stream of rte.h) */
int port_render(const rte_scene_desc* scene, const rte_camera* camera, const orc_opts* opts, orc_outputs* out) {
    if (!scene || !camera || !opts || !out) return RTE_ERR_INVALID;
    scene_t* sc = scene_build(scene);
    const int W = (int)camera->width, H = (int)camera->height;
    const int row0 = opts->row0 > 0 ? opts->row0 : 0;
    const int row1 = (opts->row1 > 0 && opts->row1 < H) ? opts->row1 : H;
    const int step = opts->row_step > 0 ? opts->row_step : 1;
    const int nrows = row1 > row0 ? (row1 - row0 + step - 1) / step : 0;
    const int max_depth = opts->max_depth > 0 ? opts->max_depth : 10;   /* RE/Scene.h:24 */
    const double bias = opts->bias > 0 ? opts->bias : 1e-3;             /* RE/Scene.h:291 */
    const int aov = opts->want_aovs;
    const int aa = camera->anti_aliasing > 1 ? camera->anti_aliasing : 1;
#ifdef _OPENMP
    if (opts->threads > 0) omp_set_num_threads(opts->threads);
    out->threads_used = omp_get_max_threads();
#else
    out->threads_used = 1;
#endif
    uint64_t shadow = 0, secondary = 0, sph = 0, pl = 0, tri = 0, queries = 0;
    const int64_t total = (int64_t)nrows * W;
    const double t0 = now_s();
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : shadow, secondary, sph, pl, tri, queries)
#endif
    for (int64_t idx = 0; idx < total; ++idx) {
        const int x = (int)(idx % W), y = row0 + (int)(idx / W) * step;
        cnt_t c; memset(&c, 0, sizeof(c)); c.count_tests = aov;
        primary_t prim; prim.have = 0;
        const ray_t ray = camera_ray(camera, (uint64_t)x, (uint64_t)y);
        /* GeneratePixelAt, RE/Scene.h:283-304: accumulatedColor(0) += colour of every sample; /= samples */
        v3 acc = V(0, 0, 0);
        for (int s = 0; s < aa; ++s) {
            const ray_t rs = camera_ray_sample(camera, (uint64_t)x, (uint64_t)y, s);
            v3 color = trace(sc, rs, 0, max_depth, bias, &c, (aov && s == 0) ? &prim : NULL);
            acc = add(acc, color);
        }
        acc = divs(acc, (double)aa);
        if (out->hdr) { out->hdr[3 * idx] = acc.x; out->hdr[3 * idx + 1] = acc.y; out->hdr[3 * idx + 2] = acc.z; }
        if (aov) {
            uint32_t pid = 0xFFFFFFFFu, lm = 0, tm = 0; double depth = INFINITY; v3 n = V(0, 0, 0), alb = background(ray);
            if (prim.have) {
                pid = ((uint32_t)prim.hit.type << 30) | prim.hit.index; depth = prim.hit.t; n = prim.hit.n;
                alb = vp(mat_of(sc, prim.hit.mat)->color); lm = prim.light_mask; tm = prim.tie_mask;
                if (primary_near_tie(sc, ray, prim.hit.t)) tm |= 1u << 31;
            }
            if (out->prim_id) out->prim_id[idx] = pid;
            if (out->depth) out->depth[idx] = depth;
            if (out->normal) { out->normal[3 * idx] = n.x; out->normal[3 * idx + 1] = n.y; out->normal[3 * idx + 2] = n.z; }
            if (out->albedo) { out->albedo[3 * idx] = alb.x; out->albedo[3 * idx + 1] = alb.y; out->albedo[3 * idx + 2] = alb.z; }
            if (out->light_mask) out->light_mask[idx] = lm;
            if (out->tie_mask) out->tie_mask[idx] = tm;
        }
        shadow += c.shadow; secondary += c.secondary; sph += c.sph; pl += c.pl; tri += c.tri; queries += c.queries;
    }
    out->render_seconds = now_s() - t0;
    out->primary_rays = (uint64_t)total * (uint64_t)aa; out->shadow_rays = shadow; out->secondary_rays = secondary;
    out->sphere_tests = sph; out->plane_tests = pl; out->triangle_tests = tri; out->closest_queries = queries;
    scene_free(sc);
    return RTE_OK;
}

int port_intersect_closest(const rte_scene_desc* scene, int64_t n, const double* origins, const double* directions, rte_hit* hits) {
    scene_t* sc = scene_build(scene);
    for (int64_t i = 0; i < n; ++i) {
        ray_t r = {vp(origins + 3 * i), vp(directions + 3 * i)};
        hit_t h; rte_hit* o = &hits[i];
        memset(o, 0, sizeof(*o));
        if (!closest(sc, r, &h)) { o->prim_id = 0xFFFFFFFFu; o->distance = INFINITY; o->material = -1; continue; }
        o->distance = h.t; o->normal[0] = h.n.x; o->normal[1] = h.n.y; o->normal[2] = h.n.z;
        o->point[0] = h.p.x; o->point[1] = h.p.y; o->point[2] = h.p.z;
        o->prim_id = ((uint32_t)h.type << 30) | h.index; o->material = h.mat;
    }
    scene_free(sc);
    return RTE_OK;
}

/* Scene::IntersectAnyBefore, RE/Scene.h:259-276 (a Model reports its closest triangle, RE/Shape.h:285-300;
 * "any model-min within" == "any triangle within" because within() is an interval test on t > 0 only when
 * the minimum qualifies: min t in (0,maxDist) <=> exists t in (0,maxDist) since all reported t > 0) */
int port_intersect_any_before(const rte_scene_desc* scene, int64_t n, const double* origins, const double* directions,
                              const double* max_dist, uint8_t* out) {
    scene_t* sc = scene_build(scene);
    for (int64_t i = 0; i < n; ++i) {
        ray_t r = {vp(origins + 3 * i), vp(directions + 3 * i)};
        double t; int any = 0;
        for (int64_t k = 0; k < sc->ns && !any; ++k) any = sphere_intersect(&sc->s[k], r, &t) && t > 0.0 && t < max_dist[i];
        for (int64_t k = 0; k < sc->np && !any; ++k) any = plane_intersect(&sc->p[k], r, &t) && t > 0.0 && t < max_dist[i];
        for (int64_t k = 0; k < sc->nt && !any; ++k) any = tri_intersect(&sc->t[k], r, &t) && t > 0.0 && t < max_dist[i];
        out[i] = (uint8_t)any;
    }
    scene_free(sc);
    return RTE_OK;
}

/* ---------------------------------------------------------------- tonemap operators --------- */
/* ClampVec3, RE/RaytracingEngine.cpp:70-76: std::min(maxVal, std::max(minVal, v)) — NaN -> minVal */
static inline double clamp_mm(double v, double lo, double hi) { return mind(hi, maxd(lo, v)); }
static v3 clamp_vec3(v3 v, double lo, double hi) { return V(clamp_mm(v.x, lo, hi), clamp_mm(v.y, lo, hi), clamp_mm(v.z, lo, hi)); }

/* uncharted2_tonemap_partial, RE/RaytracingEngine.cpp:78-87: float constants; C*B, D*E, D*F, E/F are
 * evaluated in float and widened */
static v3 uncharted2_partial(v3 x) {
    float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    float CB = C * B, DE = D * E, DF = D * F, EF = E / F;
    v3 num = adds(mulv(x, adds(mul(x, (double)A), (double)CB)), (double)DE);
    v3 den = adds(mulv(x, adds(mul(x, (double)A), (double)B)), (double)DF);
    return subs(divv(num, den), (double)EF);
}
/* aces_approx, RE/RaytracingEngine.cpp:89-98 */
static v3 aces_approx(v3 v) {
    v = mul(v, (double)0.6f);
    float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    v3 num = mulv(v, adds(mul(v, (double)a), (double)b));
    v3 den = adds(mulv(v, adds(mul(v, (double)c), (double)d)), (double)e);
    return clamp_vec3(divv(num, den), (double)0.0f, (double)1.0f);
}
/* luminance / change_luminance, RE/RaytracingEngine.cpp:100-110 */
static double luminance(v3 c) { return dot(c, V(0.2126, 0.7152, 0.0722)); }
static v3 change_luminance(v3 c_in, double l_out) { double l_in = luminance(c_in); return mul(c_in, l_out / l_in); }
/* simple, :123-131 */
static v3 tm_simple(v3 c) { return V(mind(1.0, maxd(0.0, c.x)), mind(1.0, maxd(0.0, c.y)), mind(1.0, maxd(0.0, c.z))); }
/* reinhardSimple, :133-135 */
static v3 tm_reinhard(v3 c) { return divv(c, adds(c, 1)); }
/* reinhardExtended, :137-141 */
static v3 tm_reinhard_ext(v3 c, double max_white) {
    double wsq = max_white * max_white;
    v3 numerator = mulv(c, adds(divv(c, V(wsq, wsq, wsq)), 1));
    return divv(numerator, adds(c, 1));
}
/* reinhardExtendedLuminance, :143-148 */
static v3 tm_reinhard_ext_lum(v3 c, double maxWhite) {
    double L_old = luminance(c);
    double numerator = L_old * (1 + (L_old / (maxWhite * maxWhite)));
    double l_new = numerator / (1 + L_old);
    return change_luminance(c, l_new);
}
/* reinhardJodie, :150-154 */
static v3 tm_jodie(v3 c, double a) {
    double L = luminance(c);
    double L_mapped = (a / log(2 + pow((L / 0.85), 1.7))) * log(1 + L);
    return change_luminance(c, L_mapped);
}
/* uncharted2, :156-163 */
static v3 tm_uncharted2(v3 c) {
    double exposureBias = 2.0f;
    v3 curr = uncharted2_partial(mul(c, exposureBias));
    v3 W = V(11.2, 11.2, 11.2);
    v3 whiteScale = divv(V(1, 1, 1), uncharted2_partial(W));
    return mulv(curr, whiteScale);
}
/* toColor, :113-121: clamp then truncate */
static void to_color(v3 p, uint8_t* out) {
    v3 c = clamp_vec3(p, 0.0, 1.0);
    out[0] = (uint8_t)(c.x * 255.0); out[1] = (uint8_t)(c.y * 255.0); out[2] = (uint8_t)(c.z * 255.0);
}
static v3 tm_apply(int op, v3 c, double max_white, double a) {
    switch (op) {
        case RTE_TM_SIMPLE: return tm_simple(c);
        case RTE_TM_REINHARD_SIMPLE: return tm_reinhard(c);
        case RTE_TM_REINHARD_EXTENDED: return tm_reinhard_ext(c, max_white);
        case RTE_TM_REINHARD_EXTENDED_LUMINANCE: return tm_reinhard_ext_lum(c, max_white);
        case RTE_TM_REINHARD_JODIE: return tm_jodie(c, a);
        case RTE_TM_UNCHARTED2: return tm_uncharted2(c);
        default: return aces_approx(c);
    }
}

/* tonemapAll, RE/RaytracingEngine.cpp:176-214 (max_white 5.0 :193,195; Jodie default a 0.18 :150) */
int port_tonemap_all(const double* hdr, int64_t n, uint8_t* const rgb8[RTE_TM_COUNT], double* const hdr_out[RTE_TM_COUNT]) {
    for (int64_t i = 0; i < n; ++i) {
        v3 c = vp(hdr + 3 * i);
        for (int op = 0; op < RTE_TM_COUNT; ++op) {
            v3 m = tm_apply(op, c, 5.0, 0.18);
            if (rgb8 && rgb8[op]) to_color(m, rgb8[op] + 3 * i);
            if (hdr_out && hdr_out[op]) { hdr_out[op][3 * i] = m.x; hdr_out[op][3 * i + 1] = m.y; hdr_out[op][3 * i + 2] = m.z; }
        }
    }
    return RTE_OK;
}
int port_tonemap_one(int op, double max_white, double jodie_a, const double* hdr, int64_t n, uint8_t* rgb8) {
    if (op < 0 || op >= RTE_TM_COUNT) return RTE_ERR_INVALID;
    for (int64_t i = 0; i < n; ++i) to_color(tm_apply(op, vp(hdr + 3 * i), max_white, jodie_a), rgb8 + 3 * i);
    return RTE_OK;
}

/* writePPM, RE/Image.cpp:11-31 */
int port_write_ppm(const char* filename, const uint8_t* rgb8, uint64_t width, uint64_t height) {
    FILE* f = fopen(filename, "wb");
    if (!f) return RTE_ERR_IO;
    fprintf(f, "P6\n%llu %llu\n255\n", (unsigned long long)width, (unsigned long long)height);
    size_t n = (size_t)(width * height * 3);
    int ok = fwrite(rgb8, 1, n, f) == n;
    ok = (fclose(f) == 0) && ok;
    return ok ? RTE_OK : RTE_ERR_IO;
}
