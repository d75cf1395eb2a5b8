// ref_shim.cpp — TEST INFRASTRUCTURE ONLY (see oracle_abi.h).
//
// Puts the oracle C ABI in front of the reference's own, unmodified sources.  Nothing is copied:
// the reference files are #included from where they lie (-I/root/reference/RaytracingEngine,
// see oracle/Makefile) and the build output goes to oracle/_ref/ only.
//
//   * `#define private public` is applied to the reference headers only (every standard header
//     they use is included first), so that the shim can call Scene::computeTransmittance
//     (RE/Scene.h:35) and walk Scene's primitive vectors (RE/Scene.h:17-21).  The reference
//     code that runs is byte-for-byte the upstream code.
//   * `#define main reference_main` + #include "RaytracingEngine.cpp" gives external-linkage
//     tonemapAll / the seven operators / toColor / LoadObject.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <functional>
#include <iomanip>
#include <iostream>
#include <limits>
#include <optional>
#include <random>
#include <ranges>
#include <stdexcept>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#define private public
#define main reference_main
#include "RaytracingEngine.cpp"  // from /root/reference/RaytracingEngine (includes Math/Light/Shape/Scene/Image.h)
#undef main
#undef private

#define ORC_PREFIX ref_
#include "oracle_abi.h"

namespace {

Vec3 v3(const double* p) { return Vec3(p[0], p[1], p[2]); }

Material to_material(const rte_scene_desc* s, int32_t idx) {
    Material m;  // defaults of RE/Shape.h:13-19
    if (idx >= 0 && idx < s->n_materials) {
        const rte_material& r = s->materials[idx];
        m.color = v3(r.color);
        m.shininess = r.shininess;
        m.specular = r.specular;
        m.transparency = r.transparency;
        m.refractiveIndex = r.refractive_index;
    }
    return m;
}

struct Built {
    Scene scene;
    std::vector<std::vector<Triangle>> model_tris;  // GetTrianglesFromModel per model
    explicit Built(const Camera& c) : scene(c) {}
};

Camera to_camera(const rte_camera* c) {
    Camera cam(v3(c->position), c->focal, (size_t)c->width, (size_t)c->height, c->near_plane,
               c->far_plane);
    cam.antiAliasingAmount = c->anti_aliasing > 1 ? c->anti_aliasing : 1;  // 1 = deterministic; > 1: upstream's own random jitter
    return cam;
}

Built* build(const rte_scene_desc* s, const rte_camera* c) {
    Built* b = new Built(to_camera(c));
    for (int64_t i = 0; i < s->n_spheres; ++i) {
        Sphere sp(s->spheres[i].radius, v3(s->spheres[i].center), to_material(s, s->spheres[i].material));
        b->scene.AddSphere(sp);
    }
    for (int64_t i = 0; i < s->n_planes; ++i) {
        Plane pl(v3(s->planes[i].point), v3(s->planes[i].normal), to_material(s, s->planes[i].material));
        b->scene.AddPlane(pl);
    }
    for (int64_t i = 0; i < s->n_triangles; ++i) {
        const rte_triangle& t = s->triangles[i];
        Transform tr{v3(t.translate), Vec3(0, 0, 0), Vec3(1, 1, 1)};
        Triangle tri(v3(t.v0), v3(t.v1), v3(t.v2), to_material(s, t.material), tr);
        b->scene.AddTriangle(tri);
    }
    for (int64_t i = 0; i < s->n_models; ++i) {
        const rte_model& m = s->models[i];
        std::vector<Vec3> pos;
        pos.reserve((size_t)m.vertex_count);
        for (int64_t v = 0; v < m.vertex_count; ++v) pos.push_back(v3(s->model_vertices + 3 * (m.first_vertex + v)));
        std::vector<int> idx(s->model_indices + m.first_index, s->model_indices + m.first_index + m.index_count);
        Transform tr{v3(m.translate), Vec3(0, 0, 0), Vec3(1, 1, 1)};
        Material mat = to_material(s, m.material);
        Model model(idx, tr, mat, pos);
        b->scene.AddModel(model);
        b->model_tris.push_back(model.GetTrianglesFromModel(mat));
    }
    for (int64_t i = 0; i < s->n_lights; ++i) {
        Light l(v3(s->lights[i].position), v3(s->lights[i].color), s->lights[i].intensity);
        b->scene.AddLight(l);
    }
    return b;
}

constexpr double kBias = 1e-3;  // RE/Scene.h:291

uint32_t prim_id_of(const HitInfo& h) { return ((uint32_t)h.type << 30) | (uint32_t)h.index; }

// Visit every primitive's own Intersect() result in canonical order (RE/Scene.h:218-257:
// spheres, planes, triangles, model triangles).  fn(kind, t_opt) returns false to stop.
template <class F>
void for_each_primitive(const Built& b, const Rayon& ray, F&& fn) {
    for (const auto& s : b.scene.spheres) if (!fn(0, s.Intersect(ray))) return;
    for (const auto& p : b.scene.planes) if (!fn(1, p.Intersect(ray))) return;
    for (const auto& t : b.scene.triangles) if (!fn(2, t.Intersect(ray))) return;
    for (const auto& mt : b.model_tris)
        for (const auto& t : mt) if (!fn(2, t.Intersect(ray))) return;
}

// Re-derivation of the RE/Scene.h:35-77 march with the public IntersectClosest, only to find
// out whether a `t <= bias` / `t <= 0` step was taken (tie flag) and how many queries it made.
bool march_took_bias_step(const Scene& sc, Rayon r, double maxDist, uint64_t& queries) {
    double T = 1.0, traveled = 0.0;
    int safety = 64;
    bool stepped = false;
    while (safety-- > 0 && T > 1e-4 && traveled < maxDist) {
        ++queries;
        auto h = sc.IntersectClosest(r);
        if (!h) break;
        const double t = h->distance;
        if (t <= 0.0) { r.origin = r.origin + r.direction * kBias; traveled += kBias; stepped = true; continue; }
        if (t <= kBias) { r.origin = r.pointAtDistance(t) + r.direction * kBias; traveled += t + kBias; stepped = true; continue; }
        if (traveled + t >= maxDist) break;
        T *= std::clamp(h->material.transparency, 0.0, 1.0);
        r.origin = r.pointAtDistance(t) + r.direction * kBias;
        traveled += t + kBias;
    }
    return stepped;
}

struct Counters {
    uint64_t primary = 0, shadow = 0, sph = 0, pl = 0, tri = 0, queries = 0;
};

void pixel_aovs(const Built& b, const rte_scene_desc* s, size_t x, size_t y, size_t slot, orc_outputs* out,
                Counters& cnt) {
    const Scene& sc = b.scene;
    const Rayon ray = sc.camera.getRay(x, y, false);
    const auto hitOpt = sc.IntersectClosest(ray);  // == CalculatePixelDepth(x,y,false), RE/Scene.h:278-281
    cnt.primary += 1;
    cnt.queries += 1;
    cnt.sph += sc.spheres.size();
    cnt.pl += sc.planes.size();
    uint64_t ntri = sc.triangles.size();
    for (const auto& mt : b.model_tris) ntri += mt.size();
    cnt.tri += ntri;

    uint32_t pid = 0xFFFFFFFFu, lmask = 0, tmask = 0;
    double depth = std::numeric_limits<double>::infinity();
    Vec3 n(0, 0, 0), alb = sc.backgroundColor(ray);
    if (hitOpt) {
        const HitInfo& hit = *hitOpt;
        pid = prim_id_of(hit);
        depth = hit.distance;
        n = hit.normal;
        alb = hit.material.color;
        // near-tie flag for the primary hit
        double second = std::numeric_limits<double>::infinity();
        int seen_best = 0;
        for_each_primitive(b, ray, [&](int, std::optional<double> t) {
            if (t) {
                if (*t == hit.distance && !seen_best) seen_best = 1;
                else second = std::min(second, *t);
            }
            return true;
        });
        if (second - hit.distance <= 1e-5 * std::abs(hit.distance)) tmask |= 1u << 31;

        // guards of directLightning, RE/Scene.h:79-108, with the reference's own vector ops
        const Vec3 incoming = ray.direction.normalize();
        const bool frontFace = hit.normal.dot(incoming) < 0.0;
        const Vec3 normalIn = frontFace ? hit.normal : -hit.normal;
        const Vec3 normal = normalIn.normalize();
        for (size_t j = 0; j < sc.lights.size() && j < 16; ++j) {
            const Light& light = sc.lights[j];
            Vec3 vecToLight = light.position - hit.hitPoint;
            const double d = vecToLight.length();
            if (d <= 0.0) continue;
            Vec3 L = vecToLight / d;
            const double ndl = std::max(0.0, normal.dot(L));
            if (ndl <= 0.0) continue;
            if (d <= kBias) continue;
            Rayon shadowRay{hit.hitPoint + normal * kBias, L};
            lmask |= 1u << (16 + j);
            cnt.shadow += 1;
            const double T = sc.computeTransmittance(shadowRay, d - kBias, kBias);  // the reference's own
            if (T > kBias) lmask |= 1u << j;
            if (march_took_bias_step(sc, shadowRay, d - kBias, cnt.queries)) tmask |= 1u << j;
            // algorithmic test count: canonical order up to and including the first occluder
            const double maxDist = d - kBias;
            for_each_primitive(b, shadowRay, [&](int kind, std::optional<double> t) {
                (kind == 0 ? cnt.sph : kind == 1 ? cnt.pl : cnt.tri) += 1;
                return !(t && *t > kBias && *t < maxDist);
            });
        }
    }
    (void)s;
    if (out->prim_id) out->prim_id[slot] = pid;
    if (out->depth) out->depth[slot] = depth;
    if (out->normal) { out->normal[3 * slot] = n.x; out->normal[3 * slot + 1] = n.y; out->normal[3 * slot + 2] = n.z; }
    if (out->albedo) { out->albedo[3 * slot] = alb.x; out->albedo[3 * slot + 1] = alb.y; out->albedo[3 * slot + 2] = alb.z; }
    if (out->light_mask) out->light_mask[slot] = lmask;
    if (out->tie_mask) out->tie_mask[slot] = tmask;
}

}  // namespace

extern "C" {

const char* ref_kind(void) { return "reference"; }

int ref_render(const rte_scene_desc* scene, const rte_camera* camera, const orc_opts* opts, orc_outputs* out) {
    if (!scene || !camera || !opts || !out) return RTE_ERR_INVALID;
    Built* b = build(scene, camera);
    const int W = (int)camera->width, H = (int)camera->height;
    const int row0 = std::max(0, opts->row0);
    const int row1 = opts->row1 > 0 ? std::min(H, opts->row1) : H;
    const int step = std::max(1, opts->row_step);
    const int nrows = row1 > row0 ? (row1 - row0 + step - 1) / step : 0;
#ifdef _OPENMP
    if (opts->threads > 0) omp_set_num_threads(opts->threads);
    out->threads_used = omp_get_max_threads();
#else
    out->threads_used = 1;
#endif
    const bool full = (row0 == 0 && row1 == H && step == 1);
    const auto t0 = std::chrono::high_resolution_clock::now();
    if (out->hdr) {
        if (full) {
            // the reference's own frame loop, RE/Scene.h:311-328
            std::vector<Vec3> img = b->scene.RenderImage();
            for (size_t i = 0; i < img.size(); ++i) { out->hdr[3 * i] = img[i].x; out->hdr[3 * i + 1] = img[i].y; out->hdr[3 * i + 2] = img[i].z; }
        } else {
            // same loop shape as RenderImage (dynamic schedule, chunk 1) over the selected rows
            const int total = nrows * W;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic)
#endif
            for (int idx = 0; idx < total; ++idx) {
                const int x = idx % W, slot_row = idx / W;
                const Vec3 c = b->scene.GeneratePixelAt(x, row0 + slot_row * step);  // RE/Scene.h:283
                out->hdr[3 * (size_t)idx] = c.x; out->hdr[3 * (size_t)idx + 1] = c.y; out->hdr[3 * (size_t)idx + 2] = c.z;
            }
        }
    }
    const auto t1 = std::chrono::high_resolution_clock::now();
    out->render_seconds = std::chrono::duration<double>(t1 - t0).count();

    out->primary_rays = out->shadow_rays = out->secondary_rays = 0;
    out->sphere_tests = out->plane_tests = out->triangle_tests = out->closest_queries = 0;
    if (opts->want_aovs) {
        // counters of the shim cover the primary hit only (depth 0); the port counts all depths
        uint64_t primary = 0, shadow = 0, sph = 0, pl = 0, tri = 0, queries = 0;
        const int total = nrows * W;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : primary, shadow, sph, pl, tri, queries)
#endif
        for (int idx = 0; idx < total; ++idx) {
            Counters c;
            pixel_aovs(*b, scene, (size_t)(idx % W), (size_t)(row0 + (idx / W) * step), (size_t)idx, out, c);
            primary += c.primary; shadow += c.shadow; sph += c.sph; pl += c.pl; tri += c.tri; queries += c.queries;
        }
        out->primary_rays = primary; out->shadow_rays = shadow;
        out->sphere_tests = sph; out->plane_tests = pl; out->triangle_tests = tri; out->closest_queries = queries;
    }
    delete b;
    return RTE_OK;
}

int ref_intersect_closest(const rte_scene_desc* scene, int64_t n, const double* origins, const double* directions,
                          rte_hit* hits) {
    rte_camera cam{};
    cam.width = cam.height = 1; cam.focal = 1; cam.near_plane = 1; cam.far_plane = 1000;
    Built* b = build(scene, &cam);
    for (int64_t i = 0; i < n; ++i) {
        Rayon r{v3(origins + 3 * i), v3(directions + 3 * i)};
        auto h = b->scene.IntersectClosest(r);
        rte_hit& o = hits[i];
        std::memset(&o, 0, sizeof(o));
        if (!h) { o.prim_id = 0xFFFFFFFFu; o.distance = std::numeric_limits<double>::infinity(); o.material = -1; continue; }
        o.distance = h->distance;
        o.normal[0] = h->normal.x; o.normal[1] = h->normal.y; o.normal[2] = h->normal.z;
        o.point[0] = h->hitPoint.x; o.point[1] = h->hitPoint.y; o.point[2] = h->hitPoint.z;
        o.prim_id = prim_id_of(*h);
        o.material = -1;  // the reference carries materials by value
    }
    delete b;
    return RTE_OK;
}

int ref_intersect_any_before(const rte_scene_desc* scene, int64_t n, const double* origins, const double* directions,
                             const double* max_dist, uint8_t* out) {
    rte_camera cam{};
    cam.width = cam.height = 1; cam.focal = 1; cam.near_plane = 1; cam.far_plane = 1000;
    Built* b = build(scene, &cam);
    for (int64_t i = 0; i < n; ++i) {
        Rayon r{v3(origins + 3 * i), v3(directions + 3 * i)};
        out[i] = b->scene.IntersectAnyBefore(r, max_dist[i]) ? 1 : 0;
    }
    delete b;
    return RTE_OK;
}

void ref_camera_ray(const rte_camera* camera, uint64_t px, uint64_t py, double* origin3, double* dir3) {
    const Camera cam = to_camera(camera);
    const Rayon r = cam.getRay((size_t)px, (size_t)py, false);
    origin3[0] = r.origin.x; origin3[1] = r.origin.y; origin3[2] = r.origin.z;
    dir3[0] = r.direction.x; dir3[1] = r.direction.y; dir3[2] = r.direction.z;
}

int ref_tonemap_all(const double* hdr, int64_t n_pixels, uint8_t* const rgb8[RTE_TM_COUNT],
                    double* const hdr_out[RTE_TM_COUNT]) {
    std::vector<Vec3> px((size_t)n_pixels);
    for (int64_t i = 0; i < n_pixels; ++i) px[(size_t)i] = v3(hdr + 3 * i);
    const auto all = tonemapAll(px);  // RE/RaytracingEngine.cpp:176-214
    for (int op = 0; op < RTE_TM_COUNT; ++op) {
        if (rgb8 && rgb8[op])
            for (int64_t i = 0; i < n_pixels; ++i) {
                rgb8[op][3 * i] = all[op][(size_t)i].r; rgb8[op][3 * i + 1] = all[op][(size_t)i].g; rgb8[op][3 * i + 2] = all[op][(size_t)i].b;
            }
    }
    if (hdr_out) {
        for (int64_t i = 0; i < n_pixels; ++i) {
            const Vec3& c = px[(size_t)i];
            const Vec3 r[RTE_TM_COUNT] = {simple(c), reinhardSimple(c), reinhardExtended(c, 5.0), reinhardExtendedLuminance(c, 5.0),
                                          reinhardJodie(c), uncharted2(c), aces_approx(c)};
            for (int op = 0; op < RTE_TM_COUNT; ++op)
                if (hdr_out[op]) { hdr_out[op][3 * i] = r[op].x; hdr_out[op][3 * i + 1] = r[op].y; hdr_out[op][3 * i + 2] = r[op].z; }
        }
    }
    return RTE_OK;
}

int ref_tonemap_one(int op, double max_white, double jodie_a, const double* hdr, int64_t n_pixels, uint8_t* rgb8) {
    for (int64_t i = 0; i < n_pixels; ++i) {
        const Vec3 c = v3(hdr + 3 * i);
        Vec3 m(0, 0, 0);
        switch (op) {
            case RTE_TM_SIMPLE: m = simple(c); break;
            case RTE_TM_REINHARD_SIMPLE: m = reinhardSimple(c); break;
            case RTE_TM_REINHARD_EXTENDED: m = reinhardExtended(c, max_white); break;
            case RTE_TM_REINHARD_EXTENDED_LUMINANCE: m = reinhardExtendedLuminance(c, max_white); break;
            case RTE_TM_REINHARD_JODIE: m = reinhardJodie(c, jodie_a); break;
            case RTE_TM_UNCHARTED2: m = uncharted2(c); break;
            case RTE_TM_ACES: m = aces_approx(c); break;
            default: return RTE_ERR_INVALID;
        }
        const Color q = toColor(m);  // RE/RaytracingEngine.cpp:113-121
        rgb8[3 * i] = q.r; rgb8[3 * i + 1] = q.g; rgb8[3 * i + 2] = q.b;
    }
    return RTE_OK;
}

int ref_write_ppm(const char* filename, const uint8_t* rgb8, uint64_t width, uint64_t height) {
    std::vector<Color> px((size_t)(width * height));
    for (size_t i = 0; i < px.size(); ++i) px[i] = Color(rgb8[3 * i], rgb8[3 * i + 1], rgb8[3 * i + 2]);
    try {
        std::cout.setstate(std::ios_base::failbit);  // silence "Image written to"
        writePPM(filename, px, (size_t)width, (size_t)height);  // RE/Image.cpp:11
        std::cout.clear();
    } catch (const std::exception&) {
        std::cout.clear();
        return RTE_ERR_IO;
    }
    return RTE_OK;
}

// LoadObject (RE/RaytracingEngine.cpp:15-65) through the vendored tinyobjloader, to validate the
// product's own OBJ reader.  Two-call pattern: NULL outputs => counts only.
int ref_load_obj(const char* path, int64_t* n_vertices, int64_t* n_indices, double* vertices, int32_t* indices) {
    try {
        Model m = LoadObject(path);
        if (n_vertices) *n_vertices = (int64_t)m.vertexPositions.size();
        if (n_indices) *n_indices = (int64_t)m.vertices.size();
        if (vertices)
            for (size_t i = 0; i < m.vertexPositions.size(); ++i) {
                vertices[3 * i] = m.vertexPositions[i].x; vertices[3 * i + 1] = m.vertexPositions[i].y; vertices[3 * i + 2] = m.vertexPositions[i].z;
            }
        if (indices)
            for (size_t i = 0; i < m.vertices.size(); ++i) indices[i] = m.vertices[i];
    } catch (const std::exception&) {
        return RTE_ERR_IO;
    }
    return RTE_OK;
}

// sizeof() KATs of SURVEY.md §8(c)
void ref_struct_sizes(int32_t* out10) {
    const size_t v[10] = {sizeof(Vec3), sizeof(Material), sizeof(Transform), sizeof(HitInfo), sizeof(Sphere),
                          sizeof(Plane), sizeof(Triangle), sizeof(Light), sizeof(Camera), sizeof(Color)};
    for (int i = 0; i < 10; ++i) out10[i] = (int32_t)v[i];
}

}  // extern "C"
