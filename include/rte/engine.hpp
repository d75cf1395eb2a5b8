// rte/engine.hpp — C++20 host mirror of the reference's scene-setup API over the C ABI of librte_b200.so.
//
// Same type names, constructor argument orders and defaults as the reference ("RE/" = RaytracingEngine/ upstream):
//   Vec3, Color, Camera (RE/Math.h:9-122), Transform, Material, Sphere, Plane, Triangle, Model (RE/Shape.h:7-307),
//   Light (RE/Light.h:6-15), Scene::Add*/RenderImage (RE/Scene.h:200-212,311), tonemap/tonemapAll
//   (RE/RaytracingEngine.cpp:165-214), writePPM (RE/Image.h:7).
// These classes only DESCRIBE a scene; every intersection, shading and tonemap computation happens on the GPU behind
// include/rte.h.  Errors follow the reference's convention: std::runtime_error (RE/Image.cpp:15,27).
// A reference program switches by replacing its four includes with this header and linking -lrte_b200
// (INTEGRATION.md); header-only, no CUDA headers needed.
#pragma once
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../rte.h"

namespace rte_host {

struct Vec3 {  // RE/Math.h:9-12
    double x, y, z;
    explicit Vec3(double x_ = 0, double y_ = 0, double z_ = 0) : x(x_), y(y_), z(z_) {}
};
struct Color { std::uint8_t r = 0, g = 0, b = 0; };  // RE/Math.h:73-76

struct Camera {  // RE/Math.h:85-97 (same member names; ctor: position, focal, width, height, near, far)
    Vec3 position, forward{0, 0, 1};
    std::size_t width, height;
    double focal, farPlaneDistance, nearPlaneDistance;
    std::size_t antiAliasingAmount = 32;  // RE/Math.h:94; samples 1.. are jittered from the counter-based stream of rte.h
    std::int32_t jitterSeed = 0;          // (upstream: std::random_device)
    explicit Camera(const Vec3& pos, double f = 1.0, std::size_t w = 800, std::size_t h = 600, double nearP = 1.0, double farP = 1000.0)
        : position(pos), width(w), height(h), focal(f), farPlaneDistance(farP), nearPlaneDistance(nearP) {}
};

struct Transform { Vec3 position, rotation, scale; };  // RE/Shape.h:7-11 (only position is ever applied upstream)
struct Material {                                      // RE/Shape.h:13-19
    Vec3 color;
    double shininess = 128.0, specular = 0.0, transparency = 0.0, refractiveIndex = 1.0;
};
struct Sphere {  // Sphere(r = 1, pos = 0, mat), RE/Shape.h:65
    double radius; Transform transform; Material material;
    explicit Sphere(double r = 1.0, const Vec3& pos = Vec3(), const Material& m = Material()) : radius(r), transform{pos, Vec3(), Vec3()}, material(m) {}
};
struct Plane {  // Plane(pos = (0,1,0), norm = (0,1,0), mat), RE/Shape.h:141 — the library normalises the normal like the ctor does
    Transform transform; Vec3 normal; Material material;
    explicit Plane(const Vec3& pos = Vec3(0, 1, 0), const Vec3& n = Vec3(0, 1, 0), const Material& m = Material()) : transform{pos, Vec3(), Vec3()}, normal(n), material(m) {}
};
struct Triangle {  // Triangle(v0, v1, v2, mat, transform), RE/Shape.h:194
    Vec3 v0, v1, v2; Material material; Transform transform;
    Triangle(const Vec3& a, const Vec3& b, const Vec3& c, const Material& m = Material(), const Transform& t = Transform()) : v0(a), v1(b), v2(c), material(m), transform(t) {}
};
struct Model {  // Model(indices, transform, material, vertexPositions), RE/Shape.h:256
    std::vector<int> vertices; Transform transform; Material material; std::vector<Vec3> vertexPositions;
    Model(std::vector<int> idx, const Transform& t, const Material& m, std::vector<Vec3> pos)
        : vertices(std::move(idx)), transform(t), material(m), vertexPositions(std::move(pos)) {}
};
struct Light {  // RE/Light.h:6-15
    Vec3 position, color{1, 1, 1}; double intensity = 1.0;
    Light() = default;
    Light(const Vec3& p, const Vec3& c, double i) : position(p), color(c), intensity(i) {}
};

inline void check(int rc, const rte_ctx* ctx = nullptr) {
    if (rc != RTE_OK) throw std::runtime_error(rte_last_error(ctx));
}

class Scene {  // RE/Scene.h:14-24,200-212
public:
    explicit Scene(const Camera& cam, int device = 0) : camera(cam) { check(rte_create(device, &ctx_)); }
    ~Scene() { rte_destroy(ctx_); }
    Scene(const Scene&) = delete;
    Scene& operator=(const Scene&) = delete;

    void AddSphere(Sphere& s) { spheres_.push_back({{s.transform.position.x, s.transform.position.y, s.transform.position.z}, s.radius, mat(s.material), 0}); dirty_ = true; }
    void AddPlane(Plane& p) {
        planes_.push_back({{p.transform.position.x, p.transform.position.y, p.transform.position.z}, {p.normal.x, p.normal.y, p.normal.z}, mat(p.material), 0});
        dirty_ = true;
    }
    void AddTriangle(Triangle& t) {
        triangles_.push_back({{t.v0.x, t.v0.y, t.v0.z}, {t.v1.x, t.v1.y, t.v1.z}, {t.v2.x, t.v2.y, t.v2.z},
                              {t.transform.position.x, t.transform.position.y, t.transform.position.z}, mat(t.material), 0});
        dirty_ = true;
    }
    void AddModel(Model& m) {
        rte_model d{};
        d.first_vertex = (std::int64_t)(verts_.size() / 3); d.vertex_count = (std::int64_t)m.vertexPositions.size();
        d.first_index = (std::int64_t)indices_.size(); d.index_count = (std::int64_t)m.vertices.size();
        d.translate[0] = m.transform.position.x; d.translate[1] = m.transform.position.y; d.translate[2] = m.transform.position.z;
        d.material = mat(m.material);
        for (const Vec3& v : m.vertexPositions) { verts_.push_back(v.x); verts_.push_back(v.y); verts_.push_back(v.z); }
        for (int i : m.vertices) indices_.push_back(i);
        models_.push_back(d);
        dirty_ = true;
    }
    void AddLight(Light& l) { lights_.push_back({{l.position.x, l.position.y, l.position.z}, {l.color.x, l.color.y, l.color.z}, l.intensity}); dirty_ = true; }
    std::size_t GetPixelIndex(std::size_t x, std::size_t y) const { return y * camera.width + x; }  // RE/Scene.h:214-216

    // Scene::RenderImage(), RE/Scene.h:311: linear HDR, row-major, row 0 = top — rendered on the GPU.
    std::vector<Vec3> RenderImage() {
        sync();
        const std::size_t n = camera.width * camera.height;
        std::vector<double> h(3 * n);
        rte_frame f{};
        f.hdr64 = h.data();
        check(rte_render_host(ctx_, nullptr, &f), ctx_);
        std::vector<Vec3> out(n);
        for (std::size_t i = 0; i < n; ++i) out[i] = Vec3(h[3 * i], h[3 * i + 1], h[3 * i + 2]);
        return out;
    }
    // RenderImage + tonemap(op) + toColor fused in one pass (RE/RaytracingEngine.cpp:165-174 for op = RTE_TM_ACES).
    std::vector<Color> RenderTonemapped(int op = RTE_TM_ACES) {
        sync();
        std::vector<Color> out(camera.width * camera.height);
        rte_render_opts o{};
        o.tonemap_op = op;
        rte_frame f{};
        f.rgb8 = reinterpret_cast<std::uint8_t*>(out.data());
        check(rte_render_host(ctx_, &o, &f), ctx_);
        return out;
    }
    rte_ctx* context() { return ctx_; }
    Camera camera;

private:
    int mat(const Material& m) {
        materials_.push_back({{m.color.x, m.color.y, m.color.z}, m.shininess, m.specular, m.transparency, m.refractiveIndex});
        return (int)materials_.size() - 1;
    }
    void sync() {
        if (dirty_) {
            rte_scene_desc d{};
            d.materials = materials_.data(); d.n_materials = (std::int64_t)materials_.size();
            d.spheres = spheres_.data(); d.n_spheres = (std::int64_t)spheres_.size();
            d.planes = planes_.data(); d.n_planes = (std::int64_t)planes_.size();
            d.triangles = triangles_.data(); d.n_triangles = (std::int64_t)triangles_.size();
            d.models = models_.data(); d.n_models = (std::int64_t)models_.size();
            d.model_vertices = verts_.data(); d.n_model_vertices = (std::int64_t)(verts_.size() / 3);
            d.model_indices = indices_.data(); d.n_model_indices = (std::int64_t)indices_.size();
            d.lights = lights_.data(); d.n_lights = (std::int64_t)lights_.size();
            check(rte_upload_scene(ctx_, &d), ctx_);
            dirty_ = false;
        }
        rte_camera c{};
        c.position[0] = camera.position.x; c.position[1] = camera.position.y; c.position[2] = camera.position.z;
        c.focal = camera.focal; c.width = camera.width; c.height = camera.height;
        c.near_plane = camera.nearPlaneDistance; c.far_plane = camera.farPlaneDistance;
        c.anti_aliasing = (std::int32_t)camera.antiAliasingAmount;
        c.jitter_seed = camera.jitterSeed;
        check(rte_set_camera(ctx_, &c), ctx_);
    }
    rte_ctx* ctx_ = nullptr;
    bool dirty_ = true;
    std::vector<rte_material> materials_;
    std::vector<rte_sphere> spheres_;
    std::vector<rte_plane> planes_;
    std::vector<rte_triangle> triangles_;
    std::vector<rte_model> models_;
    std::vector<double> verts_;
    std::vector<std::int32_t> indices_;
    std::vector<rte_light> lights_;
};

static_assert(sizeof(Color) == 3, "Color must stay 3 bytes: RGB8 buffers are reinterpreted (RE/Math.h:73-76)");

// tonemapAll(pixels), RE/RaytracingEngine.cpp:176-214: the seven operators in the reference's order
// (simple, reinhard_simple, reinhard_extended, reinhard_extended_luminance, reinhard_jodie, uncharted2, aces).
inline std::vector<std::vector<Color>> tonemapAll(Scene& scene, const std::vector<Vec3>& pixels, double max_white = 5.0, double jodie_a = 0.18) {
    const std::size_t n = pixels.size();
    std::vector<float> h(4 * n);
    for (std::size_t i = 0; i < n; ++i) { h[4 * i] = (float)pixels[i].x; h[4 * i + 1] = (float)pixels[i].y; h[4 * i + 2] = (float)pixels[i].z; h[4 * i + 3] = 0.f; }
    std::vector<std::vector<Color>> out(RTE_TM_COUNT, std::vector<Color>(n));
    std::uint8_t* ptr[RTE_TM_COUNT];
    for (int op = 0; op < RTE_TM_COUNT; ++op) ptr[op] = reinterpret_cast<std::uint8_t*>(out[op].data());
    check(rte_tonemap_host(scene.context(), h.data(), (std::int64_t)n, (1u << RTE_TM_COUNT) - 1u, max_white, jodie_a, ptr), scene.context());
    return out;
}
// tonemap(pixels) = ACES, RE/RaytracingEngine.cpp:165-174
inline std::vector<Color> tonemap(Scene& scene, const std::vector<Vec3>& pixels) {
    const std::size_t n = pixels.size();
    std::vector<float> h(4 * n);
    for (std::size_t i = 0; i < n; ++i) { h[4 * i] = (float)pixels[i].x; h[4 * i + 1] = (float)pixels[i].y; h[4 * i + 2] = (float)pixels[i].z; h[4 * i + 3] = 0.f; }
    std::vector<Color> out(n);
    std::uint8_t* ptr[RTE_TM_COUNT] = {};
    ptr[RTE_TM_ACES] = reinterpret_cast<std::uint8_t*>(out.data());
    check(rte_tonemap_host(scene.context(), h.data(), (std::int64_t)n, 1u << RTE_TM_ACES, 0.0, 0.0, ptr), scene.context());
    return out;
}
// writePPM(filename, pixels, width, height), RE/Image.h:7
inline void writePPM(const std::string& filename, const std::vector<Color>& pixels, std::size_t width, std::size_t height) {
    check(rte_write_ppm(filename.c_str(), reinterpret_cast<const std::uint8_t*>(pixels.data()), width, height));
}

// the .png main() gets from ffmpeg + del (RE/RaytracingEngine.cpp:317-327), written directly
inline void writePNG(const std::string& filename, const std::vector<Color>& pixels, std::size_t width, std::size_t height) {
    check(rte_write_png(filename.c_str(), reinterpret_cast<const std::uint8_t*>(pixels.data()), width, height));
}

}  // namespace rte_host
