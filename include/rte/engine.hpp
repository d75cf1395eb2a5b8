// rte/engine.hpp — C++20 host mirror of the reference's public API over the C ABI of librte_b200.so.
//
// Same type names, member names, constructor argument orders and defaults as the reference ("RE/" = RaytracingEngine/ upstream):
//   Vec3, Color, Rayon, Camera (RE/Math.h:9-122), Transform, Material, HitType, HitInfo, Sphere, Plane, Triangle, Model
//   (RE/Shape.h:7-307), Light (RE/Light.h:6-59), Scene (RE/Scene.h:200-328), LoadObject, toColor, the seven operator functions,
//   tonemap, tonemapAll (RE/RaytracingEngine.cpp:15-214), writePPM (RE/Image.h:7).
// A reference program switches by replacing its includes of Math.h / Light.h / Shape.h / Scene.h / Image.h / tiny_obj_loader.h
// with this header and linking -lrte_b200; tests/test_gpu_cpp_mirror.py compiles the body of the reference's own
// main() that way.  Header-only, no CUDA headers needed.
//
// What runs where: these classes DESCRIBE a scene (value semantics like upstream) and do the vector arithmetic a caller needs
// to set one up; every ray query, the render, the operators and tonemap()/tonemapAll() run on the GPU behind include/rte.h.
// There is no CPU render path.  Errors follow the reference's convention: std::runtime_error (RE/Image.cpp:15,27).
// The names live in namespace rte_host and are imported into the global namespace like upstream's (define
// RTE_ENGINE_NO_GLOBAL_NAMES before including to keep them qualified).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <optional>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "../rte.h"

#ifndef RTE_ENGINE_DEFAULT_AA
#define RTE_ENGINE_DEFAULT_AA 32  // Camera::antiAliasingAmount (a test harness may compile an unmodified main() with 1: deterministic render)
#endif

namespace rte_host {

// ------------------------------------------------------------------------------------------ RE/Math.h
struct Vec3 {  // RE/Math.h:9-71
    double x, y, z;
    explicit Vec3(const double x = 0.0, const double y = 0.0, const double z = 0.0) : x(x), y(y), z(z) {}

    Vec3 operator+(const Vec3& o) const noexcept { return Vec3{x + o.x, y + o.y, z + o.z}; }
    Vec3 operator+(const double s) const noexcept { return Vec3{x + s, y + s, z + s}; }
    Vec3 operator-(const Vec3& o) const noexcept { return Vec3{x - o.x, y - o.y, z - o.z}; }
    Vec3 operator-(const double s) const noexcept { return Vec3{x - s, y - s, z - s}; }
    Vec3 operator*(const double s) const noexcept { return Vec3{x * s, y * s, z * s}; }
    Vec3 operator*(const Vec3& o) const noexcept { return Vec3{x * o.x, y * o.y, z * o.z}; }
    Vec3 operator/(const Vec3& o) const noexcept { return Vec3{x / o.x, y / o.y, z / o.z}; }
    Vec3 operator/(const double s) const noexcept { return Vec3{x / s, y / s, z / s}; }
    Vec3& operator+=(const Vec3& o) noexcept { x += o.x; y += o.y; z += o.z; return *this; }
    Vec3 operator-() const noexcept { return Vec3{-x, -y, -z}; }
    Vec3& operator/=(const double s) noexcept { x /= s; y /= s; z /= s; return *this; }
    Vec3& operator*=(const double s) noexcept { x *= s; y *= s; z *= s; return *this; }

    double dot(const Vec3& o) const noexcept { return x * o.x + y * o.y + z * o.z; }
    Vec3 cross(const Vec3& o) const noexcept { return Vec3{y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x}; }
    double length() const noexcept { return std::sqrt(dot(*this)); }
    Vec3 normalize() const noexcept {  // (0,0,0) below 1e-12, RE/Math.h:31-37
        const double len = length();
        if (len <= 1e-12) return Vec3{0, 0, 0};
        return (*this) / len;
    }
    Vec3 reflect(const Vec3& normal) const noexcept { return (*this) - normal * 2.0 * this->dot(normal); }
    Vec3 refract(const Vec3& normal, double eta) const {  // RE/Math.h:43-52
        const Vec3 I = this->normalize(), N = normal.normalize();
        const double cosi = std::clamp(I.dot(N), -1.0, 1.0);
        const double k = 1.0 - eta * eta * (1.0 - cosi * cosi);
        if (k < 0.0) return Vec3{0, 0, 0};
        return I * eta - N * (eta * cosi + std::sqrt(k));
    }
    double unsafeIndex(const int index) const {
        switch (index) {
            case 0: return x;
            case 1: return y;
            case 2: return z;
            default: throw std::out_of_range("Vec3 index out of range");
        }
    }
    Vec3& lerp(const Vec3& target, const double t) noexcept {
        x = x + (target.x - x) * t; y = y + (target.y - y) * t; z = z + (target.z - z) * t;
        return *this;
    }
    friend Vec3 operator*(double s, const Vec3& v) noexcept { return v * s; }
};

struct Color {  // RE/Math.h:73-76
    std::uint8_t r, g, b;
    Color(const std::uint8_t r = 0, const std::uint8_t g = 0, const std::uint8_t b = 0) : r(r), g(g), b(b) {}
};
static_assert(sizeof(Color) == 3, "Color must stay 3 bytes: RGB8 buffers are reinterpreted (RE/Math.h:73-76)");

struct Rayon {  // RE/Math.h:78-83
    Vec3 origin, direction;
    Rayon(const Vec3& o, const Vec3& d) : origin(o), direction(d) {}
    Vec3 pointAtDistance(const double t) const noexcept { return origin + direction * t; }
};

struct Camera {  // RE/Math.h:85-122 (same members, same order)
    Vec3 position, forward;
    std::size_t width, height;
    double focal, farPlaneDistance, nearPlaneDistance;
    int antiAliasingAmount = RTE_ENGINE_DEFAULT_AA;  // 32, RE/Math.h:94; on the GPU samples 1.. are jittered from the counter-based stream of rte.h
    std::int32_t jitterSeed = 0;  // (upstream seeds its jitter from std::random_device: not reproducible)

    Camera(const Vec3& position, double focal = 1.0, std::size_t width = 800, std::size_t height = 600, double nearPlaneDistance = 1.0,
           double farPlaneDistance = 1000.0)
        : position(position), forward{0, 0, 1}, width(width), height(height), focal(focal), farPlaneDistance(farPlaneDistance),
          nearPlaneDistance(nearPlaneDistance) {}

    // One ray on the host, for single-ray queries (Scene::CalculatePixelDepth); the render generates its rays on the device.
    Rayon getRay(const std::size_t pixelX, const std::size_t pixelY, const bool aa) const {  // RE/Math.h:99-121
        auto sx = static_cast<double>(pixelX) - static_cast<double>(width) / 2.0;
        auto sy = static_cast<double>(height) / 2.0 - static_cast<double>(pixelY);
        if (aa) {
            const auto invAA = 1.0 / static_cast<double>(aa);
            thread_local static std::mt19937 gen((std::random_device())());
            thread_local static std::uniform_real_distribution<double> dist(0.0, 1.0);
            sx += dist(gen) * invAA;
            sy += dist(gen) * invAA;
        }
        const auto screenPoint = Vec3(sx, sy, position.z + focal);
        return Rayon{position, (screenPoint - position).normalize()};
    }
};

// ------------------------------------------------------------------------------------------ RE/Shape.h
struct Transform { Vec3 position, rotation, scale; };  // RE/Shape.h:7-11 (only position is ever applied upstream)
struct Material {                                      // RE/Shape.h:13-19
    Vec3 color;
    double shininess = 128.0, specular = 0.0, transparency = 0.0, refractiveIndex = 1.0;
};
enum class HitType : unsigned char { NONE, SPHERE, PLANE, TRIANGLE };  // RE/Shape.h:21-26

struct HitInfo {  // RE/Shape.h:28-57
    HitType type;
    double distance;
    std::size_t index;
    Material material;
    Vec3 normal, hitPoint;
    bool isCloserThan(const HitInfo& other) const { return distance < other.distance; }
    double normalizedDistance(const Camera& camera) const {
        return (distance - camera.nearPlaneDistance) / (camera.farPlaneDistance - camera.nearPlaneDistance);
    }
    static HitInfo getClosestIntersection(const std::vector<HitInfo>& intersections) {
        HitInfo closest = intersections[0];
        for (const auto& info : intersections)
            if (info.isCloserThan(closest)) closest = info;
        return closest;
    }
};

class Sphere {  // Sphere(r = 1, pos = 0, mat), RE/Shape.h:59-133
    double radius;
    Transform transform;
    Material material;
public:
    explicit Sphere(const double r = 1.0, const Vec3& pos = Vec3(0, 0, 0), const Material& mat = Material()) : radius(r), material(mat) {
        transform.position = pos; transform.rotation = Vec3(0, 0, 0); transform.scale = Vec3(1, 1, 1);
    }
    double getRadius() const { return radius; }
    void setRadius(double r) { radius = r; }
    Material getMaterial() const { return material; }
    Transform getTransform() const { return transform; }
    static Sphere getHitObject(const HitInfo& hit, const std::vector<Sphere>& spheres) { return spheres[hit.index]; }
};

class Plane {  // Plane(pos = (0,1,0), norm = (0,1,0), mat), RE/Shape.h:135-186
    Vec3 normal, given;  // normalised like the ctor does (RE/Shape.h:142) / as given: the library normalises the given one itself, once
    Transform transform;
    Material material;
public:
    Plane(const Vec3& pos = Vec3(0, 1, 0), const Vec3& norm = Vec3(0, 1, 0), const Material& material = Material())
        : normal(norm.normalize()), given(norm), material(material) {
        transform.position = pos; transform.rotation = Vec3(0, 0, 0); transform.scale = Vec3(1, 1, 1);
    }
    Vec3 GetNormalAt() const { return normal; }
    Vec3 GetNormal() const { return normal; }
    void SetNormal(const Vec3& norm) { normal = norm.normalize(); given = norm; }
    Material GetMaterial() const { return material; }
    Transform GetTransform() const { return transform; }
    const Vec3& givenNormal() const { return given; }
};

class Triangle {  // Triangle(v0, v1, v2, mat, transform), RE/Shape.h:188-246
    Vec3 v0, v1, v2;
    Transform transform;
    Material material;
public:
    Triangle(const Vec3& vertex0, const Vec3& vertex1, const Vec3& vertex2, const Material& mat = Material(), const Transform& t = Transform())
        : v0(vertex0), v1(vertex1), v2(vertex2), transform(t), material(mat) {}
    Vec3 tv0() const { return v0 + transform.position; }
    Vec3 tv1() const { return v1 + transform.position; }
    Vec3 tv2() const { return v2 + transform.position; }
    std::optional<Vec3> GetNormalAt() const { return (v1 - v0).cross(v2 - v0).normalize(); }  // from the untranslated vertices
    Material GetMaterial() const { return material; }
    const Vec3& vertex(int i) const { return i == 0 ? v0 : i == 1 ? v1 : v2; }
    const Transform& getTransform() const { return transform; }
};

class Model {  // Model(indices, transform, material, vertexPositions), RE/Shape.h:248-307
    std::vector<int> vertices;
    std::vector<Vec3> vertexPositions;
    Transform transform;
    Material material;
public:
    Model(const std::vector<int>& vertices, const Transform& transform = Transform(), const Material& material = Material(),
          const std::vector<Vec3>& vertexPositions = std::vector<Vec3>())
        : vertices(vertices), vertexPositions(vertexPositions), transform(transform), material(material) {}
    std::vector<Triangle> GetTrianglesFromModel(const Material& overrideMaterial) const {
        std::vector<Triangle> triangles;
        for (std::size_t i = 0; i + 2 < vertices.size(); i += 3)
            triangles.emplace_back(Triangle(vertexPositions[vertices[i]], vertexPositions[vertices[i + 1]], vertexPositions[vertices[i + 2]], overrideMaterial, transform));
        return triangles;
    }
    Transform GetTransform() const { return transform; }
    Material GetMaterial() const { return material; }
    void SetTransform(const Transform& t) { transform = t; }
    void SetMaterial(const Material& m) { material = m; }
    const std::vector<int>& indices() const { return vertices; }
    const std::vector<Vec3>& positions() const { return vertexPositions; }
};

// ------------------------------------------------------------------------------------------ RE/Light.h
struct Light {  // RE/Light.h:6-59
    Vec3 position, color;
    double intensity;
    Light(const Vec3& position, const Vec3& color, const double& intensity) : position(position), color(color), intensity(intensity) {}
    Light() : position(Vec3(0, 0, 0)), color(Vec3(1, 1, 1)), intensity(1.0) {}
    Vec3 toLightDirection(const Vec3& point) const { return position - point; }
    double distanceTo(const Vec3& point) const { return toLightDirection(point).length(); }
    Vec3 dirTo(const Vec3& point) const {
        const Vec3 v = position - point;
        const double n2 = v.x * v.x + v.y * v.y + v.z * v.z;
        if (n2 <= 1e-12 * 1e-12) return Vec3{0, 0, 0};
        return (1.0 / std::sqrt(n2)) * v;
    }
    Rayon shadowRayFrom(const Vec3& hitPoint, const double bias) const { const Vec3 L = dirTo(hitPoint); return Rayon{hitPoint + L * bias, L}; }
    Vec3 emitted() const { return color * intensity; }
    Vec3 contributionFrom(const double dist, const double NdotL) const {
        if (dist <= 1e-12 || NdotL <= 0.0) return Vec3{0, 0, 0};
        return emitted() * (1.0 / (dist * dist) * NdotL);
    }
};

inline void check(int rc, const rte_ctx* ctx = nullptr) {
    if (rc != RTE_OK) throw std::runtime_error(rte_last_error(ctx));
}

// ------------------------------------------------------------------------------------------ RE/Scene.h
class Scene {  // RE/Scene.h:14-24,200-328: value semantics (Add* copy what they are given; a copied Scene gets its own GPU context)
public:
    explicit Scene(const Camera& camera) : camera(camera) {}
    Scene(const Scene& o) : camera(o.camera), materials_(o.materials_), spheres_(o.spheres_), planes_(o.planes_), triangles_(o.triangles_),
                            models_(o.models_), verts_(o.verts_), indices_(o.indices_), lights_(o.lights_) {}
    Scene& operator=(const Scene& o) {
        if (this != &o) { release(); camera = o.camera; materials_ = o.materials_; spheres_ = o.spheres_; planes_ = o.planes_; triangles_ = o.triangles_;
                          models_ = o.models_; verts_ = o.verts_; indices_ = o.indices_; lights_ = o.lights_; dirty_ = true; }
        return *this;
    }
    ~Scene() { release(); }

    void AddSphere(Sphere& s) {
        const Vec3 c = s.getTransform().position;
        spheres_.push_back({{c.x, c.y, c.z}, s.getRadius(), mat(s.getMaterial()), 0}); dirty_ = true;
    }
    void AddPlane(Plane& p) {
        const Vec3 c = p.GetTransform().position, n = p.givenNormal();
        planes_.push_back({{c.x, c.y, c.z}, {n.x, n.y, n.z}, mat(p.GetMaterial()), 0}); dirty_ = true;
    }
    void AddLight(Light& l) { lights_.push_back({{l.position.x, l.position.y, l.position.z}, {l.color.x, l.color.y, l.color.z}, l.intensity}); dirty_ = true; }
    void AddTriangle(Triangle& t) {
        const Vec3 &a = t.vertex(0), &b = t.vertex(1), &c = t.vertex(2), &p = t.getTransform().position;
        triangles_.push_back({{a.x, a.y, a.z}, {b.x, b.y, b.z}, {c.x, c.y, c.z}, {p.x, p.y, p.z}, mat(t.GetMaterial()), 0}); dirty_ = true;
    }
    void AddModel(Model& m) {
        rte_model d{};
        d.first_vertex = (std::int64_t)(verts_.size() / 3); d.vertex_count = (std::int64_t)m.positions().size();
        d.first_index = (std::int64_t)indices_.size(); d.index_count = (std::int64_t)m.indices().size();
        const Vec3 p = m.GetTransform().position;
        d.translate[0] = p.x; d.translate[1] = p.y; d.translate[2] = p.z;
        d.material = mat(m.GetMaterial());
        for (const Vec3& v : m.positions()) { verts_.push_back(v.x); verts_.push_back(v.y); verts_.push_back(v.z); }
        for (int i : m.indices()) indices_.push_back(i);
        models_.push_back(d);
        dirty_ = true;
    }
    std::size_t GetPixelIndex(const std::size_t x, const std::size_t y) const { return y * camera.width + x; }  // RE/Scene.h:214-216

    // Scene::IntersectClosest, RE/Scene.h:218-257 (one ray through rte_intersect_closest)
    std::optional<HitInfo> IntersectClosest(const Rayon& ray) const {
        sync();
        const double o[3] = {ray.origin.x, ray.origin.y, ray.origin.z}, d[3] = {ray.direction.x, ray.direction.y, ray.direction.z};
        rte_hit h{};
        check(rte_intersect_closest(ctx_, 1, o, d, &h), ctx_);
        if (h.prim_id == 0xFFFFFFFFu) return std::nullopt;
        HitInfo r;
        r.type = static_cast<HitType>(h.prim_id >> 30);
        r.distance = h.distance;
        r.index = h.prim_id & 0x3FFFFFFFu;
        r.material = Material();
        if (h.material >= 0 && (std::size_t)h.material < materials_.size()) {
            const rte_material& m = materials_[(std::size_t)h.material];
            r.material = Material{Vec3(m.color[0], m.color[1], m.color[2]), m.shininess, m.specular, m.transparency, m.refractive_index};
        }
        r.normal = Vec3(h.normal[0], h.normal[1], h.normal[2]);
        r.hitPoint = Vec3(h.point[0], h.point[1], h.point[2]);
        return r;
    }
    // Scene::IntersectAnyBefore, RE/Scene.h:259-276
    bool IntersectAnyBefore(const Rayon& ray, const double maxDist) const {
        sync();
        const double o[3] = {ray.origin.x, ray.origin.y, ray.origin.z}, d[3] = {ray.direction.x, ray.direction.y, ray.direction.z};
        std::uint8_t out = 0;
        check(rte_intersect_any_before(ctx_, 1, o, d, &maxDist, &out), ctx_);
        return out != 0;
    }
    // Scene::CalculatePixelDepth, RE/Scene.h:278-281
    std::optional<HitInfo> CalculatePixelDepth(const std::size_t x, const std::size_t y, const bool aa) const { return IntersectClosest(camera.getRay(x, y, aa)); }
    // Scene::GeneratePixelAt, RE/Scene.h:283-304: one pixel of the render
    Vec3 GeneratePixelAt(const int x, const int y) const {
        sync();
        std::vector<double> h(3 * camera.width);
        rte_render_opts o{};
        o.row0 = y; o.row1 = y + 1; o.tonemap_op = RTE_TM_NONE;
        rte_frame f{};
        f.hdr64 = h.data();
        check(rte_render_host(ctx_, &o, &f), ctx_);
        return Vec3(h[3 * (std::size_t)x], h[3 * (std::size_t)x + 1], h[3 * (std::size_t)x + 2]);
    }
    // Scene::RenderImage(), RE/Scene.h:311: linear HDR (FP64, the reference's own values), row-major, row 0 = top — rendered on the GPU.
    std::vector<Vec3> RenderImage() const {
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
    // RenderImage + tonemap(op) + toColor in one call (RE/RaytracingEngine.cpp:165-174 for op = RTE_TM_ACES): the production path
    // (FP32 on diffuse sphere / plane scenes: RGB8 within +-1 LSB; flags = RTE_FLAG_EXACT_FP64 for the reference's own bytes).
    std::vector<Color> RenderTonemapped(int op = RTE_TM_ACES, int flags = 0) const {
        sync();
        std::vector<Color> out(camera.width * camera.height);
        rte_render_opts o{};
        o.tonemap_op = op; o.flags = flags;
        rte_frame f{};
        f.rgb8 = reinterpret_cast<std::uint8_t*>(out.data());
        check(rte_render_host(ctx_, &o, &f), ctx_);
        return out;
    }
    rte_ctx* context() const { sync(); return ctx_; }
    int device = 0;  // GPU of this scene's context (set before the first render / query)
    Camera camera;   // (private upstream; copied by the constructor like there)

private:
    int mat(const Material& m) {
        materials_.push_back({{m.color.x, m.color.y, m.color.z}, m.shininess, m.specular, m.transparency, m.refractiveIndex});
        return (int)materials_.size() - 1;
    }
    void release() { if (ctx_) { rte_destroy(ctx_); ctx_ = nullptr; } }
    void sync() const {
        if (!ctx_) { check(rte_create(device, &ctx_)); dirty_ = true; }
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
        if (std::memcmp(&c, &cam_sent_, sizeof(c)) != 0) { check(rte_set_camera(ctx_, &c), ctx_); cam_sent_ = c; }
    }
    mutable rte_ctx* ctx_ = nullptr;
    mutable bool dirty_ = true;
    mutable rte_camera cam_sent_{};
    std::vector<rte_material> materials_;
    std::vector<rte_sphere> spheres_;
    std::vector<rte_plane> planes_;
    std::vector<rte_triangle> triangles_;
    std::vector<rte_model> models_;
    std::vector<double> verts_;
    std::vector<std::int32_t> indices_;
    std::vector<rte_light> lights_;
};

// ------------------------------------------------------------------------------------------ RE/RaytracingEngine.cpp, RE/Image.h
// The free functions upstream take no Scene: they run on a process-wide context of GPU `default_device()` (created on first use).
inline int& default_device() { static int dev = 0; return dev; }
inline rte_ctx* default_context() {
    struct Holder { rte_ctx* ctx = nullptr; ~Holder() { if (ctx) rte_destroy(ctx); } };
    static Holder h;
    if (!h.ctx) check(rte_create(default_device(), &h.ctx));
    return h.ctx;
}

// ClampVec3 / toColor, RE/RaytracingEngine.cpp:70-76,113-121 (clamp via min(max, max(min, v)): NaN -> min; truncating quantiser)
inline Vec3 ClampVec3(const Vec3& v, double minVal = 0.0, double maxVal = 1.0) {
    return Vec3(std::min(maxVal, std::max(minVal, v.x)), std::min(maxVal, std::max(minVal, v.y)), std::min(maxVal, std::max(minVal, v.z)));
}
inline Color toColor(const Vec3& pixel) {
    const Vec3 c = ClampVec3(pixel);
    return Color(static_cast<std::uint8_t>(c.x * 255.0), static_cast<std::uint8_t>(c.y * 255.0), static_cast<std::uint8_t>(c.z * 255.0));
}
// The operator functions, RE/RaytracingEngine.cpp:89-163, evaluated by the library (FP64, the reference's operation order).
inline Vec3 tonemap_value(int op, const Vec3& c, double max_white = 5.0, double a = 0.18) {
    const double in[3] = {c.x, c.y, c.z};
    double out[3];
    check(rte_tonemap_values_host(default_context(), op, max_white, a, in, 1, out), default_context());
    return Vec3(out[0], out[1], out[2]);
}
inline Vec3 aces_approx(Vec3 v) { return tonemap_value(RTE_TM_ACES, v); }
inline Vec3 simple(const Vec3& color) { return tonemap_value(RTE_TM_SIMPLE, color); }
inline Vec3 reinhardSimple(const Vec3& color) { return tonemap_value(RTE_TM_REINHARD_SIMPLE, color); }
inline Vec3 reinhardExtended(const Vec3 color, double max_white) { return tonemap_value(RTE_TM_REINHARD_EXTENDED, color, max_white); }
inline Vec3 reinhardExtendedLuminance(const Vec3& color, double maxWhite) { return tonemap_value(RTE_TM_REINHARD_EXTENDED_LUMINANCE, color, maxWhite); }
inline Vec3 reinhardJodie(const Vec3& color, double a = 0.18) { return tonemap_value(RTE_TM_REINHARD_JODIE, color, 5.0, a); }
inline Vec3 uncharted2(const Vec3& color) { return tonemap_value(RTE_TM_UNCHARTED2, color); }
inline double luminance(const Vec3& color) { return color.dot(Vec3(0.2126, 0.7152, 0.0722)); }  // RE/RaytracingEngine.cpp:100-104
inline Vec3 change_luminance(Vec3 c_in, double l_out) { return c_in * (l_out / luminance(c_in)); }

inline std::vector<std::vector<Color>> tonemap_ops(const std::vector<Vec3>& pixels, std::uint32_t mask, double max_white, double jodie_a) {
    const std::size_t n = pixels.size();
    std::vector<double> h(3 * n);
    for (std::size_t i = 0; i < n; ++i) { h[3 * i] = pixels[i].x; h[3 * i + 1] = pixels[i].y; h[3 * i + 2] = pixels[i].z; }
    std::vector<std::vector<Color>> out(RTE_TM_COUNT);
    std::uint8_t* ptr[RTE_TM_COUNT] = {};
    for (int op = 0; op < RTE_TM_COUNT; ++op)
        if (mask & (1u << op)) { out[op].resize(n); ptr[op] = reinterpret_cast<std::uint8_t*>(out[op].data()); }
    check(rte_tonemap_exact_host(default_context(), h.data(), (std::int64_t)n, mask, max_white, jodie_a, ptr), default_context());
    return out;
}
// tonemap(pixels) = ACES, RE/RaytracingEngine.cpp:165-174 — exact FP64: the reference's own bytes
inline std::vector<Color> tonemap(const std::vector<Vec3>& pixels) { return tonemap_ops(pixels, 1u << RTE_TM_ACES, 5.0, 0.18)[RTE_TM_ACES]; }
// tonemapAll(pixels), RE/RaytracingEngine.cpp:176-214: simple, reinhard_simple, reinhard_extended(5.0), reinhard_extended_luminance(5.0),
// reinhard_jodie, uncharted2, aces — exact FP64
inline std::vector<std::vector<Color>> tonemapAll(const std::vector<Vec3> pixels) { return tonemap_ops(pixels, (1u << RTE_TM_COUNT) - 1u, 5.0, 0.18); }

// writePPM(filename, pixels, width, height), RE/Image.h:7 (prints "Image written to ..." like RE/Image.cpp:30)
inline void writePPM(const std::string& filename, const std::vector<Color>& pixels, const std::size_t width, const std::size_t height) {
    check(rte_write_ppm(filename.c_str(), reinterpret_cast<const std::uint8_t*>(pixels.data()), width, height));
    std::printf("Image written to %s\n", filename.c_str());
}
// the .png main() gets from ffmpeg + del (RE/RaytracingEngine.cpp:317-327), written directly
inline void writePNG(const std::string& filename, const std::vector<Color>& pixels, const std::size_t width, const std::size_t height) {
    check(rte_write_png(filename.c_str(), reinterpret_cast<const std::uint8_t*>(pixels.data()), width, height));
}

// ------------------------------------------------------------------------------------------ LoadObject, RE/RaytracingEngine.cpp:15-65
// The reference reads the file with the vendored tinyobjloader 1.0.6 (triangulate = true) and keeps only the vertex positions
// (float -> double) and the flat vertex-index list of all shapes in file order; normals, texture coordinates, materials and
// groups are ignored.  This reader restates exactly that part of the format:
//   v  x y z        numbers accumulate digit by digit into a double, then narrow to float (RE/tiny_obj_loader.h:474-596)
//   f  i[/j[/k]] ..  1-based, negative = relative to the vertices read SO FAR (fixIndex, :425-429); polygons become triangle
//                   fans (i0, i(k-1), i(k)) (:908-931)
namespace obj {
inline bool parse_double(const char* s, const char* s_end, double* result) {  // tryParseDouble, RE/tiny_obj_loader.h:474-587
    if (s >= s_end) return false;
    double mantissa = 0.0;
    int exponent = 0;
    char sign = '+', exp_sign = '+';
    const char* curr = s;
    int read = 0;
    auto digit = [](char c) { return static_cast<unsigned>(c - '0') < 10u; };
    if (*curr == '+' || *curr == '-') { sign = *curr; curr++; }
    else if (!digit(*curr)) return false;
    bool more = curr != s_end;
    while (more && digit(*curr)) { mantissa *= 10; mantissa += static_cast<int>(*curr - 0x30); curr++; read++; more = curr != s_end; }
    if (read == 0) return false;
    auto assemble = [&]() {
        *result = (sign == '+' ? 1 : -1) * (exponent ? std::ldexp(mantissa * std::pow(5.0, exponent), exponent) : mantissa);
        return true;
    };
    if (!more) return assemble();
    if (*curr == '.') {
        curr++; read = 1; more = curr != s_end;
        static const double pow_lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
        while (more && digit(*curr)) {
            mantissa += static_cast<int>(*curr - 0x30) * (read < 8 ? pow_lut[read] : std::pow(10.0, -read));
            read++; curr++; more = curr != s_end;
        }
    } else if (*curr != 'e' && *curr != 'E') {
        return assemble();
    }
    if (!more) return assemble();
    if (*curr == 'e' || *curr == 'E') {
        curr++; more = curr != s_end;
        if (more && (*curr == '+' || *curr == '-')) { exp_sign = *curr; curr++; }
        else if (!digit(*curr)) return false;
        read = 0; more = curr != s_end;
        while (more && digit(*curr)) { exponent *= 10; exponent += static_cast<int>(*curr - 0x30); curr++; read++; more = curr != s_end; }
        exponent *= (exp_sign == '+' ? 1 : -1);
        if (read == 0) return false;
    }
    return assemble();
}
inline float parse_real(const char** token) {  // parseReal, :589-596
    (*token) += std::strspn(*token, " \t");
    const char* end = (*token) + std::strcspn(*token, " \t\r");
    double val = 0.0;
    parse_double(*token, end, &val);
    *token = end;
    return static_cast<float>(val);
}
inline int fix_index(int idx, int n) { return idx > 0 ? idx - 1 : idx == 0 ? 0 : n + idx; }  // :425-429
}  // namespace obj

inline Model LoadObject(const std::string& modelName, const Transform& transform = Transform(), const Material& material = Material()) {
    std::ifstream ifs(modelName.c_str());
    if (!ifs) throw std::runtime_error("Failed to load/parse .obj.");  // RE/RaytracingEngine.cpp:37-39
    std::vector<Vec3> vertices;
    std::vector<int> indices;
    std::string line;
    while (std::getline(ifs, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        const char* token = line.c_str();
        token += std::strspn(token, " \t");
        if (token[0] == '\0' || token[0] == '#') continue;
        const auto space = [](char c) { return c == ' ' || c == '\t'; };
        if (token[0] == 'v' && space(token[1])) {
            token += 2;
            const float x = obj::parse_real(&token), y = obj::parse_real(&token), z = obj::parse_real(&token);
            vertices.emplace_back((double)x, (double)y, (double)z);
            continue;
        }
        if (token[0] == 'f' && space(token[1])) {
            token += 2;
            token += std::strspn(token, " \t");
            std::vector<int> face;
            while (token[0] != '\0' && token[0] != '\r' && token[0] != '\n') {
                face.push_back(obj::fix_index(std::atoi(token), (int)vertices.size()));  // the vertex index of i, i/j, i//k, i/j/k
                token += std::strcspn(token, " \t\r");
                token += std::strspn(token, " \t\r");
            }
            for (std::size_t k = 2; k < face.size(); ++k) { indices.push_back(face[0]); indices.push_back(face[k - 1]); indices.push_back(face[k]); }
            continue;
        }
    }
    return Model(indices, transform, material, vertices);
}

}  // namespace rte_host

#ifndef RTE_ENGINE_NO_GLOBAL_NAMES
using namespace rte_host;  // the reference's names are global (RE/Math.h, RE/Shape.h, ...)
#endif
