// Exercises the query / operator surface of include/rte/engine.hpp on the known-answer vectors of SURVEY.md §8(c): prints one
// line per answer ("key v0 v1 ..."), tests/test_gpu_cpp_mirror.py compares them with tests/golden/kat.json.  Needs a B200.
#include <cstdio>

#include "rte/engine.hpp"

static void show(const char* key, const std::optional<HitInfo>& h) {
    if (!h) { std::printf("%s miss\n", key); return; }
    std::printf("%s %d %zu %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", key, (int)h->type, h->index, h->distance, h->normal.x, h->normal.y, h->normal.z,
                h->hitPoint.x, h->hitPoint.y, h->hitPoint.z);
}

int main() {
    try {
        Camera cam(Vec3(0, 0, -25), 500, 1000, 1000, 0, 200);
        cam.antiAliasingAmount = 1;
        Material white{Vec3(1, 1, 1)};
        {   // unit sphere at the origin (RE/Shape.h:72-98)
            Scene s(cam);
            Sphere sp(1.0, Vec3(0, 0, 0), white);
            s.AddSphere(sp);
            show("sphere_front", s.IntersectClosest(Rayon(Vec3(0, 0, -5), Vec3(0, 0, 1))));
            show("sphere_inside", s.IntersectClosest(Rayon(Vec3(0, 0, 0), Vec3(0, 0, 1))));
            show("sphere_away", s.IntersectClosest(Rayon(Vec3(0, 0, -1), Vec3(0, 0, -1))));
            show("sphere_tangent", s.IntersectClosest(Rayon(Vec3(1, 0, -5), Vec3(0, 0, 1))));
            std::printf("any_before %d %d\n", (int)s.IntersectAnyBefore(Rayon(Vec3(0, 0, -5), Vec3(0, 0, 1)), 10.0),
                        (int)s.IntersectAnyBefore(Rayon(Vec3(0, 0, -5), Vec3(0, 0, 1)), 3.0));
            show("pixel_centre", s.CalculatePixelDepth(500, 500, false));   // looks along +z: t = 24
            const auto h = s.CalculatePixelDepth(500, 500, false);
            std::printf("normalized %.17g\n", h ? h->normalizedDistance(cam) : -1.0);
            Scene copy = s;                                                  // value semantics: a copy renders the same
            const Vec3 a = s.GeneratePixelAt(500, 500), b = copy.GeneratePixelAt(500, 500);
            std::printf("copy_equal %d\n", (int)(a.x == b.x && a.y == b.y && a.z == b.z));
        }
        {   // plane p0 (0,-15,0), n (0,1,0) (RE/Shape.h:149-159)
            Scene s(cam);
            Plane pl(Vec3(0, -15, 0), Vec3(0, 1, 0), white);
            s.AddPlane(pl);
            show("plane_hit", s.IntersectClosest(Rayon(Vec3(0, 0, -25), Vec3(0, -1, 1).normalize())));
            show("plane_on", s.IntersectClosest(Rayon(Vec3(3, -15, 2), Vec3(0, 1, 0))));
            show("plane_parallel", s.IntersectClosest(Rayon(Vec3(0, 0, 0), Vec3(1, 0, 0))));
        }
        {   // triangle (-1,-1,0), (1,-1,0), (0,1,0) (RE/Shape.h:202-227)
            Scene s(cam);
            Triangle t(Vec3(-1, -1, 0), Vec3(1, -1, 0), Vec3(0, 1, 0), white);
            s.AddTriangle(t);
            show("tri_front", s.IntersectClosest(Rayon(Vec3(0, 0, -3), Vec3(0, 0, 1))));
            show("tri_back", s.IntersectClosest(Rayon(Vec3(0, 0, 3), Vec3(0, 0, -1))));
            show("tri_vertex", s.IntersectClosest(Rayon(Vec3(-1, -1, -3), Vec3(0, 0, 1))));
        }
        // the operator functions + toColor on the tonemap KAT inputs (RE/RaytracingEngine.cpp:113-163)
        const double in[7][3] = {{0, 0, 0}, {0.18, 0.18, 0.18}, {1, 1, 1}, {0.5, 1, 2}, {5, 5, 5}, {10, 0.1, 0.01}, {0.003, 0.5, 30}};
        for (int i = 0; i < 7; ++i) {
            const Vec3 c(in[i][0], in[i][1], in[i][2]);
            const Vec3 r[7] = {simple(c), reinhardSimple(c), reinhardExtended(c, 5.0), reinhardExtendedLuminance(c, 5.0), reinhardJodie(c), uncharted2(c), aces_approx(c)};
            std::printf("ops %d", i);
            for (int op = 0; op < 7; ++op) { const Color q = toColor(r[op]); std::printf(" %d %d %d", q.r, q.g, q.b); }
            std::printf("\n");
        }
        std::vector<Vec3> px;
        for (int i = 0; i < 7; ++i) px.emplace_back(in[i][0], in[i][1], in[i][2]);
        const auto all = tonemapAll(px);
        const auto one = tonemap(px);
        for (int i = 0; i < 7; ++i) {
            std::printf("all %d", i);
            for (int op = 0; op < 7; ++op) std::printf(" %d %d %d", all[op][i].r, all[op][i].g, all[op][i].b);
            std::printf(" | %d %d %d\n", one[i].r, one[i].g, one[i].b);
        }
        std::printf("luminance %.17g\n", luminance(Vec3(0.5, 1, 2)));
    } catch (const std::exception& e) {
        std::printf("error %s\n", e.what());
        return 1;
    }
    return 0;
}
