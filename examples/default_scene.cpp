// The reference's main() scene (RE/RaytracingEngine.cpp:216-330) written against include/rte/engine.hpp:
// five planes, the stand-in box.obj (12 triangles; the upstream file is git-ignored), two lights,
// 1000x1000, antiAliasingAmount = 1, all seven tonemap operators to <name>.ppm.  Needs a B200.
//   g++ -std=c++20 -Iinclude examples/default_scene.cpp -Lraytracingengine_b200 -lrte_b200 -Wl,-rpath,$PWD/raytracingengine_b200
#include <chrono>
#include <cstdio>

#include "rte/engine.hpp"

using namespace rte_host;

int main(int argc, char** argv) {
    const std::size_t WIDTH = 1000, HEIGHT = 1000;  // RE/RaytracingEngine.cpp:67-68
    Camera camera(Vec3(0, 0, -25), 500, WIDTH, HEIGHT, 0, 200);
    camera.antiAliasingAmount = 1;
    try {
        Scene scene(camera);
        Material boxMat{Vec3(0, 0, 1), 128.0, 0.5, 0.0, 1.5};
        std::vector<Vec3> v = {Vec3(-1, -1, -1), Vec3(1, -1, -1), Vec3(1, 1, -1), Vec3(-1, 1, -1), Vec3(-1, -1, 1), Vec3(1, -1, 1), Vec3(1, 1, 1), Vec3(-1, 1, 1)};
        const int quads[6][4] = {{0, 3, 2, 1}, {4, 5, 6, 7}, {0, 1, 5, 4}, {2, 3, 7, 6}, {1, 2, 6, 5}, {0, 4, 7, 3}};
        std::vector<int> idx;
        for (auto& q : quads) for (int k = 1; k < 3; ++k) { idx.push_back(q[0]); idx.push_back(q[k]); idx.push_back(q[k + 1]); }  // tinyobj fan
        Model box(idx, Transform{Vec3(0, 0, 10), Vec3(), Vec3(1, 1, 1)}, boxMat, v);
        scene.AddModel(box);
        const double dirs[5][3] = {{0, 0, -1}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}};
        const double cols[5][3] = {{1, 1, 1}, {0, 1, 0}, {0, 0, 1}, {1, 1, 1}, {1, 1, 1}};
        for (int i = 0; i < 5; ++i) {
            Material m{Vec3(cols[i][0], cols[i][1], cols[i][2]), 0.128, 0.01, 0.0, 1.5};
            Plane p(Vec3(-15 * dirs[i][0], -15 * dirs[i][1], -15 * dirs[i][2]), Vec3(dirs[i][0], dirs[i][1], dirs[i][2]), m);
            scene.AddPlane(p);
        }
        Light l1(Vec3(0, 0, -5), Vec3(1, 1, 1), 150), l2(Vec3(-2, 2, -5), Vec3(1, 1, 1), 150);
        scene.AddLight(l1);
        scene.AddLight(l2);
        const auto t0 = std::chrono::high_resolution_clock::now();
        std::vector<Vec3> pixels = scene.RenderImage();
        const auto t1 = std::chrono::high_resolution_clock::now();
        std::printf("Temps de generation de l'image : %.3f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
        const char* names[7] = {"simple", "reinhard_simple", "reinhard_extended", "reinhard_extended_luminance", "reinhard_jodie", "uncharted2", "aces"};
        auto all = tonemapAll(scene, pixels);
        const std::string dir = argc > 1 ? std::string(argv[1]) + "/" : "";
        for (int i = 0; i < 7; ++i) writePPM(dir + names[i] + ".ppm", all[i], WIDTH, HEIGHT);
    } catch (const std::runtime_error& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
