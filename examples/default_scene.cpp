// The reference's main() scene (RE/RaytracingEngine.cpp:216-330) written against include/rte/engine.hpp exactly like upstream
// writes it: LoadObject("box.obj") from disk, five planes, two lights, 1000x1000, all seven tonemap operators to <name>.ppm.
// antiAliasingAmount = 1 (deterministic).  Needs a B200.
//   g++ -std=c++20 -Iinclude examples/default_scene.cpp -Lraytracingengine_b200 -lrte_b200 -Wl,-rpath,$PWD/raytracingengine_b200
//   ./default_scene <output dir> [box.obj]      (without an .obj the stand-in cube of SURVEY D5 is written to <output dir>/box.obj first)
#include <chrono>
#include <cstdio>
#include <iostream>

#include "rte/engine.hpp"

constexpr auto WIDTH = 1000;   // RE/RaytracingEngine.cpp:67-68
constexpr auto HEIGHT = 1000;

static void write_stand_in_box(const std::string& path) {  // axis-aligned cube [-1,1]^3, 8 v / 6 quads (the upstream file is git-ignored)
    std::ofstream f(path);
    f << "# stand-in for the reference's box.obj\no box\n"
         "v -1 -1 -1\nv 1 -1 -1\nv 1 1 -1\nv -1 1 -1\nv -1 -1 1\nv 1 -1 1\nv 1 1 1\nv -1 1 1\n"
         "f 1 4 3 2\nf 5 6 7 8\nf 1 2 6 5\nf 3 4 8 7\nf 2 3 7 6\nf 1 5 8 4\n";
}

int main(int argc, char** argv) {
    const std::string dir = argc > 1 ? std::string(argv[1]) + "/" : "";
    std::string objPath = argc > 2 ? argv[2] : dir + "box.obj";
    try {
        if (argc <= 2) write_stand_in_box(objPath);
        Vec3 origin(0, 0, -25);
        Camera camera(origin, 500, WIDTH, HEIGHT, 0, 200);
        camera.antiAliasingAmount = 1;
        Scene scene(camera);

        Material monkeyMaterial{.color = Vec3(0, 0, 1), .shininess = 128.0, .specular = 0.5, .transparency = 0.0, .refractiveIndex = 1.5};
        Transform monkeyTransform{Vec3(0, 0, 10), Vec3(0, 0, 0), Vec3(1, 1, 1)};
        auto monkey = LoadObject(objPath, monkeyTransform, monkeyMaterial);
        scene.AddModel(monkey);

        double distance = 15;
        std::vector<Vec3> normalDirections = {Vec3(0, 0, -1), Vec3(1, 0, 0), Vec3(-1, 0, 0), Vec3(0, 1, 0), Vec3(0, -1, 0)};
        std::vector<Vec3> planeColors = {Vec3(1, 1, 1), Vec3(0, 1, 0), Vec3(0, 0, 1), Vec3(1, 1, 1), Vec3(1, 1, 1)};
        for (size_t i = 0; i < normalDirections.size(); i++) {
            Material mat;
            mat.transparency = 0.0; mat.specular = 0.01; mat.color = planeColors[i]; mat.shininess = 0.128; mat.refractiveIndex = 1.5;
            Plane plane(normalDirections[i] * -distance, normalDirections[i], mat);
            scene.AddPlane(plane);
        }
        auto firstLight = Light(Vec3(0, 0, -5), Vec3(1, 1, 1), 150);
        auto secondLight = Light(Vec3(-2, 2, -5), Vec3(1, 1, 1), 150);
        scene.AddLight(firstLight);
        scene.AddLight(secondLight);

        const auto t0 = std::chrono::high_resolution_clock::now();
        std::vector<Vec3> pixels = scene.RenderImage();
        const auto t1 = std::chrono::high_resolution_clock::now();
        std::printf("Temps de generation de l'image : %.3f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count());

        std::vector<std::string> tonemapNames = {"simple", "reinhard_simple", "reinhard_extended", "reinhard_extended_luminance", "reinhard_jodie", "uncharted2", "aces"};
        std::vector<std::vector<Color>> allTonemapped = tonemapAll(pixels);
        for (size_t i = 0; i < allTonemapped.size(); i++) writePPM(dir + tonemapNames[i] + ".ppm", allTonemapped[i], WIDTH, HEIGHT);
    } catch (const std::runtime_error& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
