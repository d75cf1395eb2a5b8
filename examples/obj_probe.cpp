// LoadObject (include/rte/engine.hpp) on one file: "<n vertices> <n indices>", then the vertices (%.17g) and the indices.
// Host only (no GPU call); tests/test_host.py compares it with the reference's LoadObject (tinyobjloader) on the same file.
#include <cstdio>

#include "rte/engine.hpp"

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    try {
        const Model m = LoadObject(argv[1]);
        std::printf("%zu %zu\n", m.positions().size(), m.indices().size());
        for (const Vec3& v : m.positions()) std::printf("%.17g %.17g %.17g\n", v.x, v.y, v.z);
        for (int i : m.indices()) std::printf("%d\n", i);
    } catch (const std::runtime_error& e) {
        std::printf("error: %s\n", e.what());
        return 1;
    }
    return 0;
}
