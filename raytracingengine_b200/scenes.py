"""Deterministic scene fixtures for the BASELINE.json configs.

All scene scalars are rounded to FP32 before use so that the FP64 oracle and the GPU consume
identical values; the PRNG is a self-contained SplitMix64 -> (x >> 40) * 2**-24 stream, so scenes are
identical across compilers and hosts.  These functions only build descriptions (SceneArrays +
Camera); nothing here intersects or shades.
"""
from __future__ import annotations

import numpy as np

from . import _abi
from .scene import Camera, SceneArrays

MASK64 = (1 << 64) - 1


class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & MASK64

    def next_u64(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & MASK64
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
        return z ^ (z >> 31)

    def uniform(self) -> float:
        """[0,1) with 24 random bits: exactly representable in FP32."""
        return (self.next_u64() >> 40) * (1.0 / (1 << 24))

    def range(self, lo: float, hi: float) -> float:
        return float(np.float32(lo + (hi - lo) * self.uniform()))


def f32(x) -> float:
    return float(np.float32(x))


def _material(color, shininess=128.0, specular=0.0, transparency=0.0, ior=1.0):
    return (tuple(f32(c) for c in color), f32(shininess), f32(specular), f32(transparency), f32(ior))


def room_planes(mat_index_of, with_ceiling=False):
    """floor y=-15, back z=15, left x=-15, right x=15: (point, normal)."""
    planes = [((0, -15, 0), (0, 1, 0)), ((0, 0, 15), (0, 0, -1)), ((-15, 0, 0), (1, 0, 0)), ((15, 0, 0), (-1, 0, 0))]
    if with_ceiling:
        planes.append(((0, 15, 0), (0, -1, 0)))
    arr = np.zeros(len(planes), dtype=_abi.plane_dt)
    for i, (p, n) in enumerate(planes):
        arr[i] = (p, n, mat_index_of(i), 0)
    return arr


def synthetic_room(width, height, focal, n_spheres, n_lights, seed, radius_scale=1.0, fixed_light=None):
    """C2/C3/C5 family: random diffuse spheres in a 4-plane white room, point lights at y=10."""
    rng = SplitMix64(seed)
    mats = np.zeros(n_spheres + 1, dtype=_abi.material_dt)
    sph = np.zeros(n_spheres, dtype=_abi.sphere_dt)
    for i in range(n_spheres):
        c = (rng.range(-14, 14), rng.range(-14, 14), rng.range(-10, 14))
        r = f32(rng.range(0.2, 0.8) * radius_scale)
        alb = (rng.range(0.1, 1), rng.range(0.1, 1), rng.range(0.1, 1))
        mats[i] = _material(alb)
        sph[i] = (c, r, i, 0)
    mats[n_spheres] = _material((1, 1, 1))
    planes = room_planes(lambda i: n_spheres)
    lights = np.zeros(n_lights, dtype=_abi.light_dt)
    if fixed_light is not None:
        lights[0] = (fixed_light, (1, 1, 1), 150.0)
    else:
        lrng = SplitMix64(seed ^ 0x5EED11647)
        for j in range(n_lights):
            lights[j] = ((lrng.range(-8, 8), 10.0, lrng.range(-12, -4)), (1, 1, 1), 150.0)
    cam = Camera((0.0, 0.0, -25.0), float(focal), int(width), int(height), 0.0, 200.0)
    cam.antiAliasingAmount = 1
    return SceneArrays(materials=mats, spheres=sph, planes=planes, lights=lights), cam


def config2(scale=1.0):
    """C2: 3840x2160, 1024 spheres + 4 planes, 4 lights (ACES). `scale` shrinks the image only."""
    return synthetic_room(int(3840 * scale), int(2160 * scale), 1920 * scale, 1024, 4, 1234)


def config3(scale=1.0):
    """C3: 3840x2160, first 64 spheres of the C2 stream with doubled radii, 1 light (0,10,-8)."""
    return synthetic_room(int(3840 * scale), int(2160 * scale), 1920 * scale, 64, 1, 1234, radius_scale=2.0,
                          fixed_light=(0.0, 10.0, -8.0))


def config5(scale=1.0):
    """C5: 7680x4320, 4096 spheres, 8 lights, seed 5678."""
    return synthetic_room(int(7680 * scale), int(4320 * scale), 3840 * scale, 4096, 8, 5678)


# ---- meshes -------------------------------------------------------------------------------------
def box_mesh():
    """Stand-in for the missing box.obj: axis-aligned cube [-1,1]^3, 8 v / 6 quads."""
    v = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1],
                  [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=np.float64)
    quads = [[0, 3, 2, 1], [4, 5, 6, 7], [0, 1, 5, 4], [2, 3, 7, 6], [1, 2, 6, 5], [0, 4, 7, 3]]
    return v, quads


def fan_triangulate(faces):
    """Fan triangulation as tinyobjloader does it (RE/tiny_obj_loader.h:908-931)."""
    idx = []
    for f in faces:
        for k in range(1, len(f) - 1):
            idx += [f[0], f[k], f[k + 1]]
    return np.array(idx, dtype=np.int32)


def blob_mesh(level=3, radius=3.0, seed=99):
    """~1k-triangle closed mesh standing in for the missing monkey.obj: icosphere (20*4^level
    triangles) with smoothly displaced vertices; vertex coordinates are FP32-exact like OBJ input."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    verts = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
             (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    verts = [tuple(np.array(v) / np.linalg.norm(v)) for v in verts]
    faces = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2),
             (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5),
             (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(level):
        cache, new_faces = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = (np.array(verts[a]) + np.array(verts[b])) / 2
                verts.append(tuple(m / np.linalg.norm(m)))
                cache[key] = len(verts) - 1
            return cache[key]
        for a, b, c in faces:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            new_faces += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        faces = new_faces
    v = np.array(verts, dtype=np.float64)
    rng = SplitMix64(seed)
    k = np.array([[rng.range(-2, 2) for _ in range(3)] for _ in range(4)])
    ph = np.array([rng.range(0, 6.28) for _ in range(4)])
    disp = 1.0 + 0.12 * sum(np.sin(v @ k[i] * 2.0 + ph[i]) for i in range(4))
    v = (v * disp[:, None] * radius).astype(np.float32).astype(np.float64)
    return v, [list(f) for f in faces]


def write_obj(path, vertices, faces):
    """Minimal OBJ writer (1-based `v` / `f` records)."""
    with open(path, "w") as f:
        f.write("# generated by raytracingengine_b200.scenes\no mesh\n")
        for x, y, z in vertices:
            f.write(f"v {float(np.float32(x))!r} {float(np.float32(y))!r} {float(np.float32(z))!r}\n")
        for face in faces:
            f.write("f " + " ".join(str(i + 1) for i in face) + "\n")


def main_planes(specular, shininess, ior):
    """The five planes of main(), RE/RaytracingEngine.cpp:253-284: pos = -15*dir, normal = dir."""
    dirs = [(0, 0, -1), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0)]
    cols = [(1, 1, 1), (0, 1, 0), (0, 0, 1), (1, 1, 1), (1, 1, 1)]
    mats = np.zeros(5, dtype=_abi.material_dt)
    planes = np.zeros(5, dtype=_abi.plane_dt)
    for i, (d, c) in enumerate(zip(dirs, cols)):
        mats[i] = _material(c, shininess, specular, 0.0, ior)
        planes[i] = (tuple(-15.0 * x for x in d), d, i, 0)
    return mats, planes


def _with_mesh(mats, planes, verts, faces, mesh_material, lights, cam):
    idx = fan_triangulate(faces)
    allm = np.concatenate([mats, np.array([mesh_material], dtype=_abi.material_dt)])
    models = np.zeros(1, dtype=_abi.model_dt)
    models[0] = (0, len(verts), 0, len(idx), (0.0, 0.0, 10.0), len(allm) - 1, 0)
    return SceneArrays(materials=allm, planes=planes, models=models, model_vertices=verts, model_indices=idx,
                       lights=lights), cam


def config1(width=1000, height=1000):
    """C1: exactly main()'s scene with the stand-in box.obj, AA forced to 1."""
    mats, planes = main_planes(specular=0.01, shininess=0.128, ior=1.5)
    verts, faces = box_mesh()
    lights = np.zeros(2, dtype=_abi.light_dt)
    lights[0] = ((0, 0, -5), (1, 1, 1), 150.0)
    lights[1] = ((-2, 2, -5), (1, 1, 1), 150.0)
    cam = Camera((0.0, 0.0, -25.0), 500.0 * width / 1000.0, width, height, 0.0, 200.0)
    cam.antiAliasingAmount = 1
    return _with_mesh(mats, planes, verts, faces, _material((0, 0, 1), 128.0, 0.5, 0.0, 1.5), lights, cam)


def config4(mesh="blob", scale=1.0, specular=False):
    """C4: 1920x1080 OBJ-mesh scene, 5 planes of main() (diffuse unless `specular`), 1 light."""
    mats, planes = main_planes(specular=0.01 if specular else 0.0, shininess=0.128, ior=1.5)
    verts, faces = box_mesh() if mesh == "box" else blob_mesh()
    lights = np.zeros(1, dtype=_abi.light_dt)
    lights[0] = ((0, 0, -5), (1, 1, 1), 150.0)
    cam = Camera((0.0, 0.0, -25.0), 960.0 * scale, int(1920 * scale), int(1080 * scale), 0.0, 200.0)
    cam.antiAliasingAmount = 1
    return _with_mesh(mats, planes, verts, faces, _material((0, 0, 1), 128.0, 0.5 if specular else 0.0, 0.0, 1.5),
                      lights, cam)


def mixed_scene(width, height, focal, n_spheres=200, n_lights=3, seed=4242, mesh_level=2):
    """Every primitive class at once (diffuse): a synthetic room of spheres + planes, a mesh Model inside it and a few
    free Triangles — the canonical order spheres, planes, triangles, models with hits in each class."""
    arrays, cam = synthetic_room(width, height, focal, n_spheres, n_lights, seed)
    verts, faces = blob_mesh(level=mesh_level, radius=4.0, seed=seed + 1)
    idx = fan_triangulate(faces)
    mats = np.concatenate([arrays.materials, np.array([_material((0.2, 0.4, 1.0)), _material((1.0, 0.5, 0.1))], dtype=_abi.material_dt)])
    models = np.zeros(1, dtype=_abi.model_dt)
    models[0] = (0, len(verts), 0, len(idx), (2.0, -3.0, 2.0), len(mats) - 2, 0)
    rng = SplitMix64(seed + 2)
    tris = np.zeros(6, dtype=_abi.triangle_dt)
    for i in range(len(tris)):
        c = np.array([rng.range(-10, 10), rng.range(-10, 10), rng.range(-6, 10)])
        v = [tuple(f32(x) for x in c + np.array([rng.range(-3, 3), rng.range(-3, 3), rng.range(-3, 3)])) for _ in range(3)]
        tris[i] = (v[0], v[1], v[2], (0.0, 0.0, 0.0), len(mats) - 1, 0)
    return SceneArrays(materials=mats, spheres=arrays.spheres, planes=arrays.planes, triangles=tris, models=models,
                       model_vertices=verts, model_indices=idx, lights=arrays.lights), cam


def glass_scene(width=96, height=72, focal=None, max_lights=2):
    """Row N1: refraction, Fresnel reflection and coloured transmittance.  A 5-plane room (slightly
    glossy floor), glass / half-transparent / mirror / diffuse spheres, a half-transparent box model and two lights:
    every branch of TraceRay (RE/Scene.h:131-198) and computeTransmittance (RE/Scene.h:35-77) is taken."""
    focal = float(focal if focal is not None else width * 0.75)
    mats = np.array([
        _material((1, 1, 1), 32.0, 0.05, 0.0, 1.0),        # 0 floor: faint mirror
        _material((0.9, 0.3, 0.3)),                          # 1 walls
        _material((0.3, 0.9, 0.3)),                          # 2
        _material((1, 1, 1), 128.0, 0.0, 0.9, 1.5),        # 3 glass
        _material((0.2, 0.6, 1.0), 64.0, 0.0, 0.5, 1.33),  # 4 half-transparent, water-like
        _material((1, 1, 1), 256.0, 0.8, 0.0, 1.0),        # 5 mirror
        _material((1.0, 0.8, 0.2)),                          # 6 diffuse
        _material((0.8, 0.9, 1.0), 16.0, 0.0, 0.7, 1.1),   # 7 box: thin, weakly refracting
        _material((1, 1, 1), 8.0, 0.0, 1.0, 2.4),          # 8 fully transparent (transparency == 1: no local light)
    ], dtype=_abi.material_dt)
    planes = np.zeros(5, dtype=_abi.plane_dt)
    for i, (p, n, m) in enumerate([((0, -6, 0), (0, 1, 0), 0), ((0, 0, 14), (0, 0, -1), 1), ((-9, 0, 0), (1, 0, 0), 2),
                                   ((9, 0, 0), (-1, 0, 0), 1), ((0, 9, 0), (0, -1, 0), 2)]):
        planes[i] = (p, n, m, 0)
    sph = np.zeros(7, dtype=_abi.sphere_dt)
    for i, (c, r, m) in enumerate([((-3.0, -3.5, 4.0), 2.5, 3), ((3.5, -4.0, 2.0), 2.0, 4), ((0.5, 1.5, 9.0), 3.0, 5),
                                   ((-5.5, 2.0, 7.0), 1.5, 6), ((5.0, 3.0, 6.0), 1.25, 8), ((0.0, -4.5, -2.0), 1.5, 3),
                                   ((-2.5, -3.0, 4.5), 0.75, 6)]):  # the last one sits INSIDE the first glass sphere
        sph[i] = (tuple(f32(x) for x in c), f32(r), m, 0)
    verts, faces = box_mesh()
    idx = fan_triangulate(faces)
    models = np.zeros(1, dtype=_abi.model_dt)
    models[0] = (0, len(verts), 0, len(idx), (4.0, 1.0, 1.0), 7, 0)
    lights = np.zeros(2, dtype=_abi.light_dt)
    lights[0] = ((-4.0, 7.0, -3.0), (1, 1, 1), 120.0)
    lights[1] = ((5.0, 6.0, 0.0), (1.0, 0.9, 0.8), 80.0)
    cam = Camera((0.0, 0.0, -16.0), focal, int(width), int(height), 0.0, 200.0)
    cam.antiAliasingAmount = 1
    return SceneArrays(materials=mats, spheres=sph, planes=planes, models=models, model_vertices=verts, model_indices=idx,
                       lights=lights[:max_lights]), cam


CONFIGS = {"c1": config1, "c2": config2, "c3": config3, "c4": config4, "c5": config5}
