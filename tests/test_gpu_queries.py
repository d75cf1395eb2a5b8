"""GPU parity of the ray-query entry points (rte_intersect_closest / rte_intersect_any_before = Scene::IntersectClosest /
IntersectAnyBefore, RE/Scene.h:218-276) and the reference's edge-case known-answer vectors ON THE CUDA PATH
(tests/golden/kat.json, generated from the reference's own functions).

Bar: hit (type, index), FP64 distance, normal and hit point bit-exact; any-hit answers equal."""
import json
import os

import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.scene import Camera, SceneArrays

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(GOLD, "kat.json")) as f:
        return json.load(f)


def _kat_scenes():
    mats = np.zeros(1, dtype=_abi.material_dt)
    mats[0] = ((1, 1, 1), 128.0, 0.0, 0.0, 1.0)
    sph = np.zeros(1, dtype=_abi.sphere_dt); sph[0] = ((0, 0, 0), 1.0, 0, 0)
    pln = np.zeros(1, dtype=_abi.plane_dt); pln[0] = ((0, -15, 0), (0, 1, 0), 0, 0)
    tri = np.zeros(1, dtype=_abi.triangle_dt); tri[0] = ((-1, -1, 0), (1, -1, 0), (0, 1, 0), (0, 0, 0), 0, 0)
    return {"sphere": SceneArrays(materials=mats, spheres=sph), "plane": SceneArrays(materials=mats, planes=pln),
            "triangle": SceneArrays(materials=mats, triangles=tri)}


@pytest.mark.parametrize("shape", ["sphere", "plane", "triangle"])
def test_shape_kat_on_the_cuda_path(engine, kat, shape):
    """Sphere / Plane / Triangle::Intersect edge cases (RE/Shape.h:72-98,149-159,202-220): tangent ray (disc == 0) hit, origin
    at the centre -> far root, origin at (0,0,-1) looking away -> miss, origin on the plane (t == 0) hit, parallel ray miss,
    ray through vertex v0 (u = v = 0) hit, back-face hit — the reference's own answers, through closest_kernel."""
    k = kat["shapes"][shape]
    engine.upload(_kat_scenes()[shape], Camera((0.0, 0.0, -25.0), 500.0, 16, 16, 0.0, 200.0))
    h = engine.intersect_closest(k["origins"], k["directions"])
    assert h["prim_id"].tolist() == k["prim_id"]
    hit = np.array(k["prim_id"]) != _abi.PRIM_MISS
    want_t = np.array([np.inf if v is None else v for v in k["distance"]])
    assert np.array_equal(h["distance"][hit], want_t[hit])
    assert np.all(np.isinf(h["distance"][~hit]))
    assert np.array_equal(h["normal"][hit], np.array(k["normal"])[hit])
    # IntersectAnyBefore (RE/Scene.h:259-276) on the same rays: 0 < t < maxDist
    o, d = np.array(k["origins"], np.float64), np.array(k["directions"], np.float64)
    for md in (0.5, 4.5, 1e3):
        got = engine.intersect_any_before(o, d, np.full(len(o), md))
        want = (hit & (want_t > 0.0) & (want_t < md)).astype(np.uint8)
        assert got.tolist() == want.tolist(), (shape, md)


def test_shape_kat_hand_values(engine):
    """The hand-derivable rows of SURVEY §8(c), without any fixture."""
    s = _kat_scenes()
    cam = Camera((0.0, 0.0, -25.0), 500.0, 16, 16, 0.0, 200.0)
    engine.upload(s["sphere"], cam)
    h = engine.intersect_closest([(0, 0, -5), (0, 0, 0), (0, 0, -1), (1, 0, -5)], [(0, 0, 1), (0, 0, 1), (0, 0, -1), (0, 0, 1)])
    assert h["distance"][:2].tolist() == [4.0, 1.0] and h["prim_id"][2] == _abi.PRIM_MISS and h["distance"][3] == 5.0
    assert h["normal"][1].tolist() == [0.0, 0.0, 1.0]  # from the centre: far root, outward normal
    engine.upload(s["plane"], cam)
    h = engine.intersect_closest([(3, -15, 2), (0, 0, 0)], [(0, 1, 0), (1, 0, 0)])
    assert h["distance"][0] == 0.0 and h["prim_id"][0] == (_abi.HIT_PLANE << 30) and h["prim_id"][1] == _abi.PRIM_MISS
    engine.upload(s["triangle"], cam)
    h = engine.intersect_closest([(0, 0, -3), (0, 0, 3), (-1, -1, -3)], [(0, 0, 1), (0, 0, -1), (0, 0, 1)])
    assert h["distance"].tolist() == [3.0, 3.0, 3.0] and h["normal"][0].tolist() == [0.0, 0.0, 1.0]


def test_camera_kat_on_the_cuda_path(engine, oracle, kat):
    """Camera::getRay KATs (RE/Math.h:99-121; default camera of main()) through the render kernel: a sphere centred on the
    camera is hit from inside at its far root, so the primary-hit normal IS the ray direction (FP32 AOV: 1e-7), and the
    albedo of a miss is backgroundColor(dir) (RE/Scene.h:30-33); depth / hit id equal the oracle's bit for bit."""
    cam = Camera((0.0, 0.0, -25.0), 500.0, 1000, 1000, 0.0, 200.0)
    cam.antiAliasingAmount = 1
    mats = np.zeros(1, dtype=_abi.material_dt)
    mats[0] = ((1, 1, 1), 128.0, 0.0, 0.0, 1.0)
    sph = np.zeros(1, dtype=_abi.sphere_dt)
    sph[0] = ((0.0, 0.0, -25.0), 7.0, 0, 0)
    arrays = SceneArrays(materials=mats, spheres=sph)
    engine.upload(arrays, cam)
    for key, want in kat["camera"].items():
        x, y = map(int, key.split(","))
        g = engine.render_host(outputs=("normal", "depth64", "prim_id"), row0=y, row1=y + 1)
        assert g["prim_id"][x] == (_abi.HIT_SPHERE << 30)
        np.testing.assert_allclose(g["normal"][x, :3], np.array(want), rtol=0, atol=1.2e-7)
        o = oracle.render(arrays, cam, row0=y, row1=y + 1, want_aovs=True)
        assert g["depth64"][x] == o["depth"][x]
        assert np.array_equal(g["prim_id"], o["prim_id"]) and np.array_equal(g["depth64"], o["depth"])
    # the same directions through the miss colour: sky = lerp(white, (.5,.7,1), 0.5 (dir.y + 1))
    engine.upload(SceneArrays(), cam)
    for key, want in kat["camera"].items():
        x, y = map(int, key.split(","))
        a = engine.render_host(outputs=("albedo",), row0=y, row1=y + 1)["albedo"][x, :3]
        t = 0.5 * (want[1] + 1.0)
        np.testing.assert_allclose(a, np.array([1.0, 1.0, 1.0]) * (1.0 - t) + np.array([0.5, 0.7, 1.0]) * t, rtol=2e-7)


def _query_rays(arrays, n, seed):
    """Random rays plus the awkward ones: origins inside spheres and exactly on planes, directions along triangle edges and
    through triangle vertices, non-unit and axis-parallel directions."""
    rng = np.random.default_rng(seed)
    o = rng.uniform(-14, 14, (n, 3))
    d = rng.normal(size=(n, 3))
    k = n // 8
    ns = len(arrays.spheres)
    if ns:
        pick = rng.integers(0, ns, k)
        o[:k] = arrays.spheres["center"][pick] + rng.uniform(-0.4, 0.4, (k, 1)) * arrays.spheres["radius"][pick][:, None] * np.array([1.0, 0.5, 0.25])
        o[k:2 * k] = arrays.spheres["center"][rng.integers(0, ns, k)]  # exactly at a centre: far root
    if len(arrays.planes):
        p = arrays.planes[rng.integers(0, len(arrays.planes), k)]
        nrm = p["normal"] / np.linalg.norm(p["normal"], axis=1)[:, None]
        tang = np.cross(nrm, rng.normal(size=(k, 3)))
        o[2 * k:3 * k] = p["point"] + tang * 3.0  # on the plane (t == 0 is a hit, RE/Shape.h:154)
        d[2 * k:2 * k + k // 2] = tang[:k // 2]     # ... and travelling inside it (|denom| <= 1e-6: miss)
    if len(arrays.models):
        m = arrays.models[0]
        v = arrays.model_vertices[m["first_vertex"]:m["first_vertex"] + m["vertex_count"]] + m["translate"]
        idx = arrays.model_indices[m["first_index"]:m["first_index"] + m["index_count"]].reshape(-1, 3)
        t = idx[rng.integers(0, len(idx), k)]
        a, b = v[t[:, 0]], v[t[:, 1]]
        o[3 * k:4 * k] = a - (b - a) * 2.0       # along an edge of a triangle
        d[3 * k:4 * k] = b - a
        o[4 * k:5 * k] = rng.uniform(-14, 14, (k, 3))
        d[4 * k:5 * k] = v[t[:, 2]] - o[4 * k:5 * k]   # through a vertex
    d[5 * k:5 * k + 3] = np.eye(3)
    d[5 * k + 3:5 * k + 6] = -np.eye(3) * 7.5
    md = rng.uniform(0.05, 45, n)
    return o, d, md


@pytest.mark.parametrize("cfg", ["mixed", "c2", "c4_blob"])
def test_query_kernels_equal_the_reference(engine, oracle, cfg):
    """closest_kernel / any_before_kernel against the CPU checker (the reference build when present) on >= 10k rays."""
    if cfg == "mixed":
        arrays, cam = scenes.mixed_scene(64, 36, 32.0)
    elif cfg == "c2":
        arrays, cam = scenes.config2(scale=0.02)
    else:
        arrays, cam = scenes.config4("blob", scale=0.05)
    engine.upload(arrays, cam)
    o, d, md = _query_rays(arrays, 12000, 20261105)
    want = oracle.intersect_closest(arrays, o, d)
    got = engine.intersect_closest(o, d)
    assert np.array_equal(got["prim_id"], want["prim_id"]), f"{np.count_nonzero(got['prim_id'] != want['prim_id'])} hit mismatches"
    hit = want["prim_id"] != _abi.PRIM_MISS
    assert hit.any() and (~hit).any()
    for f in ("distance", "normal", "point"):
        assert np.array_equal(got[f][hit], want[f][hit]), f
    assert np.all(np.isinf(got["distance"][~hit]))
    kinds = set((want["prim_id"][hit] >> 30).tolist())
    assert kinds == ({1, 2, 3} if cfg == "mixed" else {1, 2} if cfg == "c2" else {2, 3}), kinds
    a = engine.intersect_any_before(o, d, md)
    b = oracle.intersect_any_before(arrays, o, d, md)
    assert np.array_equal(a, b), f"{np.count_nonzero(a != b)} any-hit mismatches"
    assert 0 < a.sum() < len(a)
    m = got["material"][hit]
    assert m.min() >= 0 and m.max() < len(arrays.materials)


def test_queries_empty_and_single(engine):
    arrays, cam = scenes.config3(scale=0.02)
    engine.upload(arrays, cam)
    assert len(engine.intersect_closest(np.zeros((0, 3)), np.zeros((0, 3)))) == 0
    assert len(engine.intersect_any_before(np.zeros((0, 3)), np.zeros((0, 3)), np.zeros(0))) == 0
    h = engine.intersect_closest([(0.0, 0.0, -25.0)], [(0.0, 0.0, 1.0)])
    assert h["prim_id"][0] != _abi.PRIM_MISS and h["distance"][0] > 0
    engine.upload(SceneArrays(), cam)  # IntersectAnyBefore of an empty scene is false (RE/Scene.h:260-262)
    assert engine.intersect_any_before([(0, 0, 0)], [(0, 0, 1)], [10.0]).tolist() == [0]
    assert engine.intersect_closest([(0, 0, 0)], [(0, 0, 1)])["prim_id"][0] == _abi.PRIM_MISS
