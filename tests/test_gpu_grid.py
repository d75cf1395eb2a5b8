"""GPU parity of the exact path on the direction grids (render_exact_grid_kernel, csrc/rte_grid.cuh): diffuse scenes of any size,
spheres and triangles — SURVEY.md §8(f) row N3 (mesh acceleration) without a capacity limit.

The reference loops over every primitive (RE/Scene.h:218-257, RE/Shape.h:269-283); here every ray tests the primitives of ONE cell
of its origin's direction grid, with the reference's exact FP64 arithmetic and its tie rule applied by index.  Bar: bit-exact
prim_id / light_mask / depth64 / hdr64 and identical algorithmic counters against the CPU checker."""
import time

import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.scene import SceneArrays

pytestmark = pytest.mark.gpu
COUNTERS = ("primary_rays", "shadow_rays", "sphere_tests", "plane_tests", "triangle_tests")
OUTS = ("hdr64", "prim_id", "depth64", "normal", "albedo", "light_mask")


def check_rows(engine, oracle, arrays, cam, row_pairs, flags=0):
    for r0 in row_pairs:
        g = engine.render_host(outputs=OUTS, row0=r0, row1=r0 + 2, flags=flags)
        s = engine.stats()
        o = oracle.render(arrays, cam, row0=r0, row1=r0 + 2, want_aovs=True)
        assert np.array_equal(g["prim_id"], o["prim_id"]), (r0, np.count_nonzero(g["prim_id"] != o["prim_id"]))
        assert np.array_equal(g["light_mask"], o["light_mask"]), r0
        hit = o["prim_id"] != _abi.PRIM_MISS
        assert np.array_equal(g["depth64"][hit], o["depth"][hit]), r0
        assert np.array_equal(g["hdr64"], o["hdr"]), (r0, np.abs(g["hdr64"] - o["hdr"]).max())
        np.testing.assert_allclose(g["normal"][:, :3], o["normal"], rtol=0, atol=1e-6)
        np.testing.assert_allclose(g["albedo"][:, :3], o["albedo"], rtol=1e-6, atol=1e-7)
        for k in (COUNTERS if len(arrays.lights) <= 16 else COUNTERS[:1]):  # (the checker's counting pass walks the first 16 lights: its light_mask width)
            assert s[k] == o[k], (r0, k, s[k], o[k])


@pytest.mark.parametrize("cfg", ["c2", "c3", "c5", "c4_blob", "c4_box", "mixed", "odd"])
def test_exact_grid_kernel_equals_oracle(engine, oracle, cfg):
    """RTE_FLAG_EXACT_GRID on the scenes the view-based kernels also handle: same bits as the oracle (whole small frames)."""
    if cfg == "c2":
        arrays, cam = scenes.config2(scale=0.1)
    elif cfg == "c3":
        arrays, cam = scenes.config3(scale=0.1)
    elif cfg == "c5":
        arrays, cam = scenes.config5(scale=1 / 48)
    elif cfg == "c4_blob":
        arrays, cam = scenes.config4("blob", scale=0.1)
    elif cfg == "c4_box":
        arrays, cam = scenes.config4("box", scale=0.1)
    elif cfg == "mixed":
        arrays, cam = scenes.mixed_scene(160, 90, 80.0)
    else:
        arrays, cam = scenes.synthetic_room(203, 117, 100.0, 300, 3, 77)
    engine.upload(arrays, cam)
    g = engine.render_host(outputs=OUTS + ("hdr", "rgb8"), tonemap_op=_abi.TM_ACES, flags=_abi.FLAG_EXACT_GRID)
    s = engine.stats()
    o = oracle.render(arrays, cam, want_aovs=True)
    assert np.array_equal(g["prim_id"], o["prim_id"]) and np.array_equal(g["light_mask"], o["light_mask"])
    hit = o["prim_id"] != _abi.PRIM_MISS
    assert np.array_equal(g["depth64"][hit], o["depth"][hit]) and np.all(np.isinf(g["depth64"][~hit]))
    assert np.array_equal(g["hdr64"], o["hdr"]), np.abs(g["hdr64"] - o["hdr"]).max()
    assert np.array_equal(g["rgb8"], oracle.tonemap_one(_abi.TM_ACES, o["hdr"]))
    for k in COUNTERS:
        assert s[k] == o[k], (k, s[k], o[k])
    # and equal to the view-based exact kernels
    v = engine.render_host(outputs=("hdr64", "prim_id"), flags=_abi.FLAG_EXACT_FP64)
    assert np.array_equal(v["hdr64"], g["hdr64"]) and np.array_equal(v["prim_id"], g["prim_id"])


def grid_mesh(n, size=14.0, seed=5):
    """Procedural height-field mesh of 2 n^2 triangles (a wavy sheet): vertices FP32-exact."""
    rng = scenes.SplitMix64(seed)
    k = np.array([[rng.range(0.3, 1.2) for _ in range(2)] for _ in range(3)])
    xs = np.linspace(-size, size, n + 1)
    X, Z = np.meshgrid(xs, xs, indexing="ij")
    Y = sum(0.9 * np.sin(k[i, 0] * X + k[i, 1] * Z + i) for i in range(3))
    v = np.stack([X, Z * 0.7 - 2.0, Y], axis=-1).reshape(-1, 3).astype(np.float32).astype(np.float64)  # a wavy wall facing the camera
    idx = []
    for i in range(n):
        for j in range(n):
            a, b, c, d = i * (n + 1) + j, (i + 1) * (n + 1) + j, (i + 1) * (n + 1) + j + 1, i * (n + 1) + j + 1
            idx += [a, d, c, a, c, b]
    return v, np.array(idx, dtype=np.int32)


def mesh_scene(n, width=320, height=180, n_lights=2):
    mats, planes = scenes.main_planes(specular=0.0, shininess=0.128, ior=1.5)
    v, idx = grid_mesh(n)
    allm = np.concatenate([mats, np.array([scenes._material((0.2, 0.5, 1.0))], dtype=_abi.material_dt)])
    models = np.zeros(1, dtype=_abi.model_dt)
    models[0] = (0, len(v), 0, len(idx), (0.0, 0.0, 6.0), len(allm) - 1, 0)
    lights = np.zeros(n_lights, dtype=_abi.light_dt)
    lights[0] = ((0, 8, -5), (1, 1, 1), 150.0)
    if n_lights > 1:
        lights[1] = ((-6, 3, 0), (1, 0.9, 0.8), 90.0)
    cam = scenes.Camera((0.0, 0.0, -25.0), width / 2.0, width, height, 0.0, 200.0)
    cam.antiAliasingAmount = 1
    return SceneArrays(materials=allm, planes=planes, models=models, model_vertices=v, model_indices=idx, lights=lights), cam


@pytest.mark.parametrize("n", [71, 224])
def test_large_meshes_without_a_capacity_limit(engine, oracle, n):
    """10 082- and 100 352-triangle models (round 1 refused anything beyond ~4.7 k triangles): sampled rows bit-exact against the
    reference build's brute force, and the render time grows far slower than the triangle count."""
    arrays, cam = mesh_scene(n)
    assert arrays.counts()["triangles"] == 2 * n * n
    engine.upload(arrays, cam)
    check_rows(engine, oracle, arrays, cam, (40, 95, 140) if n == 71 else (95,))
    ms = float("inf")
    for _ in range(3):  # (best of three: a timing comparison inside a test suite has to survive a noisy neighbour)
        g = engine.render_host(outputs=("hdr64", "prim_id"))
        ms = min(ms, engine.stats()["kernel_ms"])
    assert ((g["prim_id"] >> 30) == _abi.HIT_TRIANGLE).mean() > 0.15
    test_large_meshes_without_a_capacity_limit.ms = getattr(test_large_meshes_without_a_capacity_limit, "ms", {})
    test_large_meshes_without_a_capacity_limit.ms[n] = ms
    if 71 in test_large_meshes_without_a_capacity_limit.ms and n == 224:
        ratio = ms / test_large_meshes_without_a_capacity_limit.ms[71]
        assert ratio < 8.0, f"10x the triangles cost {ratio:.1f}x the time"   # sub-linear (measured 2.0x): the reference's loop is linear


def test_16k_spheres_and_many_lights(engine, oracle):
    """16 384 spheres x 21 views would need 5.5 MB of shared memory per CTA in the view kernels: the grid kernel takes it; more than
    16 lights (light_mask covers the first 16)."""
    arrays, cam = scenes.synthetic_room(240, 135, 120.0, 16384, 20, 2468, radius_scale=0.5)
    engine.upload(arrays, cam)
    check_rows(engine, oracle, arrays, cam, (30, 90))
    # the production path of the same scene: 20 lights > 16 -> the exact grid kernel as well; RGB8 byte-exact
    o = oracle.render(arrays, cam, row0=60, row1=64)
    p = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES, row0=60, row1=64)["rgb8"]
    assert np.array_equal(p, oracle.tonemap_one(_abi.TM_ACES, o["hdr"]))
