"""Randomised scenes through the FP32 production kernels against the exact path (which the other tests pin to the reference
build): every interval decision has to come out as the reference's — hit ids and visibility masks identical, bytes within 1 LSB.
Small frames, many seeds: spheres of all sizes (some around the camera or a light), free triangles, planes at odd angles, lights
close to surfaces, odd biases."""
import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes

import os

pytestmark = pytest.mark.gpu
N_DIFFUSE, N_REFLECTIVE = (int(x) for x in os.environ.get("RTE_FUZZ_SEEDS", "24,16").split(","))  # (a longer campaign: RTE_FUZZ_SEEDS=400,300)
OUTS = ("hdr", "rgb8", "prim_id", "light_mask")


def random_scene(seed, reflective):
    rng = np.random.default_rng(seed)
    W, H = int(rng.integers(40, 120)), int(rng.integers(30, 90))
    cam = scenes.Camera(tuple(rng.uniform(-3, 3, 3) + np.array([0, 0, -30.0])), float(rng.uniform(0.4, 1.2) * W), W, H, 0.0, 300.0)
    cam.antiAliasingAmount = 1
    ns = int(rng.integers(0, 18 if reflective else 60))
    nt = int(rng.integers(0, 14 if reflective else 40))
    if ns + nt == 0:
        ns = 3
    n_pl = int(rng.integers(1, 6))
    nl = int(rng.integers(1, 5))
    nm = 6
    mats = np.zeros(nm, dtype=_abi.material_dt)
    for i in range(nm):
        spec = float(rng.choice([0.0, 0.02, 0.3, 0.8])) if reflective else 0.0
        mats[i] = scenes._material(tuple(rng.uniform(0.1, 1.0, 3)), float(rng.choice([0.5, 8.0, 64.0, 200.0])), spec, 0.0, 1.5)
    if reflective:
        mats[0]["specular"] = 0.5
        if seed % 3 == 0:  # every third reflective scene has glass: refraction, Fresnel, transparent occluders
            mats[1]["transparency"] = 0.9
            mats[1]["refractive_index"] = 1.5
            mats[2]["transparency"] = float(rng.choice([0.3, 0.6, 1.0]))
            mats[2]["refractive_index"] = float(rng.choice([1.1, 1.33, 2.0]))
    spheres = np.zeros(ns, dtype=_abi.sphere_dt)
    for i in range(ns):
        r = float(np.float32(rng.choice([0.3, 1.0, 2.5, 6.0, 40.0], p=[0.2, 0.35, 0.3, 0.12, 0.03])))
        c = rng.uniform(-18, 18, 3) + np.array([0, 0, 10.0])
        spheres[i] = (tuple(np.float32(c).astype(float)), r, int(rng.integers(0, nm)), 0)
    tris = np.zeros(nt, dtype=_abi.triangle_dt)
    for i in range(nt):
        c = rng.uniform(-15, 15, 3) + np.array([0, 0, 8.0])
        v = [tuple(np.float32(c + rng.uniform(-4, 4, 3)).astype(float)) for _ in range(3)]
        tris[i] = (v[0], v[1], v[2], (0.0, 0.0, 0.0), int(rng.integers(0, nm)), 0)
    planes = np.zeros(n_pl, dtype=_abi.plane_dt)
    base = [((0, -15, 0), (0, 1, 0)), ((0, 0, 30), (0, 0, -1)), ((-20, 0, 0), (1, 0, 0)), ((20, 0, 0), (-1, 0, 0)), ((0, 18, 0), (0, -1, 0))]
    for i in range(n_pl):
        p, n = base[i]
        n = np.array(n, float) + rng.uniform(-0.2, 0.2, 3) * (rng.random() < 0.5)
        planes[i] = (p, tuple(np.float32(n).astype(float)), int(rng.integers(0, nm)), 0)
    lights = np.zeros(nl, dtype=_abi.light_dt)
    for j in range(nl):
        lights[j] = (tuple(np.float32(rng.uniform(-14, 14, 3) + np.array([0, 0, 5.0])).astype(float)), tuple(rng.uniform(0.5, 1.0, 3)), float(rng.uniform(50, 400)))
    arrays = scenes.SceneArrays(materials=mats, spheres=spheres, planes=planes, triangles=tris, lights=lights)
    bias = float(rng.choice([1e-3, 1e-3, 1e-4, 5e-3]))
    return arrays, cam, bias


@pytest.mark.parametrize("seed", range(N_DIFFUSE))
def test_random_diffuse_scenes(engine, seed):
    arrays, cam, bias = random_scene(1000 + seed, reflective=False)
    engine.upload(arrays, cam)
    fast = engine.render_host(outputs=OUTS, tonemap_op=_abi.TM_ACES, bias=bias)
    launches = engine.last_render_launches()
    exact = engine.render_host(outputs=OUTS, tonemap_op=_abi.TM_ACES, bias=bias, flags=_abi.FLAG_EXACT_FP64)
    assert launches in (1, 4), launches  # the FP32 kernel (+ the grid build)
    assert np.array_equal(fast["prim_id"], exact["prim_id"]), f"seed {seed}: {np.count_nonzero(fast['prim_id'] != exact['prim_id'])} hit mismatches"
    assert np.array_equal(fast["light_mask"], exact["light_mask"]), f"seed {seed}: {np.count_nonzero(fast['light_mask'] != exact['light_mask'])} visibility mismatches"
    d = np.abs(fast["rgb8"].astype(np.int16) - exact["rgb8"].astype(np.int16))
    assert d.max() <= 1, f"seed {seed}: {np.count_nonzero(d > 1)} bytes off by more than 1"


@pytest.mark.parametrize("seed", range(N_REFLECTIVE))
def test_random_reflective_scenes(engine, seed):
    arrays, cam, bias = random_scene(2000 + seed, reflective=True)
    engine.upload(arrays, cam)
    fast = engine.render_host(outputs=OUTS, tonemap_op=_abi.TM_ACES, bias=bias)
    launches, sf = engine.last_render_launches(), engine.stats()
    exact = engine.render_host(outputs=OUTS, tonemap_op=_abi.TM_ACES, bias=bias, flags=_abi.FLAG_EXACT_FP64)
    assert launches == 2, launches  # render32r_kernel + the (empty) exact pass
    assert np.array_equal(fast["prim_id"], exact["prim_id"]), f"seed {seed}: {np.count_nonzero(fast['prim_id'] != exact['prim_id'])} hit mismatches"
    assert np.array_equal(fast["light_mask"], exact["light_mask"]), f"seed {seed}: {np.count_nonzero(fast['light_mask'] != exact['light_mask'])} visibility mismatches"
    d = np.abs(fast["rgb8"].astype(np.int16) - exact["rgb8"].astype(np.int16))
    assert d.max() <= 1, f"seed {seed}: {np.count_nonzero(d > 1)} bytes off by more than 1 (max {d.max()})"
    np.testing.assert_allclose(fast["hdr"][:, :3], exact["hdr"][:, :3], rtol=5e-4, atol=5e-4)
