"""GPU parity of the FP32 PRODUCTION path (csrc/rte_render32.cu incl. its exact fix-up batches): calls that ask for hdr / rgb8 /
prim_id / light_mask only, on diffuse sphere / plane scenes.

Bars (BASELINE.json north_star): nearest-hit (type, index) and shadow visibility bit-exact, RGB8 within +-1 LSB of the
reference; here also: HDR to 2e-5 relative, ray / test counters identical to the exact path's (which equal the oracle's)."""
import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes

pytestmark = pytest.mark.gpu
COUNTERS = ("primary_rays", "shadow_rays", "sphere_tests", "plane_tests", "triangle_tests")
OUTS = ("hdr", "rgb8", "prim_id", "light_mask")


def both_paths(engine, op=_abi.TM_ACES, outs=OUTS, **opts):
    fast = engine.render_host(outputs=outs, tonemap_op=op, **opts)
    sf, lf = engine.stats(), engine.last_render_launches()
    exact = engine.render_host(outputs=outs, tonemap_op=op, flags=_abi.FLAG_EXACT_FP64, **opts)
    se = engine.stats()
    return fast, sf, lf, exact, se


def check(fast, sf, exact, se, n_px):
    assert np.array_equal(fast["prim_id"], exact["prim_id"]), f"{np.count_nonzero(fast['prim_id'] != exact['prim_id'])} nearest-hit mismatches"
    assert np.array_equal(fast["light_mask"], exact["light_mask"]), f"{np.count_nonzero(fast['light_mask'] != exact['light_mask'])} visibility mismatches"
    d = np.abs(fast["rgb8"].astype(np.int16) - exact["rgb8"].astype(np.int16))
    assert d.max() <= 1, f"{np.count_nonzero(d > 1)} bytes differ by more than 1 LSB"
    np.testing.assert_allclose(fast["hdr"][:, :3], exact["hdr"][:, :3], rtol=2e-5, atol=1e-5)
    for k in COUNTERS:
        assert sf[k] == se[k], f"counter {k}: fp32 path {sf[k]}, exact path {se[k]}"
    assert sf["fp64_confirms"] <= 0.02 * n_px, f"{sf['fp64_confirms']} of {n_px} pixels went through the exact fix-up"


@pytest.mark.parametrize("cfg", ["c2", "c2_quarter", "c3", "odd", "big_resident", "camera_inside", "c4_blob", "c4_box", "mixed", "mesh_10k"])
def test_fp32_path_against_exact_path_and_oracle(engine, oracle, cfg):
    if cfg == "c2":
        arrays, cam = scenes.config2(scale=0.125)
    elif cfg == "c2_quarter":
        arrays, cam = scenes.config2(scale=0.25)
    elif cfg == "c3":
        arrays, cam = scenes.config3(scale=0.125)
    elif cfg == "big_resident":  # 2048 spheres x 5 views = 160 KB resident, no room for the cull scratch
        arrays, cam = scenes.synthetic_room(240, 135, 120.0, 2048, 4, 4321)
    elif cfg == "c4_blob":  # diffuse OBJ-mesh scenes: triangles through the same grids (tri_interval)
        arrays, cam = scenes.config4("blob", scale=0.25)
    elif cfg == "c4_box":
        arrays, cam = scenes.config4("box", scale=0.2)
    elif cfg == "mixed":  # spheres, planes, free triangles and a model in one scene
        arrays, cam = scenes.mixed_scene(224, 126, 130.0)
    elif cfg == "mesh_10k":  # a 5 120-triangle model
        arrays, cam = scenes.mixed_scene(200, 112, 120.0, n_spheres=40, mesh_level=4)
    elif cfg == "camera_inside":  # the camera sits inside a big sphere: far roots, inward normals
        arrays, cam = scenes.synthetic_room(160, 90, 80.0, 200, 3, 99)
        arrays.spheres[7] = ((0.0, 0.0, -25.0), 60.0, 7, 0)
    else:  # width / height not multiples of the tile, width not a multiple of 4 (byte stores)
        arrays, cam = scenes.synthetic_room(203, 117, 100.0, 300, 3, 77)
    engine.upload(arrays, cam)
    fast, sf, lf, exact, se = both_paths(engine)
    assert lf in (1, 4), lf  # (3 grid-build kernels +) the FP32 kernel
    n = cam.width * cam.height
    check(fast, sf, exact, se, n)
    o = oracle.render(arrays, cam, want_aovs=True)
    assert np.array_equal(fast["prim_id"], o["prim_id"]) and np.array_equal(fast["light_mask"], o["light_mask"])
    want = oracle.tonemap_one(_abi.TM_ACES, o["hdr"])
    assert np.abs(fast["rgb8"].astype(np.int16) - want.astype(np.int16)).max() <= 1
    assert sf["primary_rays"] == o["primary_rays"] and sf["shadow_rays"] == o["shadow_rays"]
    assert sf["sphere_tests"] == o["sphere_tests"] and sf["plane_tests"] == o["plane_tests"] and sf["triangle_tests"] == o["triangle_tests"]


def test_fp32_path_all_operators_bands_stripes(engine, oracle):
    from raytracingengine_b200 import sharding
    arrays, cam = scenes.config3(scale=0.1)
    engine.upload(arrays, cam)
    o = oracle.render(arrays, cam)
    want = oracle.tonemap_all(o["hdr"])
    full = {}
    for op in range(_abi.TM_COUNT):
        g = engine.render_host(outputs=("rgb8",), tonemap_op=op)
        assert engine.last_render_launches() in (1, 4)  # (+ the grid build after an upload)
        assert np.abs(g["rgb8"].astype(np.int16) - want[op].astype(np.int16)).max() <= 1, _abi.TM_NAMES[op]
        full[op] = g["rgb8"]
    H, W = cam.height, cam.width
    ref = engine.render_host(outputs=("hdr", "rgb8"), tonemap_op=_abi.TM_ACES)
    cuts = [0, H // 3, H // 3 + 1, H - 5, H]
    parts = [engine.render_host(outputs=("hdr", "rgb8"), tonemap_op=_abi.TM_ACES, row0=a, row1=b) for a, b in zip(cuts[:-1], cuts[1:])]
    for k in ("hdr", "rgb8"):
        assert np.array_equal(np.concatenate([p[k] for p in parts]), ref[k]), k
    for r in range(3):
        so = sharding.stripe_opts(r, 3, H, 8)
        rows = sharding.stripe_row_list(r, 3, H, 8)
        part = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES, **so)["rgb8"]
        assert np.array_equal(part.reshape(len(rows), W, 3), ref["rgb8"].reshape(H, W, 3)[rows])


def test_fp32_path_multisample(engine, port_oracle):
    arrays, cam = scenes.config3(scale=0.05)
    cam.antiAliasingAmount, cam.jitterSeed = 4, 2026
    engine.upload(arrays, cam)
    g = engine.render_host(outputs=("hdr", "rgb8"), tonemap_op=_abi.TM_ACES)
    s = engine.stats()
    o = port_oracle.render(arrays, cam)
    np.testing.assert_allclose(g["hdr"][:, :3], o["hdr"], rtol=2e-5, atol=1e-5)
    assert s["primary_rays"] == 4 * cam.width * cam.height == o["primary_rays"] and s["shadow_rays"] == o["shadow_rays"]
    want = port_oracle.tonemap_one(_abi.TM_ACES, o["hdr"])
    assert np.abs(g["rgb8"].astype(np.int16) - want.astype(np.int16)).max() <= 1


def test_exact_flag_and_fp64_outputs_keep_the_exact_path(engine, oracle):
    arrays, cam = scenes.config2(scale=0.05)
    engine.upload(arrays, cam)
    o = oracle.render(arrays, cam)
    want = oracle.tonemap_one(_abi.TM_ACES, o["hdr"])
    a = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES, flags=_abi.FLAG_EXACT_FP64)["rgb8"]
    b = engine.render_host(outputs=("rgb8", "hdr64"), tonemap_op=_abi.TM_ACES)
    assert np.array_equal(a, want) and np.array_equal(b["rgb8"], want) and np.array_equal(b["hdr64"], o["hdr"])


def test_fp32_path_on_grid_aligned_edge_cases(engine, oracle):
    """Rays that pass EXACTLY through triangle vertices and edges, graze spheres tangentially, run parallel to a plane and look
    along the line where two planes meet (integer scene coordinates against integer pixel offsets: camera at the origin,
    focal = 64, so the point (x, y, 64) is the centre of pixel (x + W/2, H/2 - y)).  The intervals cannot decide those pixels;
    the exact batches must, like the reference: hit ids, visibility and bytes against the reference build."""
    W, H, F = 96, 64, 64.0
    cam = scenes.Camera((0.0, 0.0, 0.0), F, W, H, 0.0, 200.0)
    cam.antiAliasingAmount = 1
    mats = np.array([scenes._material(c) for c in ((1, 1, 1), (1, 0.3, 0.2), (0.2, 0.4, 1), (0.3, 1, 0.3), (1, 1, 0.2))], dtype=_abi.material_dt)
    planes = np.zeros(3, dtype=_abi.plane_dt)
    planes[0] = ((0, -20, 0), (0, 1, 0), 0, 0)        # floor: meets the back wall along a pixel row (y / z = -20 / 128 = -10 / 64)
    planes[1] = ((0, 0, 128), (0, 0, -1), 0, 0)       # back wall
    planes[2] = ((24, 0, 0), (-1, 0, 0), 3, 0)        # right wall: column x = 24 / 128 * 64 = 12 looks along its edge with the back wall
    tris = np.zeros(3, dtype=_abi.triangle_dt)
    tris[0] = ((-16, -8, 64), (0, 8, 64), (-16, 8, 64), (0, 0, 0), 1, 0)    # vertices on pixel centres, one edge along a pixel diagonal
    tris[1] = ((-16, -8, 64), (0, -8, 64), (0, 8, 64), (0, 0, 0), 2, 0)     # shares that diagonal (the first triangle wins the tie)
    tris[2] = ((4, -4, 96), (20, -4, 96), (4, 12, 96), (0, 0, 0), 4, 0)     # back-facing from the camera? no: counter-clockwise, seen from behind
    spheres = np.zeros(2, dtype=_abi.sphere_dt)
    spheres[0] = ((10.0, -14.0, 70.0), 6.0, 1, 0)     # rests on... above the floor, its shadow falls on the floor
    spheres[1] = ((-30.0, 10.0, 100.0), 10.0, 2, 0)
    lights = np.zeros(2, dtype=_abi.light_dt)
    lights[0] = ((0.0, 40.0, 64.0), (1, 1, 1), 4000.0)
    lights[1] = ((-16.0, 8.0, 20.0), (1, 0.9, 0.8), 1500.0)   # on the line camera -> vertex (-16, 8, 64): shadow rays along triangle edges
    arrays = scenes.SceneArrays(materials=mats, spheres=spheres, planes=planes, triangles=tris, lights=lights)
    engine.upload(arrays, cam)
    fast, sf, lf, exact, se = both_paths(engine)
    o = oracle.render(arrays, cam, want_aovs=True)
    assert np.array_equal(exact["prim_id"], o["prim_id"]) and np.array_equal(exact["light_mask"], o["light_mask"])
    assert np.array_equal(fast["prim_id"], o["prim_id"]), f"{np.count_nonzero(fast['prim_id'] != o['prim_id'])} nearest-hit mismatches"
    assert np.array_equal(fast["light_mask"], o["light_mask"]), f"{np.count_nonzero(fast['light_mask'] != o['light_mask'])} visibility mismatches"
    want = oracle.tonemap_one(_abi.TM_ACES, o["hdr"])
    assert np.abs(fast["rgb8"].astype(np.int16) - want.astype(np.int16)).max() <= 1
    for k in COUNTERS:
        assert sf[k] == se[k], k
    assert sf["fp64_confirms"] >= W  # at least the row along the floor / wall edge went through the exact batches
    kinds = set((o["prim_id"][o["prim_id"] != _abi.PRIM_MISS] >> 30).tolist())
    assert kinds == {_abi.HIT_SPHERE, _abi.HIT_PLANE, _abi.HIT_TRIANGLE}
