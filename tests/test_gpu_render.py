"""GPU parity of the render path (through the C ABI) against the CPU oracle.

Bars (BASELINE.json north_star): nearest-hit (type,index) and shadow visibility bit-exact, depth and
normals <= 1e-5 relative, 8-bit pixels within +-1 LSB.  The CUDA path confirms every FP32-filter candidate
in FP64 with the reference's operation order, so the tests below ask for MORE than the bar: identical
FP64 depth / HDR for diffuse scenes, identical algorithmic counters.
"""
import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes

pytestmark = pytest.mark.gpu

COUNTERS = ("primary_rays", "shadow_rays", "secondary_rays", "sphere_tests", "plane_tests", "triangle_tests")
AOVS = ("hdr64", "hdr", "prim_id", "depth64", "depth", "normal", "albedo", "light_mask")


def render_both(engine, oracle, arrays, cam, row0=0, row1=0, **opts):
    engine.upload(arrays, cam)
    g = engine.render_host(outputs=AOVS, row0=row0, row1=row1, **opts)
    g["stats"] = engine.stats()
    o = oracle.render(arrays, cam, row0=row0, row1=row1, want_aovs=True, max_depth=opts.get("max_depth", 0))
    if oracle.kind == "reference" and o["secondary_rays"] == 0 and g["stats"]["secondary_rays"] > 0:
        # the reference-built checker cannot instrument the private TraceRay recursion: its counters cover
        # the primary hit only.  The plain-C port (bit-identical outputs, tests/test_oracle.py) counts all depths.
        from oracle import pyoracle
        c = pyoracle.load_port().render(arrays, cam, row0=row0, row1=row1, want_aovs=True, hdr=False, max_depth=opts.get("max_depth", 0))
        for k in COUNTERS:
            o[k] = c[k]
    return g, o


def assert_parity(g, o, exact_hdr=True, hdr_rtol=1e-12):
    assert np.array_equal(g["prim_id"], o["prim_id"]), f"{np.count_nonzero(g['prim_id'] != o['prim_id'])} nearest-hit mismatches"
    assert np.array_equal(g["light_mask"], o["light_mask"]), f"{np.count_nonzero(g['light_mask'] != o['light_mask'])} visibility mismatches"
    hit = o["prim_id"] != _abi.PRIM_MISS
    assert np.array_equal(g["depth64"][hit], o["depth"][hit]), "FP64 depth differs from the oracle"
    assert np.all(np.isinf(g["depth64"][~hit]))
    np.testing.assert_allclose(g["depth"][hit], o["depth"][hit], rtol=1e-5)   # the stated bar, FP32 AOV
    np.testing.assert_allclose(g["normal"][:, :3], o["normal"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(g["albedo"][:, :3], o["albedo"], rtol=1e-6, atol=1e-7)
    if exact_hdr:
        assert np.array_equal(g["hdr64"], o["hdr"]), f"max |dHDR| = {np.abs(g['hdr64'] - o['hdr']).max()}"
    else:
        np.testing.assert_allclose(g["hdr64"], o["hdr"], rtol=hdr_rtol, atol=1e-14)
    np.testing.assert_allclose(g["hdr"][:, :3], o["hdr"], rtol=1e-6, atol=1e-30)
    s = g["stats"]
    for k in COUNTERS:
        assert s[k] == o[k], f"counter {k}: gpu {s[k]} != oracle {o[k]}"


def test_config2_small(engine, oracle):
    """C2 family (1024 spheres + 4 planes, 4 lights) at 1/8 resolution."""
    arrays, cam = scenes.config2(scale=0.125)
    g, o = render_both(engine, oracle, arrays, cam)
    assert_parity(g, o)
    assert g["stats"]["fp64_confirms"] > 0


def test_config3_small(engine, oracle):
    arrays, cam = scenes.config3(scale=0.125)
    g, o = render_both(engine, oracle, arrays, cam)
    assert_parity(g, o)


def test_config5_small(engine, oracle):
    """C5 family: 4096 spheres (64 KB of filter data in shared memory), 8 lights; a band of rows."""
    arrays, cam = scenes.config5(scale=1 / 32)
    g, o = render_both(engine, oracle, arrays, cam, row0=40, row1=104)
    assert_parity(g, o)


def test_config4_mesh_diffuse(engine, oracle):
    """C4: 1280-triangle model + 5 planes, diffuse."""
    arrays, cam = scenes.config4(mesh="blob", scale=0.125)
    g, o = render_both(engine, oracle, arrays, cam)
    assert_parity(g, o)
    tri = (g["prim_id"] >> 30) == _abi.HIT_TRIANGLE
    assert tri.any() and np.all((g["prim_id"][tri] & 0x3FFFFFFF) == 0)  # Model hits report the MODEL index


def test_config4_box(engine, oracle):
    arrays, cam = scenes.config4(mesh="box", scale=0.125)
    g, o = render_both(engine, oracle, arrays, cam)
    assert_parity(g, o)


def test_config1_reflections(engine, oracle):
    """C1: main()'s scene (specular planes + box, 2 lights), reflection chain to depth 10.
    Blinn-Phong uses pow(): CUDA's and glibc's pow differ in the last ulp, so HDR is compared to 1e-9."""
    arrays, cam = scenes.config1(200, 200)
    g, o = render_both(engine, oracle, arrays, cam)
    assert_parity(g, o, exact_hdr=False, hdr_rtol=1e-9)
    assert g["stats"]["secondary_rays"] > 0


def test_config4_specular(engine, oracle):
    arrays, cam = scenes.config4(mesh="blob", scale=0.1, specular=True)
    g, o = render_both(engine, oracle, arrays, cam)
    assert_parity(g, o, exact_hdr=False, hdr_rtol=1e-9)


def test_mixed_scene_all_primitive_classes(engine, oracle):
    """Spheres, planes, free triangles and a mesh model in one diffuse scene: canonical order across the classes."""
    arrays, cam = scenes.mixed_scene(160, 90, 80.0)
    g, o = render_both(engine, oracle, arrays, cam)
    kinds = set((o["prim_id"][o["prim_id"] != 0xFFFFFFFF] >> 30).tolist())
    assert kinds == {1, 2, 3}, kinds  # RTE_HIT_SPHERE, _PLANE and _TRIANGLE all occur
    assert_parity(g, o)


def test_glass_scene_refraction_fresnel_transmittance(engine, oracle):
    """Row N1: transparent materials — refraction rays, Fresnel-weighted reflection, total internal reflection and the
    transmittance march through partially transparent occluders (RE/Scene.h:35-77,161-195), evaluated on the
    per-pixel stack in the reference's order.  libm's pow() (Fresnel, Blinn-Phong) is the only inexact step."""
    arrays, cam = scenes.glass_scene(96, 72)
    g, o = render_both(engine, oracle, arrays, cam)
    assert_parity(g, o, exact_hdr=False, hdr_rtol=1e-9)
    assert g["stats"]["secondary_rays"] > 2 * g["stats"]["primary_rays"]  # ray trees (refraction + reflection)
    assert g["stats"]["exact_marches"] > 0


def test_glass_scene_recursion_limits(engine, port_oracle):
    arrays, cam = scenes.glass_scene(48, 36)
    engine.upload(arrays, cam)
    for md in (1, 2, 5, _abi.MAX_DEPTH):
        g = engine.render_host(outputs=("hdr64",), max_depth=md)
        s = engine.stats()
        o = port_oracle.render(arrays, cam, max_depth=md)
        np.testing.assert_allclose(g["hdr64"], o["hdr"], rtol=1e-9, atol=1e-14)
        assert s["secondary_rays"] == o["secondary_rays"] and s["shadow_rays"] == o["shadow_rays"], md


@pytest.mark.parametrize("cfg", ["c3_fast", "c3_general", "c5_streamed", "c4_mesh", "glass"])
def test_multisample_antialiasing(engine, port_oracle, cfg):
    """Row N2: camera.antiAliasingAmount samples per pixel (sample 0 unjittered, the others jittered by the
    counter-based stream of rte.h) — the same samples, summed and divided like GeneratePixelAt, as the port oracle."""
    if cfg == "c5_streamed":
        arrays, cam = scenes.config5(scale=1 / 48)
    elif cfg == "c4_mesh":
        arrays, cam = scenes.config4("blob", scale=0.06)
    elif cfg == "glass":
        arrays, cam = scenes.glass_scene(40, 30)
    else:
        arrays, cam = scenes.config3(scale=0.05)
    cam.antiAliasingAmount, cam.jitterSeed = 5, 2026
    engine.upload(arrays, cam)
    outs = ("hdr64", "prim_id") if cfg in ("c3_general", "glass") else ("hdr64",)
    g = engine.render_host(outputs=outs)
    s = engine.stats()
    o = port_oracle.render(arrays, cam, want_aovs=True)
    if cfg == "glass":
        np.testing.assert_allclose(g["hdr64"], o["hdr"], rtol=1e-9, atol=1e-14)
    else:
        assert np.array_equal(g["hdr64"], o["hdr"]), f"max |dHDR| = {np.abs(g['hdr64'] - o['hdr']).max()}"
    if "prim_id" in g:
        assert np.array_equal(g["prim_id"], o["prim_id"])  # AOVs describe the unjittered sample
    assert s["primary_rays"] == 5 * cam.width * cam.height == o["primary_rays"]
    assert s["shadow_rays"] == o["shadow_rays"]
    cam.jitterSeed = 2027
    engine.upload(arrays, cam)
    assert not np.array_equal(engine.render_host(outputs=("hdr64",))["hdr64"], g["hdr64"])


def test_max_depth_and_bias_options(engine, port_oracle):
    """max_depth / bias are constants in the reference (RE/Scene.h:24,291): only the port checker takes them."""
    arrays, cam = scenes.config1(96, 96)
    engine.upload(arrays, cam)
    g = engine.render_host(outputs=("hdr64",), max_depth=3)
    o = port_oracle.render(arrays, cam, max_depth=3)
    np.testing.assert_allclose(g["hdr64"], o["hdr"], rtol=1e-9, atol=1e-14)
    g = engine.render_host(outputs=("hdr64", "light_mask"), bias=5e-3)
    o = port_oracle.render(arrays, cam, bias=5e-3, want_aovs=True)
    assert np.array_equal(g["light_mask"], o["light_mask"])
    np.testing.assert_allclose(g["hdr64"], o["hdr"], rtol=1e-9, atol=1e-14)


@pytest.mark.parametrize("flag", [_abi.FLAG_EXACT_SHADOWS, _abi.FLAG_NO_FILTER, _abi.FLAG_EXACT_SHADOWS | _abi.FLAG_NO_FILTER])
def test_validation_switches_are_identical(engine, flag):
    """The any-hit shadow query == the literal march, and the FP32 filters never reject a real hit:
    switching either off must not change a single bit."""
    for arrays, cam in (scenes.config2(scale=0.05), scenes.config4(mesh="blob", scale=0.08), scenes.config1(64, 64)):
        engine.upload(arrays, cam)
        a = engine.render_host(outputs=("hdr64", "prim_id", "depth64", "light_mask"))
        b = engine.render_host(outputs=("hdr64", "prim_id", "depth64", "light_mask"), flags=flag)
        for k in a:
            assert np.array_equal(a[k], b[k], equal_nan=True), k


def test_fused_tonemap_is_bit_exact(engine, oracle):
    """The FP64 epilogue (tonemap + toColor fused in the render pass) reproduces tonemapAll byte for byte."""
    arrays, cam = scenes.config2(scale=0.1)
    engine.upload(arrays, cam)
    ref = None
    for op in range(_abi.TM_COUNT):
        g = engine.render_host(outputs=("rgb8", "hdr64"), tonemap_op=op)
        if ref is None:
            ref = oracle.tonemap_all(g["hdr64"])
        assert np.array_equal(g["rgb8"], ref[op]), _abi.TM_NAMES[op]


def test_row_bands_tile_the_frame(engine):
    """Row-band sharding: bands rendered separately concatenate to the full frame."""
    arrays, cam = scenes.config3(scale=0.1)
    engine.upload(arrays, cam)
    full = engine.render_host(outputs=("hdr64", "prim_id"))
    H = cam.height
    cuts = [0, H // 3, H // 3 + 1, H - 5, H]
    parts = [engine.render_host(outputs=("hdr64", "prim_id"), row0=a, row1=b) for a, b in zip(cuts[:-1], cuts[1:])]
    for k in ("hdr64", "prim_id"):
        assert np.array_equal(np.concatenate([p[k] for p in parts]), full[k])


def test_empty_scene_is_sky(engine, oracle):
    from raytracingengine_b200.scene import Camera, SceneArrays
    cam = Camera((0.0, 0.0, -25.0), 50.0, 33, 17, 0.0, 200.0)
    cam.antiAliasingAmount = 1
    arrays = SceneArrays()
    g, o = render_both(engine, oracle, arrays, cam)
    assert np.all(g["prim_id"] == _abi.PRIM_MISS)
    assert_parity(g, o)


def test_errors(engine):
    from raytracingengine_b200.engine import RteError
    arrays, cam = scenes.config3(scale=0.05)
    engine.upload(arrays, cam)
    with pytest.raises(RteError):
        engine.render_host(outputs=("hdr",), row0=5, row1=cam.height + 1)
    with pytest.raises(RteError):
        engine.render_host(outputs=("rgb8",))  # rgb8 without an operator
    with pytest.raises(RteError):
        engine.render_host(outputs=("hdr",), max_depth=_abi.MAX_DEPTH + 1)
    bad = scenes.config3(scale=0.05)[0]
    bad.materials = bad.materials[:3]  # sphere materials out of range are fine (default Material), NULL arrays are not
    bad.lights = np.zeros(4097, dtype=_abi.light_dt)
    with pytest.raises(RteError):
        engine.upload(bad, cam)  # more lights than the documented capacity (4096; light_mask covers the first 16)
    engine.upload(arrays, cam)  # the ctx stays usable after argument errors
    assert engine.render_host(outputs=("prim_id",))["prim_id"].size == cam.width * cam.height
    # a frame the device cannot hold (2^38 pixels x 16 B of staging): cudaMalloc fails, the call reports RTE_ERR_LIMIT before it
    # touches the destination, and the ctx is NOT poisoned (recoverable CUDA errors are not sticky)
    import ctypes as C
    import dataclasses
    huge = dataclasses.replace(cam, width=1 << 19, height=1 << 19)
    engine.set_camera(huge)
    f = _abi.Frame()
    dummy = np.zeros(4, np.float32)
    f.hdr = dummy.ctypes.data
    rc = engine.lib.rte_render_host(engine.ctx, None, C.byref(f))
    assert rc == _abi.RTE_ERR_LIMIT, (rc, engine.lib.rte_last_error(engine.ctx))
    engine.set_camera(cam)
    assert engine.render_host(outputs=("prim_id",))["prim_id"].size == cam.width * cam.height


@pytest.mark.parametrize("cfg", ["c2", "c3", "c2_band", "c2_odd", "c5_streamed", "c5_streamed_band", "big_resident", "c4_blob", "c4_box", "mixed"])
def test_fast_kernel_is_bit_identical(engine, oracle, cfg):
    """Without AOV outputs diffuse sphere/plane scenes take the lean multi-pixel kernel (render_fast_kernel):
    same HDR bits, same fused RGB8, same algorithmic counters as the oracle and as the general kernel."""
    opts = {}
    if cfg == "c2":
        arrays, cam = scenes.config2(scale=0.125)
    elif cfg == "c3":
        arrays, cam = scenes.config3(scale=0.125)
    elif cfg == "c2_band":
        arrays, cam = scenes.config2(scale=0.1)
        opts = dict(row0=37, row1=150)
    elif cfg == "c5_streamed":  # 9 views x 64 KB do not fit shared memory: streamed through two TMA-filled buffers
        arrays, cam = scenes.config5(scale=1 / 32)
    elif cfg == "c5_streamed_band":  # fewer tiles than warps in the last batch, odd band
        arrays, cam = scenes.config5(scale=1 / 48)
        opts = dict(row0=11, row1=64)
    elif cfg == "c4_blob":  # triangles through the views of their bounding spheres (no spheres in the scene)
        arrays, cam = scenes.config4("blob", scale=0.1)
    elif cfg == "c4_box":
        arrays, cam = scenes.config4("box", scale=0.1)
    elif cfg == "mixed":  # spheres + planes + free triangles + a mesh model, all hit
        arrays, cam = scenes.mixed_scene(192, 108, 96.0)
    elif cfg == "big_resident":  # 2048 spheres x 5 views = 160 KB: resident, one 512-thread CTA per SM
        arrays, cam = scenes.synthetic_room(240, 135, 120.0, 2048, 4, 4321)
    else:  # width / height not multiples of the 8 x 8 warp tile
        arrays, cam = scenes.synthetic_room(203, 117, 100.0, 300, 3, 77)
    engine.upload(arrays, cam)
    fast = engine.render_host(outputs=("hdr64", "hdr", "rgb8"), tonemap_op=_abi.TM_ACES, **opts)
    sf = engine.stats()
    gen = engine.render_host(outputs=("hdr64", "hdr", "rgb8"), tonemap_op=_abi.TM_ACES, flags=_abi.FLAG_GENERAL_KERNEL, **opts)
    sg = engine.stats()
    o = oracle.render(arrays, cam, want_aovs=True, **opts)
    assert np.array_equal(fast["hdr64"], o["hdr"]), f"max |dHDR| = {np.abs(fast['hdr64'] - o['hdr']).max()}"
    for k in ("hdr64", "hdr", "rgb8"):
        assert np.array_equal(fast[k], gen[k]), k
    assert np.array_equal(fast["rgb8"], oracle.tonemap_all(o["hdr"])[_abi.TM_ACES])
    for k in COUNTERS:
        assert sf[k] == o[k] == sg[k], f"counter {k}: fast {sf[k]}, general {sg[k]}, oracle {o[k]}"
    assert sf["sphere_tests_executed"] > 0 or arrays.counts()["spheres"] == 0
    # one launch (primary rays + shading) and two launches (nearest hits through a device buffer, then shading): same bits.
    # Where the two-launch path does not apply (streamed views, triangles) RTE_FLAG_TWO_PASS is ignored.
    launches = {}
    for flag in (_abi.FLAG_SINGLE_PASS, _abi.FLAG_TWO_PASS):
        alt = engine.render_host(outputs=("hdr64", "hdr", "rgb8"), tonemap_op=_abi.TM_ACES, flags=flag, **opts)
        launches[flag] = engine.last_render_launches()  # (views are built already: render launch(es) + the RGB8 pass)
        sa = engine.stats()
        for k in ("hdr64", "hdr", "rgb8"):
            assert np.array_equal(fast[k], alt[k]), (flag, k)
        for k in COUNTERS:
            assert sa[k] == sf[k], (flag, k)
    two_pass_applies = cfg in ("c2", "c2_band", "c2_odd", "c3", "big_resident", "c5_streamed", "c5_streamed_band")  # sphere views, no triangles
    assert launches[_abi.FLAG_SINGLE_PASS] == 2
    assert launches[_abi.FLAG_TWO_PASS] == (3 if two_pass_applies else 2), launches


@pytest.mark.parametrize("cfg", ["c2", "c5_streamed", "c4_blob", "mixed"])
def test_tile_culling_changes_nothing(engine, cfg):
    """The lean kernel's per-tile culled mini-views (cone-cone test of every sphere against the warp's rays) against
    RTE_FLAG_NO_CULL (every ray scans the full view): identical bits and counters, fewer executed tests."""
    if cfg == "c2":
        arrays, cam = scenes.config2(scale=0.2)
    elif cfg == "c5_streamed":
        arrays, cam = scenes.config5(scale=1 / 24)
    elif cfg == "c4_blob":
        arrays, cam = scenes.config4("blob", scale=0.15)
    else:
        arrays, cam = scenes.mixed_scene(256, 144, 128.0, n_spheres=400)
    engine.upload(arrays, cam)
    a = engine.render_host(outputs=("hdr64", "rgb8"), tonemap_op=_abi.TM_ACES)
    sa = engine.stats()
    b = engine.render_host(outputs=("hdr64", "rgb8"), tonemap_op=_abi.TM_ACES, flags=_abi.FLAG_NO_CULL)
    sb = engine.stats()
    assert np.array_equal(a["hdr64"], b["hdr64"]) and np.array_equal(a["rgb8"], b["rgb8"])
    for k in COUNTERS:
        assert sa[k] == sb[k], k
    if arrays.counts()["spheres"] >= 128:  # (tiles of a small image span wide cones: the saving grows with the resolution)
        assert sa["sphere_tests_executed"] < sb["sphere_tests_executed"], (sa["sphere_tests_executed"], sb["sphere_tests_executed"])


@pytest.mark.parametrize("cfg", ["c2_stripe8", "c5_streamed_stripe8", "c5_streamed_stripe16", "c5_streamed_stripe4"])
def test_fast_kernel_stripes(engine, cfg):
    """Interleaved stripes through the fast kernel == the same rows of the full frame (streamed views: the CTA's batch of
    tiles stays inside one stripe, 16 x 1, 8 x 2 or 4 x 4 tiles)."""
    from raytracingengine_b200 import sharding
    arrays, cam = scenes.config2(scale=0.1) if cfg == "c2_stripe8" else scenes.config5(scale=1 / 24)
    engine.upload(arrays, cam)
    full = engine.render_host(outputs=("hdr64",))["hdr64"].reshape(cam.height, cam.width, 3)
    world, S = 3, int(cfg.split("stripe")[1])
    for r in range(world):
        o = sharding.stripe_opts(r, world, cam.height, S)
        rows = sharding.stripe_row_list(r, world, cam.height, S)
        for flag in (0, _abi.FLAG_TWO_PASS):  # (two launches: the primary launch of streamed views works on CTA batches inside a stripe)
            part = engine.render_host(outputs=("hdr64",), flags=flag, **o)["hdr64"]
            assert np.array_equal(part.reshape(len(rows), cam.width, 3), full[rows]), flag


@pytest.mark.parametrize("cfg", ["rows", "stripes", "aovs"])
def test_render_host_bands_equal_one_launch(engine, cfg):
    """rte_render_host renders calls of >= 1024 rows as two bands (device->host copy of the first under the kernels of the
    second): outputs and statistics must equal those of the same rows rendered band by band in single-launch calls."""
    arrays, cam = scenes.config2(scale=0.25)
    cam.height = 1288 if cfg != "stripes" else 2600   # >= 1024 rows per call; not a multiple of the band unit
    engine.upload(arrays, cam)
    outs = ("hdr64", "rgb8") if cfg != "aovs" else AOVS
    opts = dict(tonemap_op=_abi.TM_ACES) if cfg != "aovs" else {}
    if cfg == "stripes":
        from raytracingengine_b200 import sharding
        o = sharding.stripe_opts(1, 2, cam.height, 8)
        rows = sharding.stripe_row_list(1, 2, cam.height, 8)
        assert len(rows) >= 1024
        got = engine.render_host(outputs=outs, **o, **opts)
        st = engine.stats()
        assert st["primary_rays"] == len(rows) * cam.width
        for k in outs:  # the same rows out of the full frame, rendered 500 rows per (single-launch) call
            full = np.concatenate([engine.render_host(outputs=(k,), row0=a, row1=min(a + 500, cam.height), flags=_abi.FLAG_EXACT_FP64, **opts)[k]
                                       for a in range(0, cam.height, 500)])  # (the striped call asked for hdr64: the exact path on both sides)
            full = full.reshape(cam.height, cam.width, -1)
            assert np.array_equal(got[k].reshape(len(rows), cam.width, -1), full[rows]), k
        return
    got = engine.render_host(outputs=outs, **opts)
    st = engine.stats()
    want = {k: [] for k in outs}
    tot = {k: 0 for k in COUNTERS}
    for a in range(0, cam.height, 500):
        part = engine.render_host(outputs=outs, row0=a, row1=min(a + 500, cam.height), **opts)
        s1 = engine.stats()
        for k in outs:
            want[k].append(part[k])
        for k in COUNTERS:
            tot[k] += s1[k]
    for k in outs:
        assert np.array_equal(got[k], np.concatenate(want[k]), equal_nan=(got[k].dtype.kind == "f")), k
    for k in COUNTERS:
        assert st[k] == tot[k], k


@pytest.mark.parametrize("cfg", ["c2_fast", "c2_general", "c2_odd_width", "c4_general"])
def test_rgb8_peer_frames_equal_the_band_output(engine, cfg):
    """rte_set_rgb8_peers: every rank's stripes, stored by the render kernel straight into the registered full frames
    (here two buffers of the same GPU stand in for the peers), assemble the frame a single full render produces."""
    from raytracingengine_b200 import sharding
    if cfg == "c4_general":
        arrays, cam = scenes.config4("box", scale=0.1)
    else:
        arrays, cam = scenes.config2(scale=0.1)
    if cfg == "c2_odd_width":  # width not a multiple of 4: rows are not word aligned, tiles overhang the right edge
        arrays, cam = scenes.synthetic_room(203, 117, 100.0, 300, 3, 77)
    flags = _abi.FLAG_GENERAL_KERNEL if cfg == "c2_general" else 0
    engine.upload(arrays, cam)
    n = cam.width * cam.height
    full = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES, flags=flags)["rgb8"]
    frames = [engine.malloc(3 * n + 16) for _ in range(2)]
    band = engine.malloc(3 * n + 16)
    engine.h2d(frames[0], np.full(3 * n, 0xAB, np.uint8))
    engine.h2d(frames[1], np.full(3 * n, 0xCD, np.uint8))
    engine.set_rgb8_peers(frames)
    try:
        world, S = 3, 8
        for r in range(world):
            o = sharding.stripe_opts(r, world, cam.height, S)
            out = {"rgb8": band} if r == 1 else {}  # with and without the local band buffer
            if r != 1:
                out = {"hdr": engine.malloc(16 * n)}
            engine.render_device(out, tonemap_op=_abi.TM_ACES, flags=flags, **o)
            if r == 1:
                rows = sharding.stripe_row_list(r, world, cam.height, S)
                b = np.zeros((len(rows) * cam.width, 3), np.uint8)
                engine.d2h(b, band)
                assert np.array_equal(b.reshape(len(rows), cam.width, 3), full.reshape(cam.height, cam.width, 3)[rows])
        for f in frames:
            got = np.zeros((n, 3), np.uint8)
            engine.d2h(got, f)
            assert np.array_equal(got, full), f"{np.count_nonzero((got != full).any(axis=1))} pixels differ"
    finally:
        engine.set_rgb8_peers([])


@pytest.mark.parametrize("cfg", ["c2_4k", "c5_4k"])
def test_full_size_properties(engine, oracle, cfg):
    """BASELINE.json configs[1] at its full size (3840x2160, 1024 spheres, 4 lights) and the C5 scene (4096 spheres, 8
    lights, streamed views) at 3840x2160: the production path (two launches, tile / CTA culling, two-band host copy)
    against size-independent properties — the frame equals its row bands rendered by the single-launch kernel (one of
    them without culling), the counters add up, and sampled rows equal the CPU oracle bit for bit."""
    arrays, cam = scenes.config2() if cfg == "c2_4k" else scenes.config5(0.5)
    engine.upload(arrays, cam)
    full = engine.render_host(outputs=("hdr64", "rgb8"), tonemap_op=_abi.TM_ACES)
    sf = engine.stats()
    assert engine.last_render_launches() >= 3  # primary-ray launch + shading launch + RGB8 pass (per band of the host copy)
    assert sf["primary_rays"] == cam.width * cam.height
    H, W = cam.height, cam.width
    tot = {k: 0 for k in COUNTERS}
    for a in range(0, H, 540):
        flags = _abi.FLAG_SINGLE_PASS | (_abi.FLAG_NO_CULL if a == 540 else 0)
        part = engine.render_host(outputs=("hdr64", "rgb8"), tonemap_op=_abi.TM_ACES, row0=a, row1=a + 540, flags=flags)
        sp = engine.stats()
        for k in COUNTERS:
            tot[k] += sp[k]
        for k in ("hdr64", "rgb8"):
            assert np.array_equal(full[k][a * W:(a + 540) * W], part[k]), (a, k)
    for k in COUNTERS:
        assert sf[k] == tot[k], k
    for r0 in (0, 1079, 2158):  # top, middle, bottom rows against the oracle
        o = oracle.render(arrays, cam, row0=r0, row1=r0 + 2, want_aovs=False)
        assert np.array_equal(full["hdr64"][r0 * W:(r0 + 2) * W], o["hdr"]), r0
        assert np.array_equal(full["rgb8"][r0 * W:(r0 + 2) * W], oracle.tonemap_all(o["hdr"])[_abi.TM_ACES]), r0


def test_frame_fence_flags(engine):
    """rte_set_fence_peers / rte_frame_fence: the kernel that does a render's peer stores ends by writing the render's
    sequence number into every rank's flag array; the fence waits (device side) for all ranks and times out when one is
    missing.  One process, one GPU: "rank 1" is a second flag array whose flag this test writes itself."""
    from raytracingengine_b200.engine import RteError
    arrays, cam = scenes.config2(scale=0.05)
    engine.upload(arrays, cam)
    n = cam.width * cam.height
    frame = engine.malloc(n * 3)
    f0, f1 = engine.malloc(4 * _abi.MAX_PEERS), engine.malloc(4 * _abi.MAX_PEERS)
    engine.h2d(f1, np.zeros(_abi.MAX_PEERS, np.uint32))
    engine.set_rgb8_peers([frame])
    try:
        # one rank: its own flag arrives with the frame
        engine.set_fence_peers(0, [f0])
        for seq in (1, 2):
            engine.render_device({}, tonemap_op=_abi.TM_ACES)
            engine.frame_fence()
            engine.synchronize()
            got = np.zeros(_abi.MAX_PEERS, np.uint32)
            engine.d2h(got, f0)
            assert got[0] == seq and not got[1:].any()
        ref = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES)["rgb8"]
        out = np.zeros((n, 3), np.uint8)
        engine.d2h(out, frame)
        assert np.array_equal(out, ref)
        # two ranks, rank 1 silent: both arrays get rank 0's flag, the fence times out
        engine.set_fence_peers(0, [f0, f1])
        engine.render_device({}, tonemap_op=_abi.TM_ACES)
        engine.frame_fence(50.0)
        with pytest.raises(RteError, match="did not deliver"):
            engine.synchronize()
        got0, got1 = np.zeros(_abi.MAX_PEERS, np.uint32), np.zeros(_abi.MAX_PEERS, np.uint32)
        engine.d2h(got0, f0); engine.d2h(got1, f1)
        assert got0[0] == 1 and got1[0] == 1 and got0[1] == 0
        # "rank 1" delivers frame 2 before this rank asks: no wait, no error; an empty band still signals
        two = np.zeros(_abi.MAX_PEERS, np.uint32); two[0] = 1; two[1] = 2
        engine.h2d(f0, two)
        engine.render_device({}, tonemap_op=_abi.TM_ACES, row0=cam.height, row1=cam.height)
        engine.frame_fence(50.0)
        engine.synchronize()
        engine.d2h(got0, f0)
        assert got0[0] == 2 and got0[1] == 2
    finally:
        engine.set_fence_peers(0, [])
        engine.set_rgb8_peers([])


def test_ipc_export_open_roundtrip(engine):
    """rte_ipc_export / rte_ipc_open / rte_ipc_close: a second process maps this process' buffer and stores into it."""
    import os, subprocess, sys, textwrap
    n = 1 << 16
    dev = engine.malloc(n)
    engine.h2d(dev, np.zeros(n, np.uint8))
    handle = engine.ipc_export(dev)
    assert len(handle) == _abi.IPC_HANDLE_BYTES
    child = textwrap.dedent(f"""
        import sys, numpy as np
        sys.path.insert(0, {os.path.dirname(os.path.dirname(os.path.abspath(__file__)))!r})
        from raytracingengine_b200.engine import Engine
        with Engine(0) as e:
            p = e.ipc_open(bytes.fromhex({handle.hex()!r}))
            e.h2d(p, (np.arange({n}) % 251).astype(np.uint8))
            e.synchronize()
            e.ipc_close(p)
    """)
    r = subprocess.run([sys.executable, "-c", child], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    got = np.zeros(n, np.uint8)
    engine.d2h(got, dev)
    assert np.array_equal(got, (np.arange(n) % 251).astype(np.uint8))
