"""GPU parity AT THE NAMED SIZES of BASELINE.json's configs: whole frames against the CPU checker
(the reference's own code where oracle/_ref was built), not scaled-down copies.

  C1  1000x1000    main()'s scene (5 specular planes + box model, 2 lights), antiAliasingAmount = 1
  C2  3840x2160    1024 spheres + 4 planes, 4 lights, ACES
  C3  3840x2160    64 spheres, 1 light, all seven operators
  C4  1920x1080    OBJ-mesh scene: blob (1280 triangles) and box (12 triangles), 1 light
  C5  7680x4320    4096 spheres, 8 lights: 16 row pairs spread over the frame (a whole frame is ~10^12 tests on the CPU)

Bars (north_star): hit (type, index) and shadow visibility bit-exact, depth / normals <= 1e-5, RGB8 +-1 LSB.  The exact
FP64 outputs (hdr64, depth64) are asked for more: identical bits on diffuse scenes, 1e-9 where pow() is on the path."""
import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes

pytestmark = pytest.mark.gpu


def lsb_diff(a, b):
    return np.abs(a.astype(np.int16) - b.astype(np.int16))


def sampled_aovs(engine, oracle, arrays, cam, rows):
    """prim_id / light_mask / depth of single rows against the checker's AOV pass."""
    for r0 in rows:
        g = engine.render_host(outputs=("prim_id", "light_mask", "depth64", "normal"), row0=r0, row1=r0 + 1)
        o = oracle.render(arrays, cam, row0=r0, row1=r0 + 1, want_aovs=True, hdr=False)
        assert np.array_equal(g["prim_id"], o["prim_id"]), r0
        assert np.array_equal(g["light_mask"], o["light_mask"]), r0
        hit = o["prim_id"] != _abi.PRIM_MISS
        assert np.array_equal(g["depth64"][hit], o["depth"][hit]), r0
        np.testing.assert_allclose(g["normal"][:, :3], o["normal"], rtol=0, atol=1e-5)


def test_c2_whole_frame(engine, oracle):
    """BASELINE.json configs[1], the bench workload: every pixel of the 3840x2160 frame."""
    arrays, cam = scenes.config2()
    engine.upload(arrays, cam)
    o = oracle.render(arrays, cam)  # RenderImage() of the whole frame
    want = oracle.tonemap_one(_abi.TM_ACES, o["hdr"])
    exact = engine.render_host(outputs=("hdr64", "rgb8"), tonemap_op=_abi.TM_ACES)
    assert np.array_equal(exact["hdr64"], o["hdr"]), f"{np.count_nonzero((exact['hdr64'] != o['hdr']).any(axis=1))} pixels differ"
    assert np.array_equal(exact["rgb8"], want)
    prod = engine.render_host(outputs=("hdr", "rgb8"), tonemap_op=_abi.TM_ACES)  # the production path of bench.py (no FP64 outputs)
    st = engine.stats()
    d = lsb_diff(prod["rgb8"], want)
    assert d.max() <= 1, f"{np.count_nonzero(d > 1)} bytes differ by more than 1 LSB"
    np.testing.assert_allclose(prod["hdr"][:, :3], o["hdr"], rtol=2e-5, atol=1e-5)
    assert st["primary_rays"] == cam.width * cam.height
    sampled_aovs(engine, oracle, arrays, cam, (0, 700, 1081, 1500, 2159))


def test_c3_whole_frame_all_operators(engine, oracle):
    """C3: the seven operators at 3840x2160 — fused in the render call (byte-exact) and as the stand-alone FP32 pass (+-1 LSB)."""
    arrays, cam = scenes.config3()
    engine.upload(arrays, cam)
    o = oracle.render(arrays, cam, want_aovs=True)
    want = oracle.tonemap_all(o["hdr"])
    g = engine.render_host(outputs=("hdr64", "hdr", "prim_id", "light_mask"))
    assert np.array_equal(g["hdr64"], o["hdr"])
    assert np.array_equal(g["prim_id"], o["prim_id"]) and np.array_equal(g["light_mask"], o["light_mask"])
    for op in range(_abi.TM_COUNT):
        f = engine.render_host(outputs=("rgb8", "hdr64"), tonemap_op=op)["rgb8"]
        assert np.array_equal(f, want[op]), _abi.TM_NAMES[op]
        p = engine.render_host(outputs=("rgb8",), tonemap_op=op)["rgb8"]  # production path
        assert lsb_diff(p, want[op]).max() <= 1, _abi.TM_NAMES[op]
    tm = engine.tonemap_host(g["hdr"])
    for op in range(_abi.TM_COUNT):
        assert lsb_diff(tm[op], want[op]).max() <= 1, _abi.TM_NAMES[op]


@pytest.mark.parametrize("mesh", ["blob", "box"])
def test_c4_whole_frame(engine, oracle, mesh):
    """C4 at 1920x1080: the mesh model (flattened to triangles in canonical order) + the five planes of main(), diffuse."""
    arrays, cam = scenes.config4(mesh)
    assert (cam.width, cam.height) == (1920, 1080)
    engine.upload(arrays, cam)
    o = oracle.render(arrays, cam, want_aovs=(mesh == "box"))
    g = engine.render_host(outputs=("hdr64", "prim_id", "light_mask", "depth64"))
    assert np.array_equal(g["hdr64"], o["hdr"]), f"{np.count_nonzero((g['hdr64'] != o['hdr']).any(axis=1))} pixels differ"
    tri = (g["prim_id"] >> 30) == _abi.HIT_TRIANGLE
    assert tri.any() and np.all((g["prim_id"][tri] & 0x3FFFFFFF) == 0)  # Model hits report the MODEL index (RE/Shape.h:276)
    if mesh == "box":
        assert np.array_equal(g["prim_id"], o["prim_id"]) and np.array_equal(g["light_mask"], o["light_mask"])
        hit = o["prim_id"] != _abi.PRIM_MISS
        assert np.array_equal(g["depth64"][hit], o["depth"][hit])
    else:
        sampled_aovs(engine, oracle, arrays, cam, (300, 540, 700))
    want = oracle.tonemap_one(_abi.TM_ACES, o["hdr"])
    p = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES)["rgb8"]
    assert lsb_diff(p, want).max() <= 1


def test_c1_whole_frame(engine, oracle):
    """C1: main()'s scene at 1000x1000 through the Python mirror: reflection chains to depth 10; pow() (Blinn-Phong) makes the
    HDR agree to 1e-9 instead of bit for bit (CUDA's and glibc's pow differ in the last ulp)."""
    arrays, cam = scenes.config1()
    assert (cam.width, cam.height) == (1000, 1000)
    engine.upload(arrays, cam)
    o = oracle.render(arrays, cam, want_aovs=True)
    g = engine.render_host(outputs=("hdr64", "prim_id", "light_mask", "depth64"))
    assert np.array_equal(g["prim_id"], o["prim_id"]) and np.array_equal(g["light_mask"], o["light_mask"])
    hit = o["prim_id"] != _abi.PRIM_MISS
    assert np.array_equal(g["depth64"][hit], o["depth"][hit])
    np.testing.assert_allclose(g["hdr64"], o["hdr"], rtol=1e-9, atol=1e-14)
    want = oracle.tonemap_all(o["hdr"])
    for op in range(_abi.TM_COUNT):  # the production path of this scene: the reflective FP32 kernel (rte_render32r.cu)
        p = engine.render_host(outputs=("rgb8",), tonemap_op=op)["rgb8"]
        assert lsb_diff(p, want[op]).max() <= 1, _abi.TM_NAMES[op]
    p = engine.render_host(outputs=("hdr", "prim_id", "light_mask"))
    assert engine.last_render_launches() == 2 and engine.stats()["fp64_confirms"] == 0  # that kernel, and no tile left to the exact one
    assert np.array_equal(p["prim_id"], o["prim_id"]) and np.array_equal(p["light_mask"], o["light_mask"])
    np.testing.assert_allclose(p["hdr"][:, :3], o["hdr"], rtol=2e-4, atol=2e-4)


def test_c5_8k_sampled_rows(engine, oracle):
    """C5 at its named size 7680x4320 (4096 spheres, 8 lights): 16 row pairs spread over the frame, bit for bit, through the
    production kernels of the full-height frame (row bands of the same camera)."""
    arrays, cam = scenes.config5()
    assert (cam.width, cam.height) == (7680, 4320)
    engine.upload(arrays, cam)
    W = cam.width
    for i in range(16):
        r0 = 3 + i * 287
        o = oracle.render(arrays, cam, row0=r0, row1=r0 + 2)
        want = oracle.tonemap_one(_abi.TM_REINHARD_EXT_LUM, o["hdr"])
        g = engine.render_host(outputs=("hdr64", "rgb8"), tonemap_op=_abi.TM_REINHARD_EXT_LUM, row0=r0, row1=r0 + 2)
        assert np.array_equal(g["hdr64"], o["hdr"]), r0
        assert np.array_equal(g["rgb8"], want), r0
        p = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_REINHARD_EXT_LUM, row0=r0, row1=r0 + 2)["rgb8"]
        assert lsb_diff(p, want).max() <= 1, r0
    # one band of 256 rows through the same call the bench makes: its first / last rows equal single-row renders
    band = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_REINHARD_EXT_LUM, row0=2000, row1=2256)["rgb8"]
    for r0 in (2000, 2255):
        one = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_REINHARD_EXT_LUM, row0=r0, row1=r0 + 1)["rgb8"]
        assert np.array_equal(band[(r0 - 2000) * W:(r0 - 1999) * W], one), r0
