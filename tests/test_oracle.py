"""CPU suite: pins the oracle.  The plain-C port (oracle/oracle_port.c) must reproduce, bit for bit,
(1) the golden fixtures generated from the reference's own headers (tests/golden/make_golden.py),
(2) the known-answer vectors of SURVEY.md §8(c), and (3) — where oracle/_ref was built (this container) —
the reference-built checker itself on fresh scenes."""
import json
import os

import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.scene import Camera, SceneArrays

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
from tests.golden.make_golden import CASES  # noqa: E402

AOVS = ("hdr", "prim_id", "depth", "normal", "albedo", "light_mask")


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(GOLD, "kat.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("name", sorted(CASES))
def test_port_reproduces_reference_goldens(port_oracle, name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    arrays, cam = CASES[name]()
    assert (cam.width, cam.height) == (int(g["width"]), int(g["height"]))
    r = port_oracle.render(arrays, cam, want_aovs=True)
    for k in AOVS:
        assert np.array_equal(r[k], g[k], equal_nan=True), f"{name}: {k} differs from the reference golden"
    assert r["primary_rays"] == int(g["primary_rays"])
    rgb = port_oracle.tonemap_all(r["hdr"])
    for op in range(_abi.TM_COUNT):
        assert np.array_equal(rgb[op], g["rgb8"][op]), f"{name}: {_abi.TM_NAMES[op]}"


def test_camera_kat(port_oracle, kat):
    """Camera::getRay KATs, RE/Math.h:99-121 (default camera of main(), RE/RaytracingEngine.cpp:223-224)."""
    cam = Camera((0.0, 0.0, -25.0), 500.0, 1000, 1000, 0.0, 200.0)
    for key, want in kat["camera"].items():
        x, y = map(int, key.split(","))
        o, d = port_oracle.camera_ray(cam, x, y)
        assert d.tolist() == want and o.tolist() == [0.0, 0.0, -25.0]
    # hand-derivable: the centre pixel looks along +Z, the corner is (-1,1,1)/sqrt(3)
    assert port_oracle.camera_ray(cam, 500, 500)[1].tolist() == [0.0, 0.0, 1.0]
    np.testing.assert_allclose(port_oracle.camera_ray(cam, 0, 0)[1], np.array([-1, 1, 1]) / np.sqrt(3), rtol=1e-15)


def _kat_scenes():
    mats = np.zeros(1, dtype=_abi.material_dt)
    mats[0] = ((1, 1, 1), 128.0, 0.0, 0.0, 1.0)
    sph = np.zeros(1, dtype=_abi.sphere_dt); sph[0] = ((0, 0, 0), 1.0, 0, 0)
    pln = np.zeros(1, dtype=_abi.plane_dt); pln[0] = ((0, -15, 0), (0, 1, 0), 0, 0)
    tri = np.zeros(1, dtype=_abi.triangle_dt); tri[0] = ((-1, -1, 0), (1, -1, 0), (0, 1, 0), (0, 0, 0), 0, 0)
    return {"sphere": SceneArrays(materials=mats, spheres=sph), "plane": SceneArrays(materials=mats, planes=pln),
            "triangle": SceneArrays(materials=mats, triangles=tri)}


@pytest.mark.parametrize("shape", ["sphere", "plane", "triangle"])
def test_shape_kat(port_oracle, kat, shape):
    """Sphere/Plane/Triangle::Intersect KATs (RE/Shape.h:72-98,149-159,202-220): far root from inside, tangent
    hit, t == 0 on the plane, back-face and through-vertex triangle hits, misses."""
    k = kat["shapes"][shape]
    h = port_oracle.intersect_closest(_kat_scenes()[shape], k["origins"], k["directions"])
    assert h["prim_id"].tolist() == k["prim_id"]
    for got, want, pid in zip(h["distance"], k["distance"], k["prim_id"]):
        if pid != _abi.PRIM_MISS:
            assert got == want
    hit = np.array(k["prim_id"]) != _abi.PRIM_MISS
    assert np.array_equal(h["normal"][hit], np.array(k["normal"])[hit])


def test_shape_kat_hand_values(port_oracle):
    s = _kat_scenes()
    h = port_oracle.intersect_closest(s["sphere"], [(0, 0, -5), (0, 0, 0), (0, 0, -1), (1, 0, -5)], [(0, 0, 1), (0, 0, 1), (0, 0, -1), (0, 0, 1)])
    assert h["distance"][:2].tolist() == [4.0, 1.0] and h["prim_id"][2] == _abi.PRIM_MISS and h["distance"][3] == 5.0
    h = port_oracle.intersect_closest(s["triangle"], [(0, 0, -3), (0, 0, 3)], [(0, 0, 1), (0, 0, -1)])
    assert h["distance"].tolist() == [3.0, 3.0] and h["normal"][0].tolist() == [0.0, 0.0, 1.0]


def test_tonemap_kat(port_oracle, kat):
    """tonemapAll + toColor KATs (RE/RaytracingEngine.cpp:113-121,176-214); NaN / negative -> 0, >= 1 -> 255."""
    hdr = np.array([[np.nan if v is None else v for v in row] for row in kat["tonemap"]["hdr"]], dtype=np.float64)
    rgb = port_oracle.tonemap_all(hdr)
    for op in range(_abi.TM_COUNT):
        assert rgb[op].tolist() == kat["tonemap"]["rgb8"][op], _abi.TM_NAMES[op]
    # SURVEY §8(c) table rows (simple, reinhard, ext, ext-lum, jodie, uncharted2, aces)
    assert [rgb[op][1, 0] for op in range(7)] == [45, 38, 39, 39, 10, 32, 35]
    assert [rgb[op][2, 0] for op in range(7)] == [255, 127, 132, 132, 26, 125, 171]
    assert [tuple(rgb[op][3]) for op in range(7)] == [(127, 255, 255), (85, 127, 170), (86, 132, 183), (67, 134, 255), (13, 27, 54), (77, 125, 181), (111, 171, 214)]
    for op in range(_abi.TM_COUNT):
        assert np.array_equal(port_oracle.tonemap_one(op, hdr), rgb[op])


def test_ppm_bytes(port_oracle, tmp_path):
    """writePPM, RE/Image.cpp:11-31: 'P6\\n<W> <H>\\n255\\n' + RGB; 1000x1000 -> 3 000 017 bytes like RE/output.ppm."""
    rgb = (np.arange(5 * 3 * 3) % 256).astype(np.uint8)
    p = tmp_path / "a.ppm"
    assert port_oracle.write_ppm(str(p), rgb, 5, 3) == 0
    assert p.read_bytes() == b"P6\n5 3\n255\n" + rgb.tobytes()
    big = np.zeros(1000 * 1000 * 3, dtype=np.uint8)
    port_oracle.write_ppm(str(p), big, 1000, 1000)
    assert p.stat().st_size == 3000017


def test_port_equals_reference_build(port_oracle, ref_oracle):
    """Fresh scenes (not in the goldens): both checkers bit-identical, incl. any-hit queries."""
    for arrays, cam in (scenes.config2(scale=0.03), scenes.config1(40, 40), scenes.config4("blob", 0.04, specular=True)):
        a = ref_oracle.render(arrays, cam, want_aovs=True)
        b = port_oracle.render(arrays, cam, want_aovs=True)
        for k in AOVS:
            assert np.array_equal(a[k], b[k], equal_nan=True), k
    arrays, cam = scenes.config2(scale=0.03)
    rng = np.random.default_rng(7)
    o = rng.uniform(-14, 14, (500, 3)); d = rng.normal(size=(500, 3)); md = rng.uniform(0.5, 40, 500)
    ha, hb = ref_oracle.intersect_closest(arrays, o, d), port_oracle.intersect_closest(arrays, o, d)
    for f in ("prim_id", "distance", "normal", "point"):
        assert np.array_equal(ha[f], hb[f], equal_nan=True)
    assert np.array_equal(ref_oracle.intersect_any_before(arrays, o, d, md), port_oracle.intersect_any_before(arrays, o, d, md))


def test_row_subsets_and_threads(port_oracle):
    """The bounded CPU-baseline sample (every S-th row) is a subset of the full frame; thread count is irrelevant."""
    arrays, cam = scenes.config3(scale=0.03)
    full = port_oracle.render(arrays, cam)["hdr"].reshape(cam.height, cam.width, 3)
    sub = port_oracle.render(arrays, cam, row0=3, row1=cam.height - 2, row_step=5, threads=1)
    rows = list(range(3, cam.height - 2, 5))
    assert sub["rows"] == len(rows)
    assert np.array_equal(sub["hdr"].reshape(len(rows), cam.width, 3), full[rows])


def test_shadow_ray_count_definition(port_oracle):
    """rays = W*H + sum over hit pixels of #{lights: d > 0, N.L > 0, d > bias} == popcount of the cast bits."""
    arrays, cam = scenes.config2(scale=0.03)
    r = port_oracle.render(arrays, cam, want_aovs=True)
    cast = (r["light_mask"] >> 16) & 0xFFFF
    assert int(np.bitwise_count(cast).sum()) == r["shadow_rays"]
    assert r["primary_rays"] == cam.width * cam.height and r["secondary_rays"] == 0


def _jitter01(seed, pixel, s, axis):
    """The jitter stream include/rte.h defines for rte_camera, restated in Python integers."""
    M = (1 << 64) - 1
    z = ((seed & 0xFFFFFFFF) << 32) ^ ((pixel * 8192 + 2 * s + axis) & M)
    z = (z + 0x9E3779B97F4A7C15) & M
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
    z ^= z >> 31
    return (z >> 11) * 2.0 ** -53


def test_multisample_jitter_stream(port_oracle):
    """Row N2: anti_aliasing = n renders the mean of n samples, sample 0 unjittered, samples 1.. jittered by the
    counter-based stream of rte.h.  A sky-only scene makes the pixel a closed form of the ray directions."""
    arrays = scenes.SceneArrays()
    cam = scenes.Camera((0.0, 0.0, -25.0), 40.0, 8, 6, 0.0, 200.0)
    cam.antiAliasingAmount, cam.jitterSeed = 3, 77
    got = port_oracle.render(arrays, cam)
    assert got["primary_rays"] == 3 * 8 * 6
    want = np.zeros((6, 8, 3))
    for y in range(6):
        for x in range(8):
            acc = np.zeros(3)
            for s in range(3):
                jx = _jitter01(77, y * 8 + x, s, 0) if s else 0.0
                jy = _jitter01(77, y * 8 + x, s, 1) if s else 0.0
                d = np.array([x - 4.0 + jx, 3.0 - y + jy, 40.0])
                t = 0.5 * (d[1] / np.sqrt(d @ d) + 1.0)
                acc = acc + (np.array([1.0, 1.0, 1.0]) * (1.0 - t) + np.array([0.5, 0.7, 1.0]) * t)
            want[y, x] = acc / 3.0
    np.testing.assert_allclose(got["hdr"].reshape(6, 8, 3), want, rtol=1e-14)
    cam.jitterSeed = 78
    assert not np.array_equal(port_oracle.render(arrays, cam)["hdr"], got["hdr"])
    cam.antiAliasingAmount = 1
    one = port_oracle.render(arrays, cam)["hdr"]
    cam.jitterSeed = 77
    assert np.array_equal(port_oracle.render(arrays, cam)["hdr"], one)  # a single sample never reads the stream


def test_multisample_is_statistically_the_reference(port_oracle, ref_oracle):
    """The reference's own 16-sample render (mt19937 seeded by random_device: different jitter on every run) and the
    port's 16-sample render with the counter-based stream estimate the same image: their difference is sampling
    noise — far smaller than the difference to the 1-sample image, and unbiased."""
    arrays, cam = scenes.config3(scale=0.03)
    cam.antiAliasingAmount = 16
    a = ref_oracle.render(arrays, cam)["hdr"]
    b = port_oracle.render(arrays, cam)["hdr"]
    cam.antiAliasingAmount = 1
    one = port_oracle.render(arrays, cam)["hdr"]
    noise = np.abs(a - b).mean()
    shift = np.abs(b - one).mean()   # the jitter is one-sided ([0,1)^2): the multi-sample image is shifted by half a pixel
    assert noise < 0.5 * shift, (noise, shift)
    assert abs((a - b).mean()) < 0.1 * shift
