"""Out-of-bounds writes: every output buffer of a render / tonemap call sits between two guard zones filled with a pattern;
after the call the guards are intact and the payload equals what rte_render_host returns for the same call.
(compute-sanitizer is closed on the GPU pool, profiles/r02_sanitizer.txt; this is the bounds check of our own.)"""
import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import FRAME_FIELDS

pytestmark = pytest.mark.gpu
GUARD = 4096
PATTERN = 0xA5


class Guarded:
    def __init__(self, eng, nbytes):
        self.eng, self.n = eng, nbytes
        self.pad = (-nbytes) % 256  # keeps the rear guard 256-byte aligned like the front one
        self.base = eng.malloc(GUARD + nbytes + self.pad + GUARD)
        eng.h2d(self.base, np.full(GUARD + nbytes + self.pad + GUARD, PATTERN, np.uint8))
        self.ptr = self.base + GUARD

    def check(self, what):
        raw = np.empty(GUARD + self.n + self.pad + GUARD, np.uint8)
        self.eng.d2h(raw, self.base)
        assert (raw[:GUARD] == PATTERN).all(), f"{what}: bytes in front of the buffer were written"
        assert (raw[GUARD + self.n:] == PATTERN).all(), f"{what}: bytes behind the buffer were written"
        return raw[GUARD:GUARD + self.n]


CASES = {
    # name: (scene, outputs, opts)
    "fp32_odd": (lambda: scenes.synthetic_room(203, 117, 150.0, 300, 3, 11), ("hdr", "rgb8", "prim_id", "light_mask"), dict(tonemap_op=_abi.TM_ACES)),
    "fp32_band": (lambda: scenes.config2(scale=0.1), ("hdr", "rgb8"), dict(tonemap_op=_abi.TM_REINHARD_JODIE, row0=13, row1=150)),
    "fp32_stripes": (lambda: scenes.config2(scale=0.1), ("rgb8", "prim_id"), dict(tonemap_op=_abi.TM_ACES, row0=8, row1=216, stripe_rows=8, stripe_period=24)),
    "exact_views": (lambda: scenes.config2(scale=0.1), ("hdr", "hdr64", "rgb8", "prim_id", "depth", "depth64", "depth_norm", "normal", "albedo", "light_mask"),
                    dict(tonemap_op=_abi.TM_ACES)),
    "exact_grid": (lambda: scenes.mixed_scene(187, 101, 140.0), ("hdr64", "rgb8", "prim_id", "depth64", "normal", "light_mask"),
                   dict(tonemap_op=_abi.TM_UNCHARTED2, flags=_abi.FLAG_EXACT_GRID)),
    "reflective": (lambda: scenes.config1(121, 77), ("hdr64", "rgb8", "prim_id", "depth_norm"), dict(tonemap_op=_abi.TM_ACES)),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_render_outputs_stay_inside_their_buffers(engine, case):
    make, outs, opts = CASES[case]
    arrays, cam = make()
    engine.upload(arrays, cam)
    n = engine.band_pixels(opts.get("row0", 0), opts.get("row1", 0), opts.get("stripe_rows", 0), opts.get("stripe_period", 0))
    ref = engine.render_host(outputs=outs, **opts)
    bufs = {}
    for name in outs:
        dt, tail = FRAME_FIELDS[name]
        bufs[name] = Guarded(engine, n * int(np.prod(tail, dtype=np.int64)) * np.dtype(dt).itemsize)
    engine.render_device({k: b.ptr for k, b in bufs.items()}, **opts)
    engine.synchronize()
    for name in outs:
        dt, _ = FRAME_FIELDS[name]
        got = bufs[name].check(f"{case}.{name}").view(dt)
        want = ref[name].reshape(-1)
        if name in ("hdr", "normal", "albedo"):  # float4: the fourth component is padding
            assert np.array_equal(got.reshape(-1, 4)[:, :3], want.reshape(-1, 4)[:, :3], equal_nan=True), f"{case}.{name}"
        else:
            assert np.array_equal(got, want, equal_nan=True), f"{case}.{name}"


@pytest.mark.parametrize("n", [1, 3, 4, 1021, 4099, 65537])
def test_tonemap_outputs_stay_inside_their_buffers(engine, n):
    rng = np.random.default_rng(n)
    hdr = (rng.random((n, 4), dtype=np.float32) * 3.0).astype(np.float32)
    src = Guarded(engine, hdr.nbytes)
    engine.h2d(src.ptr, hdr)
    outs = [Guarded(engine, 3 * n) for _ in range(_abi.TM_COUNT)]
    engine.tonemap_device(src.ptr, n, [o.ptr for o in outs])
    engine.synchronize()
    want = engine.tonemap_host(hdr)
    for op, o in enumerate(outs):
        got = o.check(f"tonemap op {op}, n = {n}").reshape(-1, 3)
        assert np.array_equal(got, want[op]), f"op {op}"
    white = Guarded(engine, 4)
    engine.max_luminance_device(src.ptr, n, white.ptr)
    engine.synchronize()
    white.check("max luminance")


@pytest.mark.parametrize("case", ["c1", "c2", "mixed_grid"])
def test_depth_norm_is_normalized_distance(engine, case):
    """frame.depth_norm = HitInfo::normalizedDistance(camera) of the primary hit (RE/Shape.h:40-42), FP64 arithmetic rounded to
    FP32, +inf on a miss; asking for it alone (no depth64) works too."""
    flags = 0
    if case == "c1":
        arrays, cam = scenes.config1(160, 120)
    elif case == "c2":
        arrays, cam = scenes.config2(scale=0.05)
        cam.nearPlaneDistance, cam.farPlaneDistance = 2.5, 750.0
    else:
        arrays, cam = scenes.mixed_scene(150, 90, 120.0)
        cam.nearPlaneDistance, cam.farPlaneDistance = 1.0, 300.0
        flags = _abi.FLAG_EXACT_GRID
    engine.upload(arrays, cam)
    both = engine.render_host(outputs=("depth64", "depth_norm"), flags=flags)
    alone = engine.render_host(outputs=("depth_norm",), flags=flags)
    with np.errstate(invalid="ignore"):
        want = ((both["depth64"] - cam.nearPlaneDistance) / (cam.farPlaneDistance - cam.nearPlaneDistance)).astype(np.float32)
    assert np.isfinite(both["depth64"]).any()
    assert np.array_equal(both["depth_norm"], want) and np.array_equal(alone["depth_norm"], want)
    assert np.all(np.isinf(want[np.isinf(both["depth64"])]))
