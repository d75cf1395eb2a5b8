"""GPU parity against the COMMITTED golden fixtures (tests/golden/*.npz, written by tests/golden/make_golden.py from the
reference's own headers compiled unmodified): no checker library is needed at run time, only the fixtures.

Bars (BASELINE.json north_star): nearest-hit index and shadow visibility bit-exact, depth / normals <= 1e-5, 8-bit
pixels +-1 LSB.  Diffuse scenes are asked for more: identical FP64 HDR and identical fused RGB8."""
import os

import numpy as np
import pytest

from raytracingengine_b200 import _abi
from tests.golden.make_golden import CASES

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
INEXACT = {"c1_default_64", "c4_box_specular_96x54", "glass_64x48"}  # pow() (Blinn-Phong, Fresnel) on the path


@pytest.mark.parametrize("name", sorted(CASES))
def test_render_matches_reference_golden(engine, name):
    arrays, cam = CASES[name]()
    gold = np.load(os.path.join(GOLD, name + ".npz"))
    engine.upload(arrays, cam)
    g = engine.render_host(outputs=("hdr64", "prim_id", "depth64", "normal", "light_mask"))
    s = engine.stats()
    assert (cam.width, cam.height) == (int(gold["width"]), int(gold["height"]))
    assert np.array_equal(g["prim_id"], gold["prim_id"])
    assert np.array_equal(g["light_mask"], gold["light_mask"])
    hit = g["prim_id"] != _abi.PRIM_MISS
    assert np.array_equal(g["depth64"][hit], gold["depth"][hit])
    np.testing.assert_allclose(g["normal"][:, :3], gold["normal"], rtol=0, atol=1e-5)
    if name in INEXACT:
        np.testing.assert_allclose(g["hdr64"], gold["hdr"], rtol=1e-9, atol=1e-14)
    else:
        assert np.array_equal(g["hdr64"], gold["hdr"])
    assert s["primary_rays"] == int(gold["primary_rays"])
    if name not in INEXACT:  # (the reference-built checker counts the shadow rays of the primary hits only)
        assert s["shadow_rays"] == int(gold["shadow_rays"])
    # the seven operators over the golden HDR (stand-alone FP32 pass): +-1 LSB against the reference's tonemapAll
    h4 = np.zeros((len(gold["hdr"]), 4), np.float32)
    h4[:, :3] = gold["hdr"]
    tm = engine.tonemap_host(h4)
    for op in range(_abi.TM_COUNT):
        d = np.abs(tm[op].astype(np.int16) - gold["rgb8"][op].astype(np.int16))
        # the FP32 pass sees the HDR rounded to FP32: a value within 1e-7 of a quantisation step may fall on either side
        assert d.max() <= 1, f"{name}: {_abi.TM_NAMES[op]} differs by {d.max()} LSB"
