"""GPU parity of the FP32 production path for REFLECTIVE scenes with few primitives (csrc/rte_render32r.cu + the exact kernel
over the tiles it flags): the reference's default scene (C1) and relatives.

Bars (BASELINE.json north_star): nearest-hit (type, index) and shadow visibility of the primary hit bit-exact, RGB8 within
+-1 LSB of the reference (the exact path, itself checked against the reference build in test_gpu_render.py / test_gpu_golden.py)."""
import dataclasses

import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes

pytestmark = pytest.mark.gpu
OUTS = ("hdr", "rgb8", "prim_id", "light_mask")


def mirror_room(width, height, n_spheres=12, seed=77):
    """A synthetic room whose spheres are mirrors of various strengths (specular 0.05 .. 0.9) and whose planes are the slightly
    specular planes of main(): chains of sphere-to-sphere reflections that keep their weight."""
    arrays, cam = scenes.synthetic_room(width, height, width * 0.6, n_spheres, 3, seed, radius_scale=2.5)
    rng = np.random.default_rng(seed)
    mats = arrays.materials.copy()
    mats["specular"] = rng.uniform(0.05, 0.9, len(mats))
    mats["shininess"] = rng.uniform(0.5, 200.0, len(mats))
    arrays.materials = mats
    return arrays, cam


CASES = {
    "c1": lambda: scenes.config1(240, 180),
    "c1_square": lambda: scenes.config1(160, 160),
    "c4_specular": lambda: scenes.config4("box", scale=0.12, specular=True),
    "mirror_room": lambda: mirror_room(200, 120),
    "mirror_room_28": lambda: mirror_room(200, 120, n_spheres=28, seed=5),   # near the limit of 32 primitives
    "glass": lambda: scenes.glass_scene(160, 120),        # refraction, Fresnel, total internal reflection, transparent occluders: TraceRay is a tree
    "glass_small": lambda: scenes.glass_scene(61, 47),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_reflective_fp32_path_against_exact_path(engine, case):
    arrays, cam = CASES[case]()
    engine.upload(arrays, cam)
    fast = engine.render_host(outputs=OUTS, tonemap_op=_abi.TM_ACES)
    sf, lf = engine.stats(), engine.last_render_launches()
    exact = engine.render_host(outputs=OUTS, tonemap_op=_abi.TM_ACES, flags=_abi.FLAG_EXACT_FP64)
    se = engine.stats()
    assert lf == 2, "the reflective FP32 kernel + the exact kernel over its flagged tiles"
    n_px = cam.width * cam.height
    assert np.array_equal(fast["prim_id"], exact["prim_id"]), f"{np.count_nonzero(fast['prim_id'] != exact['prim_id'])} nearest-hit mismatches"
    assert np.array_equal(fast["light_mask"], exact["light_mask"]), f"{np.count_nonzero(fast['light_mask'] != exact['light_mask'])} visibility mismatches"
    d = np.abs(fast["rgb8"].astype(np.int16) - exact["rgb8"].astype(np.int16))
    assert d.max() <= 1, f"{np.count_nonzero(d > 1)} bytes differ by more than 1 LSB (max {d.max()})"
    np.testing.assert_allclose(fast["hdr"][:, :3], exact["hdr"][:, :3], rtol=2e-4, atol=2e-4)
    assert sf["primary_rays"] == se["primary_rays"] == n_px
    for k in ("shadow_rays", "secondary_rays"):  # (branches end below a weight of 1e-7: the production path traces fewer rays than the reference)
        assert 0 < sf[k] <= se[k], f"counter {k}: fp32 path {sf[k]}, exact path {se[k]}"
    frac = sf["fp64_confirms"] / n_px
    print(f"{case}: {100 * frac:.1f} % of the pixels in flagged tiles; kernel {sf['kernel_ms']:.3f} ms vs exact {se['kernel_ms']:.3f} ms")
    assert frac == 0.0  # (inconclusive interval decisions are settled in the kernel by the reference's own tests: no tile is left to the exact kernel)


def test_reflective_fp32_path_multisample_bands_and_operators(engine):
    arrays, cam = scenes.config1(96, 72)
    cam = dataclasses.replace(cam, antiAliasingAmount=4, jitterSeed=5)
    engine.upload(arrays, cam)
    for op in (_abi.TM_REINHARD_JODIE, _abi.TM_UNCHARTED2):
        fast = engine.render_host(outputs=("rgb8", "prim_id"), tonemap_op=op, row0=8, row1=60)
        exact = engine.render_host(outputs=("rgb8", "prim_id"), tonemap_op=op, row0=8, row1=60, flags=_abi.FLAG_EXACT_FP64)
        assert np.array_equal(fast["prim_id"], exact["prim_id"])
        assert np.abs(fast["rgb8"].astype(np.int16) - exact["rgb8"].astype(np.int16)).max() <= 1
    # stripes, max_depth 0 / 1 / 3 (recursion limit -> sky), bias option
    for md in (0, 1, 3):
        fast = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES, max_depth=md, row0=4, row1=68, stripe_rows=4, stripe_period=12)
        exact = engine.render_host(outputs=("rgb8",), tonemap_op=_abi.TM_ACES, max_depth=md, row0=4, row1=68, stripe_rows=4, stripe_period=12, flags=_abi.FLAG_EXACT_FP64)
        assert np.abs(fast["rgb8"].astype(np.int16) - exact["rgb8"].astype(np.int16)).max() <= 1, f"max_depth {md}"
