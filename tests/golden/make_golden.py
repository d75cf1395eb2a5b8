"""Generates tests/golden/*.npz + kat.json from the REFERENCE-BUILT checker (oracle/_ref/libref_oracle.so,
i.e. the reference's own headers compiled unmodified from /root/reference).  Run in the authoring container:

    make -C oracle ref && python tests/golden/make_golden.py

The fixtures pin oracle/oracle_port.c (tests/test_oracle.py) and, on the GPU box where /root/reference does
not exist, the CUDA path (tests/test_gpu_golden.py).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import pyoracle  # noqa: E402
from raytracingengine_b200 import _abi, scenes  # noqa: E402
from raytracingengine_b200.scene import Camera, SceneArrays  # noqa: E402

CASES = {
    "c1_default_64": lambda: scenes.config1(64, 64),
    "c2_spheres_96x54": lambda: scenes.config2(scale=0.025),
    "c3_64spheres_96x54": lambda: scenes.config3(scale=0.025),
    "c4_blob_96x54": lambda: scenes.config4("blob", scale=0.05),
    "c4_box_specular_96x54": lambda: scenes.config4("box", scale=0.05, specular=True),
    "c5_4096_60x33": lambda: scenes.config5(scale=1 / 128),
    "mixed_96x54": lambda: scenes.mixed_scene(96, 54, 48.0),
    "glass_64x48": lambda: scenes.glass_scene(64, 48),   # row N1: refraction / Fresnel / transmittance
}


def kat_scene():
    """unit sphere @0, plane p0(0,-15,0) n(0,1,0), triangle (-1,-1,0),(1,-1,0),(0,1,0) at z+100."""
    mats = np.zeros(1, dtype=_abi.material_dt)
    mats[0] = ((1, 1, 1), 128.0, 0.0, 0.0, 1.0)
    return mats


def main():
    ref = pyoracle.load_ref()
    assert ref is not None and ref.kind == "reference", "build oracle/_ref first (needs /root/reference)"
    only_missing = "--missing" in sys.argv
    for name, make in CASES.items():
        if only_missing and os.path.exists(os.path.join(HERE, name + ".npz")):
            continue
        arrays, cam = make()
        r = ref.render(arrays, cam, want_aovs=True)
        rgb = ref.tonemap_all(r["hdr"])
        np.savez_compressed(os.path.join(HERE, name + ".npz"), hdr=r["hdr"], prim_id=r["prim_id"], depth=r["depth"],
                            normal=r["normal"], albedo=r["albedo"], light_mask=r["light_mask"],
                            rgb8=np.stack(rgb), width=cam.width, height=cam.height,
                            primary_rays=r["primary_rays"], shadow_rays=r["shadow_rays"])
        print(name, cam.width, cam.height, r["primary_rays"], r["shadow_rays"])

    if only_missing:
        return
    kat = {}
    cam = Camera((0.0, 0.0, -25.0), 500.0, 1000, 1000, 0.0, 200.0)
    kat["camera"] = {f"{x},{y}": ref.camera_ray(cam, x, y)[1].tolist() for x, y in [(500, 500), (0, 0), (999, 999), (250, 750), (1, 998)]}
    # ray queries against single primitives (Scene::IntersectClosest over a one-primitive scene)
    mats = kat_scene()
    sph = np.zeros(1, dtype=_abi.sphere_dt); sph[0] = ((0, 0, 0), 1.0, 0, 0)
    pln = np.zeros(1, dtype=_abi.plane_dt); pln[0] = ((0, -15, 0), (0, 1, 0), 0, 0)
    tri = np.zeros(1, dtype=_abi.triangle_dt); tri[0] = ((-1, -1, 0), (1, -1, 0), (0, 1, 0), (0, 0, 0), 0, 0)
    s2 = 2.0 ** -0.5
    queries = {
        "sphere": (SceneArrays(materials=mats, spheres=sph),
                   [((0, 0, -5), (0, 0, 1)), ((0, 0, 0), (0, 0, 1)), ((0, 0, -1), (0, 0, -1)), ((1, 0, -5), (0, 0, 1)), ((0, 0, -5), (0, 0, 2))]),
        "plane": (SceneArrays(materials=mats, planes=pln),
                  [((0, 0, -25), (0, -s2, s2)), ((3, -15, 2), (0, -1, 0)), ((0, 0, 0), (1, 0, 0)), ((0, -20, 0), (0, 1, 0)), ((0, 0, 0), (0, 1, 0))]),
        "triangle": (SceneArrays(materials=mats, triangles=tri),
                     [((0, 0, -3), (0, 0, 1)), ((0, 0, 3), (0, 0, -1)), ((-1, -1, -3), (0, 0, 1)), ((2, 2, -3), (0, 0, 1)), ((0, 0, -3), (0, 0, -1))]),
    }
    kat["shapes"] = {}
    for name, (arr, rays) in queries.items():
        o = np.array([r[0] for r in rays], dtype=np.float64)
        d = np.array([r[1] for r in rays], dtype=np.float64)
        h = ref.intersect_closest(arr, o, d)
        kat["shapes"][name] = {"origins": o.tolist(), "directions": d.tolist(), "prim_id": h["prim_id"].tolist(),
                               "distance": [None if p == _abi.PRIM_MISS else float(t) for p, t in zip(h["prim_id"], h["distance"])],
                               "normal": h["normal"].tolist()}
    hdr = np.array([[0, 0, 0], [.18, .18, .18], [1, 1, 1], [.5, 1, 2], [5, 5, 5], [10, .1, .01], [.003, .5, 30], [-1, np.nan, 1e30]], dtype=np.float64)
    kat["tonemap"] = {"hdr": [[None if np.isnan(v) else v for v in row] for row in hdr.tolist()],
                      "rgb8": [a.tolist() for a in ref.tonemap_all(hdr)]}
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(kat, f, indent=1)
    print("kat.json written")


if __name__ == "__main__":
    main()
