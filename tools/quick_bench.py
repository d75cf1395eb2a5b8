"""Developer probe (not the bench contract): device time of the render / tonemap kernels per config."""
import os, sys, time, json
import numpy as np
sys.path.insert(0, ".")
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine

def main():
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c2", "c3", "c4", "c5", "c1"]
    e = Engine(0)
    for name in which:
        arrays, cam = {"c2": lambda: scenes.config2(), "c3": lambda: scenes.config3(), "c4": lambda: scenes.config4("blob"),
                       "c5": lambda: scenes.config5(0.5), "c1": lambda: scenes.config1()}[name]()
        e.upload(arrays, cam)
        n = cam.width * cam.height
        hdr = e.malloc(n * 16)
        rgb = [e.malloc(n * 3 + 16) for _ in range(7)]
        full = "--rgb8" in sys.argv
        for it in range(3):
            e.render_device({"hdr": hdr, "rgb8": rgb[0]} if full else {"hdr": hdr}, flags=int(os.environ.get("RTE_FLAGS", "0")), **({"tonemap_op": _abi.TM_ACES} if full else {}))
            s = e.stats()
        flops = _abi.algorithmic_flops(s["sphere_tests"], s["plane_tests"], s["triangle_tests"])
        rays = s["primary_rays"] + s["shadow_rays"] + s["secondary_rays"]
        for it in range(3):
            e.tonemap_device(hdr, n, rgb)
            t = e.stats()["tonemap_ms"]
        print(json.dumps({"cfg": name, "W": cam.width, "H": cam.height, "kernel_ms": round(s["kernel_ms"], 3),
                          "Mrays/s": round(rays / s["kernel_ms"] / 1e3, 1), "alg_TFLOPs": round(flops / s["kernel_ms"] / 1e9, 2),
                          "confirms_per_ray": round(s["fp64_confirms"] / rays, 2), "exec_frac": round(s["sphere_tests_executed"] / max(1, s["sphere_tests"]), 3), "marches": s["exact_marches"],
                          "tonemap7_ms": round(t, 4), "tonemap_GBs": round(n * 33 / t / 1e6, 1)}), flush=True)

main()
