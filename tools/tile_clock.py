import sys; sys.path.insert(0, ".")
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine
import torch
e = Engine(0)
for name, mk, op in (("c2", scenes.config2, _abi.TM_ACES), ("c5_8k", scenes.config5, _abi.TM_REINHARD_EXT_LUM)):
    arrays, cam = mk(); e.upload(arrays, cam)
    n = cam.width * cam.height
    hdr = e.malloc(n * 16); rgb = e.malloc(n * 3 + 16)
    for i in range(3):
        e.render_device({"hdr": hdr, "rgb8": rgb}, tonemap_op=op); s = e.stats()
    print(name, "kernel_ms", round(s["kernel_ms"], 3), "slowest tile us", s["secondary_rays"] / 1e3, "slowest exact batch (ns * 1000 + pixels)", s["triangle_tests"])
