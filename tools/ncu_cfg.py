"""Developer probe for ncu: render one config a few times (no bench machinery).  python tools/ncu_cfg.py c5 0.25 [n]"""
import sys
sys.path.insert(0, ".")
from raytracingengine_b200 import scenes
from raytracingengine_b200.engine import Engine

name = sys.argv[1]
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
n_it = int(sys.argv[3]) if len(sys.argv) > 3 else 3
arrays, cam = {"c2": scenes.config2, "c3": scenes.config3, "c5": scenes.config5}[name](scale)
e = Engine(0)
e.upload(arrays, cam)
hdr = e.malloc(cam.width * cam.height * 16)
for _ in range(n_it):
    e.render_device({"hdr": hdr})
    s = e.stats()
print(name, cam.width, cam.height, s["kernel_ms"])
