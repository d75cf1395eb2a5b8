"""Developer probe for ncu: the default scene (C1) a few times."""
import sys
sys.path.insert(0, ".")
from raytracingengine_b200 import scenes
from raytracingengine_b200.engine import Engine
arrays, cam = scenes.config1()
e = Engine(0)
e.upload(arrays, cam)
hdr = e.malloc(cam.width * cam.height * 16)
for _ in range(3):
    e.render_device({"hdr": hdr})
    s = e.stats()
print("c1", cam.width, cam.height, s["kernel_ms"], s["secondary_rays"], s["primary_rays"], s["shadow_rays"])
