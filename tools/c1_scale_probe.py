import sys
sys.path.insert(0, ".")
from raytracingengine_b200 import scenes
from raytracingengine_b200.engine import Engine
e = Engine(0)
for w in (500, 1000, 2000, 4000):
    arrays, cam = scenes.config1(w, w)
    e.upload(arrays, cam)
    hdr = e.malloc(w * w * 16)
    ts = []
    for _ in range(4):
        e.render_device({"hdr": hdr}); s = e.stats(); ts.append(s["kernel_ms"])
    rays = s["primary_rays"] + s["shadow_rays"] + s["secondary_rays"]
    print(f"c1 {w}x{w}: {min(ts[1:]):.3f} ms, {rays / min(ts[1:]) / 1e3:.0f} Mrays/s, {min(ts[1:]) * 1e6 / (w * w):.2f} ns/px")
