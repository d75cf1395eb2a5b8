"""Developer probe: one launch vs two launches of the fast path on 1/N of the C2 frame (threshold of two_pass_wanted)."""
import sys
sys.path.insert(0, ".")
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine

arrays, cam = scenes.config2()
e = Engine(0)
e.upload(arrays, cam)
n = cam.width * cam.height
hdr = e.malloc(n * 16)
rgb = e.malloc(n * 3 + 16)
for world in (1, 2, 3, 4, 6, 8):
    ropts = dict(row0=0, row1=cam.height) if world == 1 else dict(row0=0, row1=cam.height, stripe_rows=8, stripe_period=8 * world)
    res = {}
    for name, flag in (("one", _abi.FLAG_SINGLE_PASS), ("two", _abi.FLAG_TWO_PASS)):
        ts = []
        for it in range(7):
            e.render_device({"hdr": hdr, "rgb8": rgb}, tonemap_op=_abi.TM_ACES, flags=flag, **ropts)
            ts.append(e.stats()["kernel_ms"])
        res[name] = min(ts[2:])
    print(f"1/{world} of the frame ({n // world / 1e6:.2f} Mpx): one launch {res['one']:.3f} ms, two launches {res['two']:.3f} ms")
