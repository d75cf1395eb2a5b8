"""Developer probe: device time of one rank's share of a striped C2 frame (strong-scaling fixed costs)."""
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
full = None
for world in (1, 2, 4, 8):
    ropts = dict(row0=0, row1=cam.height) if world == 1 else dict(row0=0, row1=cam.height, stripe_rows=8, stripe_period=8 * world)
    ts = []
    for it in range(6):
        e.render_device({"hdr": hdr, "rgb8": rgb}, tonemap_op=_abi.TM_ACES, **ropts)
        ts.append(e.stats()["kernel_ms"])
    t = min(ts[2:])
    full = full or t
    print(f"1/{world} of the frame: {t:.3f} ms  (ideal {full / world:.3f}, efficiency {full / world / t:.3f})")
ropts = dict(row0=0, row1=cam.height, stripe_rows=8, stripe_period=64)
for name, frame, kw in (("1/8, HDR only (no tonemap64 pass)", {"hdr": hdr}, {}), ("1/8, HDR + RGB8", {"hdr": hdr, "rgb8": rgb}, {"tonemap_op": _abi.TM_ACES})):
    ts = []
    for it in range(6):
        e.render_device(frame, **kw, **ropts)
        ts.append(e.stats()["kernel_ms"])
    print(f"{name}: {min(ts[2:]):.3f} ms")
ts = []
for it in range(6):
    e.render_device({"hdr": hdr, "rgb8": rgb}, tonemap_op=_abi.TM_ACES, row0=0, row1=8)
    ts.append(e.stats()["kernel_ms"])
print(f"8 rows (launch floor): {min(ts[2:]):.3f} ms")
