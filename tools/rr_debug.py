"""Developer probe for rte_render32r.cu built with -DRTE_RR_DEBUG: histogram of the first inconclusive decision per pixel."""
import sys, collections
sys.path.insert(0, ".")
import numpy as np
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine
e = Engine(0)
arrays, cam = scenes.config1(400, 300)
e.upload(arrays, cam)
r = e.render_host(outputs=("light_mask", "prim_id"), tonemap_op=_abi.TM_ACES)
lm = r["light_mask"]
bad = lm[(lm & 0x80000000) != 0] & 0x7fffffff
print("pixels", lm.size, "flagged", bad.size)
c = collections.Counter((int(v) & 15, int(v) >> 4) for v in bad)
for (why, depth), n in sorted(c.items(), key=lambda x: -x[1])[:20]:
    print(f"  reason {why} at depth {depth}: {n}")
W = cam.width
idx = np.nonzero((lm & 0x80000000) != 0)[0]
for why, depth in ((2, 0), (2, 1)):
    sel = idx[(lm[idx] & 0x7fffffff) == why + 16 * depth]
    ys, xs = sel // W, sel % W
    print(f"reason {why} depth {depth}: {sel.size} px, rows {ys.min()}..{ys.max()}, cols {xs.min()}..{xs.max()}")
    rows = collections.Counter(ys.tolist()).most_common(5)
    print("   busiest rows", rows)
    for y, _ in rows[:2]:
        print("   row", y, "cols", sorted(xs[ys == y].tolist())[:40])
    # prim ids around one flagged pixel (exact path)
ex = e.render_host(outputs=("prim_id",), flags=_abi.FLAG_EXACT_FP64)["prim_id"].reshape(cam.height, W)
sel = idx[(lm[idx] & 0x7fffffff) == 2]
for s_ in sel[:: max(1, sel.size // 6)][:6]:
    y, x = int(s_ // W), int(s_ % W)
    print("   px", x, y, "prim ids row", [hex(int(v)) for v in ex[y, max(0, x - 2):x + 3]])
