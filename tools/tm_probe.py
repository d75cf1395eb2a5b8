"""Developer probe: the 7-operator tonemap pass over a 4K float4 buffer, median of 40 launches (two rotating inputs).
    [RTE_TM_NULL=1|2] python tools/tm_probe.py   (1: the pass without arithmetic, 2: the same with full-sector stores)"""
import os, sys, json
sys.path.insert(0, ".")
import numpy as np
from raytracingengine_b200 import _abi
from raytracingengine_b200.engine import Engine
e = Engine(0)
n = 3840 * 2160
rng = np.random.default_rng(1)
h = (rng.random((n, 4), dtype=np.float32) * 4.0).astype(np.float32)
a, b = e.malloc(n * 16), e.malloc(n * 16)
e.h2d(a, h); e.h2d(b, h[::-1].copy())
out = [e.malloc(n * 3 + 16) for _ in range(7)]
ts = []
for i in range(45):
    e.tonemap_device(a if i % 2 else b, n, out)
    t = e.stats()["tonemap_ms"]
    if i >= 5: ts.append(t)
ts.sort()
med = ts[len(ts) // 2]
print(json.dumps({"null_probe": os.environ.get("RTE_TM_NULL", ""), "median_us": round(med * 1e3, 2), "min_us": round(ts[0] * 1e3, 2), "GBs": round(n * 33 / med / 1e6, 1)}))
