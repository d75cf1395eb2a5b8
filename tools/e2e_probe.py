"""Developer probe: wall-clock split of the end-to-end call sequence (upload, render_host) on C2."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine, PinnedBuffer

arrays, cam = scenes.config2()
e = Engine(0)
pin = PinnedBuffer(cam.width * cam.height * 3)
host = pin.view(np.uint8, (cam.width * cam.height, 3))
for _ in range(3):
    e.upload(arrays, cam)
    e.render_host(outputs=("rgb8",), into={"rgb8": host}, tonemap_op=_abi.TM_ACES)
N = 20
tu = tr = 0.0
for _ in range(N):
    t0 = time.perf_counter(); e.upload(arrays, cam); t1 = time.perf_counter()
    e.render_host(outputs=("rgb8",), into={"rgb8": host}, tonemap_op=_abi.TM_ACES); t2 = time.perf_counter()
    tu += t1 - t0; tr += t2 - t1
print(f"upload {tu / N * 1e3:.3f} ms, render_host {tr / N * 1e3:.3f} ms, kernel {e.stats()['kernel_ms']:.3f} ms, checksum {int(host.sum())}")
