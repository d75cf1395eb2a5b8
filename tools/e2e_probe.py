"""Developer probe: where the end-to-end step (rte_upload_scene + rte_render_host into pinned RGB8) spends its wall time."""
import sys, time, json
sys.path.insert(0, ".")
import numpy as np
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine, PinnedBuffer
arrays, cam = scenes.config2()
e = Engine(0); e.upload(arrays, cam)
n = cam.width * cam.height
pin = PinnedBuffer(n * 3); host = pin.view(np.uint8, (n, 3))
def t(fn, k=30):
    for _ in range(3): fn()
    e.synchronize(); t0 = time.perf_counter()
    for _ in range(k): fn()
    e.synchronize(); return (time.perf_counter() - t0) / k * 1e3
up = t(lambda: (e.upload(arrays, cam), e.synchronize()))
rh_cached = t(lambda: e.render_host(outputs=("rgb8",), into={"rgb8": host}, tonemap_op=_abi.TM_ACES))
rh_rebuild = t(lambda: (e.set_camera(cam), e.render_host(outputs=("rgb8",), into={"rgb8": host}, tonemap_op=_abi.TM_ACES)))
both = t(lambda: (e.upload(arrays, cam), e.render_host(outputs=("rgb8",), into={"rgb8": host}, tonemap_op=_abi.TM_ACES)))
print(json.dumps({"upload_ms": round(up, 4), "render_host_cached_grids_ms": round(rh_cached, 4), "render_host_rebuild_ms": round(rh_rebuild, 4), "upload_plus_render_host_ms": round(both, 4)}))
