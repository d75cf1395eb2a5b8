"""Developer probe: cost of rebuilding the direction grids per frame, by G and by band height (device time through CUDA events).

    RTE_GRID_G=64 python tools/grid_probe.py [c2|c5] [rows_divisor]
"""
import os, sys, json
sys.path.insert(0, ".")
import torch
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
div = int(sys.argv[2]) if len(sys.argv) > 2 else 1
arrays, cam = {"c2": scenes.config2, "c5": lambda: scenes.config5(0.5)}[name]()
torch.cuda.set_device(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
e = Engine(0); e.set_stream(st.cuda_stream); e.upload(arrays, cam)
rows = cam.height // div
rgb = torch.zeros(rows * cam.width * 3 + 16, dtype=torch.uint8, device="cuda")
hdr = torch.zeros(rows * cam.width * 4, dtype=torch.float32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(rebuild, n=30):
    tot = 0.0
    for i in range(n + 3):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        if rebuild: e.set_camera(cam)
        e.render_device({"hdr": hdr.data_ptr(), "rgb8": rgb.data_ptr()}, tonemap_op=_abi.TM_ACES, row0=0, row1=rows)
        b.record(st); torch.cuda.synchronize()
        if i >= 3: tot += a.elapsed_time(b)
    return tot / n
print(json.dumps({"cfg": name, "G": os.environ.get("RTE_GRID_G", "default"), "rows": rows, "cached_ms": round(run(False), 4), "rebuild_ms": round(run(True), 4)}), flush=True)
