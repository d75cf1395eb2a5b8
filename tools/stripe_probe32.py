"""Developer probe: one rank's share (1/8) of C5 at 8K on one GPU, as interleaved stripes of different heights (device time)."""
import sys, json
sys.path.insert(0, ".")
import torch
from raytracingengine_b200 import _abi, scenes
from raytracingengine_b200.engine import Engine
arrays, cam = scenes.config5()
torch.cuda.set_device(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
e = Engine(0); e.set_stream(st.cuda_stream); e.upload(arrays, cam)
W, H = cam.width, cam.height
rgb = torch.zeros(H * W * 3 + 16, dtype=torch.uint8, device="cuda")
hdr = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(**o):
    ts = []
    for i in range(8):
        flush.zero_()
        e.render_device({"hdr": hdr.data_ptr(), "rgb8": rgb.data_ptr()}, tonemap_op=_abi.TM_REINHARD_EXT_LUM, **o)
        k = e.stats()["kernel_ms"]
        if i >= 3: ts.append(k)
    return round(sum(ts) / len(ts), 4)
full = run(row0=0, row1=H)
print(json.dumps({"fractions_stripes8": {f"1/{n}": run(row0=0, row1=H, stripe_rows=8, stripe_period=8 * n) for n in (2, 4, 8, 16, 32, 64)}}))
print(json.dumps({"full_frame_ms": full, "ideal_eighth": round(full / 8, 4),
                  **{f"stripes_{s}": run(row0=3 * s, row1=H, stripe_rows=s, stripe_period=8 * s) for s in (4, 8, 16, 32, 64, 540)}}))
