"""Writes tests/golden/frame_crc.json: CRC-32 of the RGB8 frame the CPU checker (the reference's own code when
oracle/_ref is built) produces for every bench workload at its NAMED size.  bench.py compares the frame its exact
path renders with these values ("frame_ok").  Run once in the authoring container:

    python tools/make_frame_crc.py [c2 c3 c5]

C5 at 7680x4320 is ~10^12 ray-sphere tests: about ten minutes on eight cores.
"""
import json
import os
import sys
import time
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import pyoracle  # noqa: E402
from raytracingengine_b200 import _abi, scenes  # noqa: E402

WORKLOADS = {  # name -> (scene, operator of the bench's render call)
    "c2": (scenes.config2, _abi.TM_ACES),
    "c3": (scenes.config3, _abi.TM_ACES),
    "c5": (scenes.config5, _abi.TM_REINHARD_EXT_LUM),  # max_white = 5.0 as at the reference's call sites
}


def main():
    path = os.path.join(ROOT, "tests", "golden", "frame_crc.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    orc = pyoracle.load_best()
    for name in (sys.argv[1:] or list(WORKLOADS)):
        make, op = WORKLOADS[name]
        arrays, cam = make()
        t0 = time.time()
        crc, rows_done = 0, 0
        chunk = 135  # rows per oracle call (bounded memory: C5's FP64 frame is 796 MB)
        while rows_done < cam.height:
            r1 = min(cam.height, rows_done + chunk)
            hdr = orc.render(arrays, cam, row0=rows_done, row1=r1)["hdr"]
            crc = zlib.crc32(orc.tonemap_one(op, hdr).tobytes(), crc)
            rows_done = r1
        out[name] = {"width": cam.width, "height": cam.height, "operator": _abi.TM_NAMES[op], "crc32": crc, "oracle": orc.kind,
                     **arrays.counts()}
        print(name, out[name], f"{time.time() - t0:.1f} s", flush=True)
        with open(path, "w") as f:
            json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
