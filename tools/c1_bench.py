"""C1 — the reference's default workload (main()'s scene, 1000x1000, reflective planes + the box mesh, 2 lights) on both arms,
at antiAliasingAmount = 1 and at the reference's default 32 samples per pixel (RE/Math.h:94).  One JSON line.

    python tools/c1_bench.py [--cpu-seconds 10] > profiles/r02_c1.json

GPU arm: rte_render (device HDR + RGB8 with ACES), CUDA events, 256 MiB L2 flush between frames, clocks sampled with NVML.
CPU arm: the reference's own RenderImage (oracle/_ref) on all host cores over every S-th row (bounded), extrapolated to the frame.
"""
import argparse, dataclasses, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (ClockSampler, cpu_reference_sample, host_threads)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    from raytracingengine_b200 import _abi, scenes
    from raytracingengine_b200.engine import Engine
    torch.cuda.set_device(0)
    st = torch.cuda.Stream(); torch.cuda.set_stream(st)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {"workload": "C1: main()'s scene at 1000x1000 (5 reflective planes, box mesh, 2 lights, maxRecursion 10)", "gpu": {}, "cpu": {}}
    for aa in (1, 32):
        arrays, cam = scenes.config1()
        cam = dataclasses.replace(cam, antiAliasingAmount=aa, jitterSeed=7)
        e = Engine(0); e.set_stream(st.cuda_stream); e.upload(arrays, cam)
        n = cam.width * cam.height
        hdr = torch.empty(n * 4, dtype=torch.float32, device="cuda"); rgb = torch.empty(n * 3 + 16, dtype=torch.uint8, device="cuda")
        steps = args.steps if aa == 1 else max(3, args.steps // 3)
        ts = []
        sampler = bench.ClockSampler(bench.visible_index(0))
        with sampler:
            for i in range(3 + steps):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(st)
                e.set_camera(cam)
                e.render_device({"hdr": hdr.data_ptr(), "rgb8": rgb.data_ptr()}, tonemap_op=_abi.TM_ACES)
                b.record(st); torch.cuda.synchronize()
                if i >= 3: ts.append(a.elapsed_time(b))
        s = e.stats()
        rays = s["primary_rays"] + s["shadow_rays"] + s["secondary_rays"]
        ms = sum(ts) / len(ts)
        out["gpu"][f"aa{aa}"] = {"ms_per_frame": round(ms, 4), "min_ms": round(min(ts), 4), "rays_per_frame": rays, "Mrays_per_s": round(rays / ms / 1e3, 1),
                                 "steps": steps, "launches": e.last_render_launches(), "clocks": sampler.summary()}
        e.close()
        if not args.no_cpu:
            orc, row_step = bench.cpu_reference_sample(arrays, cam, args.cpu_seconds)
            r = orc.render(arrays, cam, row_step=row_step, threads=bench.host_threads())
            frame_s = r["render_seconds"] * cam.height / max(1, r["rows"])
            out["cpu"][f"aa{aa}"] = {"kind": orc.kind, "cores": r["threads_used"], "sample": f"every {row_step}th row ({r['rows']} of {cam.height})",
                                     "sample_seconds": round(r["render_seconds"], 3), "ms_per_frame_extrapolated": round(frame_s * 1e3, 1),
                                     "speedup_gpu": round(frame_s * 1e3 / ms, 1),
                                     "note": "the reference build draws its own random jitter (std::random_device) at aa > 1: same work, different samples"}
    print(json.dumps(out))


main()
