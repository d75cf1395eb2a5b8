#!/usr/bin/env python
"""bench.py — the per-pixel render path on N B200s, against its rooflines and the reference's CPU build.

Contract (one JSON line on stdout from rank 0):
    python bench.py --gpus N --steps K --warmup W            (N > 1: launched with torch.distributed.run)
    python bench.py --impl reference ...                     (the reference's own CPU code, same metric)

Workload (BASELINE.json configs[1], SURVEY.md §8(d) "C2"): 3840x2160, 1024 random spheres + 4 planes,
4 point lights with shadow rays, ACES tonemap.  One "step" = one frame: primary rays -> nearest hit ->
point-light shading with shadow rays -> ACES -> RGB8 (HDR float4 + RGB8 written to HBM).
    value  = (primary + shadow rays of the frame) / device time, scene and framebuffers resident in HBM;
    e2e    = the same through rte_upload_scene + rte_render_host with HOST buffers (scene H2D and RGB8 D2H
             inside the timed region, wall clock);
    N > 1  = the frame is sharded by interleaved row stripes (strong scaling), no data-path collective in the
             render; the RGB8 stripes are all-gathered over NCCL into the PPM writer's buffer inside the step.
Only the `cpu_baseline` leg and `--impl reference` touch oracle/ (the CPU checker); the product path is
librte_b200.so through its C ABI and raises when the library or a GPU is missing.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_OUT = sys.stdout
METRIC = "Mrays/s (primary+shadow) at 3840x2160"
UNIT = "Mrays/s"


def workload(name: str, scale: float):
    from raytracingengine_b200 import scenes
    if name == "c2":
        arrays, cam = scenes.config2(scale)
        label = "C2: 3840x2160, 1024 spheres + 4 planes, 4 point lights + shadow rays, ACES"
    elif name == "c5":
        arrays, cam = scenes.config5(scale)
        label = "C5: 7680x4320, 4096 spheres + 4 planes, 8 point lights + shadow rays, Reinhard-extended-luminance"
    elif name == "c3":
        arrays, cam = scenes.config3(scale)
        label = "C3: 3840x2160, 64 spheres + 4 planes, 1 light, ACES"
    else:
        raise SystemExit(f"unknown workload {name}")
    if scale != 1.0:
        label += f" (image scaled x{scale}: NOT the named config)"
    return arrays, cam, label


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "MEASURED_PEAKS.json"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU with NVML while the timed region runs."""

    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}
    NOTE = {"sw_power_cap": 0x4}

    def __init__(self, index: int, period_s: float = 0.02):
        self.samples, self.reasons, self.power = [], 0, 0.0
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.period = period_s

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                self.reasons |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.power = max(self.power, nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread:
            self._thread.join()

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        s = sorted(self.samples)
        reasons = [k for k, bit in {**self.BAD, **self.NOTE}.items() if self.reasons & bit]
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(s), "power_w_max": round(self.power, 1)}


def ncu_traffic(kernel_prefix: str):
    """DRAM bytes (read + write) of one launch of a kernel from the newest committed `ncu --set full` capture
    (profiles/*_traffic.json, written by tools/profile_summary.py for THIS workload), or None."""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")), reverse=True):
        try:
            with open(path) as f:
                d = json.load(f)
        except (OSError, ValueError):
            continue
        hit = {name: v for name, v in d.items() if kernel_prefix in name}  # (the render step is two launches: <..., 1> and <..., 2>)
        if hit:
            return {"bytes": int(sum(v["dram_bytes_read"] + v["dram_bytes_write"] for v in hit.values())), "source": os.path.relpath(path, ROOT),
                    "kernel": " + ".join(sorted(hit))}
    return None


def visible_index(local_rank: int) -> int:
    """NVML index of cuda:local_rank (CUDA_VISIBLE_DEVICES may remap)."""
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# ------------------------------------------------------------------------------------------ reference arm
def host_threads() -> int:
    """All host cores this process may use — passed explicitly: torchrun exports OMP_NUM_THREADS=1 to its ranks."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_reference_sample(arrays, cam, target_s: float, threads: int = 0):
    """Times the CPU checker (the reference's own code when oracle/_ref was built) on every S-th row of the
    workload, S chosen so that one pass takes about `target_s`.  Returns (oracle, row_step, result dict)."""
    from oracle import pyoracle
    orc = pyoracle.load_best()
    probe_step = max(1, cam.height // 24)
    r = orc.render(arrays, cam, row_step=probe_step, threads=threads or host_threads())
    per_row = r["render_seconds"] / max(1, r["rows"])
    rows = int(min(cam.height, max(8, target_s / max(per_row, 1e-9))))
    step = max(1, cam.height // rows)
    return orc, step


def sample_ray_count(arrays, cam, row_step: int) -> int:
    """primary + shadow rays of the sampled rows (SURVEY §8(d) definition).  The reference-built checker cannot
    count inside the reference's private TraceRay, so the count comes from one untimed pass of the plain-C port
    (bit-identical output, tests/test_oracle.py)."""
    from oracle import pyoracle
    port = pyoracle.load_port()
    if port is None:
        pyoracle.build("port")
        port = pyoracle.load_port()
    r = port.render(arrays, cam, row_step=row_step, threads=host_threads())
    return int(r["primary_rays"] + r["shadow_rays"])


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arrays, cam, label = workload(args.workload, args.scale)
    # bounded sample: the whole --steps/--warmup run stays within a few minutes
    per_pass = min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup))
    orc, row_step = cpu_reference_sample(arrays, cam, per_pass)
    times, threads = [], 0
    rays = sample_ray_count(arrays, cam, row_step)
    for it in range(args.warmup + args.steps):
        r = orc.render(arrays, cam, row_step=row_step, threads=host_threads())
        t = r["render_seconds"]
        t0 = time.perf_counter()
        orc.tonemap_one(6, r["hdr"])  # tonemap() = ACES + toColor, RE/RaytracingEngine.cpp:165-174 (single-threaded upstream)
        t += time.perf_counter() - t0
        if it >= args.warmup:
            times.append(t)
        threads = r["threads_used"]
    total = sum(times)
    value = rays * len(times) / total / 1e6
    frame_rays_est = rays * (cam.height / max(1, r["rows"]))
    sample = f"every {row_step}th row ({r['rows']} of {cam.height} rows, {rays} rays) per step; RenderImage (OpenMP, all host threads) + ACES tonemap (serial, as upstream)"
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(total / len(times) * 1e3, 3), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": label, "width": cam.width, "height": cam.height, **arrays.counts(), "seed": 1234,
                   "sampled_rows": r["rows"], "row_step": row_step},
        "ms_per_frame_extrapolated": round(frame_rays_est / (value * 1e6) * 1e3, 1),
        "cpu_baseline": {"value": round(value, 4), "unit": UNIT, "cores": threads, "kind": orc.kind, "sample": sample},
        "e2e": {"value": round(value, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_OUT, flush=True)


# ------------------------------------------------------------------------------------------ GPU arm
def run_gpu(args) -> None:
    import numpy as np
    import torch

    from raytracingengine_b200 import _abi
    from raytracingengine_b200.engine import Engine, PinnedBuffer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    arrays, cam, label = workload(args.workload, args.scale)
    W, H = cam.width, cam.height
    tm_op = _abi.TM_ACES if args.workload != "c5" else 3
    eng = Engine(local_rank)
    stream = torch.cuda.Stream(device=dev)  # a real (non-default) stream: handle 0 would mean "ctx-private" to the library
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)  # the library enqueues on torch's current stream: torch events and NCCL see its work
    eng.upload(arrays, cam)

    # ---- row assignment: the whole frame at N=1; interleaved stripes of `--stripe` rows at N>1, one launch per rank
    if world == 1:
        ropts = dict(row0=0, row1=H)
    else:
        ropts = dict(row0=rank * args.stripe, row1=H, stripe_rows=args.stripe, stripe_period=world * args.stripe)
    my_rows = Engine.striped_rows(ropts["row0"], H, ropts.get("stripe_rows", 0), ropts.get("stripe_period", 0)) if ropts["row0"] < H else 0
    rows_max = my_rows
    if dist is not None:
        t = torch.tensor([my_rows], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rows_max = int(t.item())
    hdr = torch.empty((max(my_rows, 1) * W, 4), dtype=torch.float32, device=dev)
    rgb_local = torch.zeros((rows_max * W * 3,), dtype=torch.uint8, device=dev)
    rgb_all = torch.empty((world * rows_max * W * 3,), dtype=torch.uint8, device=dev) if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    auto_white = args.workload == "c5" and not args.fixed_white

    # ---- N>1 output path.  "peer" (default): every rank's render kernel stores its RGB8 stripes straight into the
    # full-frame buffer of EVERY rank over NVLink (rte_set_rgb8_peers; buffers exchanged as CUDA IPC handles), and a
    # 4-byte NCCL all-reduce fences the frame.  "nccl": the stripes are all-gathered by NCCL after the kernel.
    gather = "none" if dist is None else ("nccl" if auto_white else args.gather)
    use_flag_fence = False
    peer_note = ""
    frame_all = 0
    opened = []
    if gather == "peer":
        try:
            frame_all = eng.malloc(W * H * 3)
            mine = torch.frombuffer(bytearray(eng.ipc_export(frame_all)), dtype=torch.uint8).to(dev)
            handles = torch.empty(world * _abi.IPC_HANDLE_BYTES, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(handles, mine)
            handles = handles.cpu().numpy().reshape(world, _abi.IPC_HANDLE_BYTES)
            frames = []
            for r in range(world):
                if r == rank:
                    frames.append(frame_all)
                else:
                    opened.append(eng.ipc_open(handles[r].tobytes()))
                    frames.append(opened[-1])
            eng.set_rgb8_peers(frames)
            if args.fence == "flags":
                # frame fence without NCCL: every rank's flag array mapped into every rank like the frames (rte_set_fence_peers)
                my_flags = eng.malloc(4 * _abi.MAX_PEERS)
                mine = torch.frombuffer(bytearray(eng.ipc_export(my_flags)), dtype=torch.uint8).to(dev)
                fh = torch.empty(world * _abi.IPC_HANDLE_BYTES, dtype=torch.uint8, device=dev)
                dist.all_gather_into_tensor(fh, mine)
                fh = fh.cpu().numpy().reshape(world, _abi.IPC_HANDLE_BYTES)
                flags = []
                for r in range(world):
                    if r == rank:
                        flags.append(my_flags)
                    else:
                        opened.append(eng.ipc_open(fh[r].tobytes()))
                        flags.append(opened[-1])
                eng.set_fence_peers(rank, flags)
                use_flag_fence = True
            ok = torch.ones(1, device=dev)
        except Exception as exc:  # no peer mapping on this box: say so and use the NCCL all-gather
            peer_note = f" (peer frames unavailable: {exc})"
            ok = torch.zeros(1, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if ok.item() < 1:
            eng.set_rgb8_peers([])
            if use_flag_fence:
                eng.set_fence_peers(rank, [])
            use_flag_fence = False
            gather = "nccl"
        dist.barrier()  # every rank has zeroed its flags before anybody renders
    fence = torch.zeros(1, dtype=torch.float32, device=dev)
    white = torch.zeros(1, dtype=torch.float32, device=dev)
    tm_ptrs = [rgb_local.data_ptr() if op == tm_op else None for op in range(_abi.TM_COUNT)]

    def render_step():
        if auto_white:
            # C5's auto max-white pipeline: HDR stripes -> per-rank max luminance (device float) ->
            # NCCL all-reduce MAX -> Reinhard-extended-luminance with that white point -> NCCL all-gather of RGB8
            eng.render_device({"hdr": hdr.data_ptr()}, **ropts)
            eng.max_luminance_device(hdr.data_ptr(), my_rows * W, white.data_ptr())
            if dist is not None:
                dist.all_reduce(white, op=dist.ReduceOp.MAX)
            eng.tonemap_auto_white(hdr.data_ptr(), my_rows * W, tm_ptrs, white.data_ptr(), ops=[tm_op])
        elif gather == "peer":
            eng.render_device({"hdr": hdr.data_ptr()}, tonemap_op=tm_op, **ropts)  # RGB8 goes to every rank's frame
            if use_flag_fence:
                eng.frame_fence()            # device-side wait for every rank's "frame delivered" flag (written over NVLink)
            else:
                dist.all_reduce(fence)       # frame fence across ranks through NCCL
            return
        else:
            eng.render_device({"hdr": hdr.data_ptr(), "rgb8": rgb_local.data_ptr()}, tonemap_op=tm_op, **ropts)
        if dist is not None:  # the RGB8 stripes of every rank, into the PPM writer's buffer (de-interleaved at write time)
            dist.all_gather_into_tensor(rgb_all, rgb_local)

    if gather == "peer":
        # one untimed check: the frame the peer stores assemble == the frame NCCL's all-gather assembles
        from raytracingengine_b200 import sharding
        render_step()
        torch.cuda.synchronize()
        eng.render_device({"hdr": hdr.data_ptr(), "rgb8": rgb_local.data_ptr()}, tonemap_op=tm_op, **ropts)
        dist.all_gather_into_tensor(rgb_all, rgb_local)
        torch.cuda.synchronize()
        if rank == 0:
            got = np.zeros((H * W, 3), np.uint8)
            eng.d2h(got, frame_all)
            want = sharding.deinterleave(rgb_all.cpu().numpy().reshape(-1, 3), world, H, W, args.stripe, rows_max)
            if not np.array_equal(got, want):
                raise SystemExit("bench.py: peer-stored frame differs from the NCCL all-gather of the same stripes")

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    fp32_peak_tf, _ = eng.measure_fp32_peak(15.0)
    fp32x2_peak_tf, _ = eng.measure_fp32_peak(15.0, packed=True)

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        render_step()
    sync_all()
    st0 = eng.stats()

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms = []
    sampler = ClockSampler(visible_index(local_rank))
    sync_all()
    wall0 = time.perf_counter()
    with sampler:
        for i in range(args.steps):
            flush.zero_()              # L2 flush between timed iterations (outside the event pair)
            ev[i][0].record(stream)
            render_step()
            ev[i][1].record(stream)
            if world == 1:
                kernel_ms.append(eng.stats()["kernel_ms"])  # syncs; per-launch time of the render kernel alone
        sync_all()
    wall = time.perf_counter() - wall0
    launches_per_render = eng.last_render_launches()  # rte_render of the last timed step: render launch(es) + the RGB8 pass
    step_ms = [a.elapsed_time(b) for a, b in ev]
    my_ms = sum(step_ms)
    st = eng.stats()
    if world > 1:
        kernel_ms = [st["kernel_ms"]]
    counters = torch.tensor([st["primary_rays"] + st["shadow_rays"], st["sphere_tests"], st["plane_tests"], st["triangle_tests"],
                             st["fp64_confirms"], st["exact_marches"], st["sphere_tests_executed"]], dtype=torch.float64, device=dev)
    tmax = torch.tensor([my_ms, sum(kernel_ms) / len(kernel_ms)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(counters, op=dist.ReduceOp.SUM)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    rays, sph_t, pl_t, tri_t, confirms, marches, sph_exec = [float(x) for x in counters.tolist()]
    total_ms, kern_ms = [float(x) for x in tmax.tolist()]
    ms_per_step = total_ms / args.steps
    value = rays / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: C-ABI with HOST buffers — scene description H2D (rte_upload_scene) + render + RGB8 D2H, wall clock
    pin = PinnedBuffer(max(my_rows, 1) * W * 3)
    host_rgb = pin.view(np.uint8, (max(my_rows, 1) * W, 3))
    scene_bytes = sum(a.nbytes for a in (arrays.materials, arrays.spheres, arrays.planes, arrays.triangles, arrays.models,
                                         arrays.model_vertices, arrays.model_indices, arrays.lights))

    if opened or frame_all:
        sync_all()
        eng.set_rgb8_peers([])  # the end-to-end leg returns each rank's stripes to its own host buffer
        if use_flag_fence:
            eng.set_fence_peers(rank, [])

    def e2e_step():
        eng.upload(arrays, cam)
        eng.render_host(outputs=("rgb8",), into={"rgb8": host_rgb}, tonemap_op=tm_op, **ropts)

    for _ in range(2):
        e2e_step()
    sync_all()
    e0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e_ms = (time.perf_counter() - e0) * 1e3
    te = torch.tensor([e_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms = float(te.item()) / args.steps
    e2e_value = rays / (e2e_ms * 1e-3) / 1e6

    # ---- the HBM-bound pass of the path: all seven tonemap operators over the (band) HDR buffer (C3's pass)
    n_px = my_rows * W
    tm_out = [torch.empty(n_px * 3 + 16, dtype=torch.uint8, device=dev) for _ in range(_abi.TM_COUNT)]
    tm_ms = []
    hdr2 = hdr.clone()  # two rotating inputs: 2 x 133 MB in + 174 MB out per launch stream through the 126 MB L2 (no flush needed)
    for i in range(3 + 20):
        eng.tonemap_device((hdr if i % 2 else hdr2).data_ptr(), n_px, [t.data_ptr() for t in tm_out])
        if i >= 3:
            tm_ms.append(eng.stats()["tonemap_ms"])
    tm_med = sorted(tm_ms)[len(tm_ms) // 2]

    if opened:
        sync_all()
        eng.set_rgb8_peers([])
        for p in opened:
            eng.ipc_close(p)
        sync_all()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    pk, pk_src = peaks()
    clocks = sampler.summary()
    tr_render, tr_tm = ncu_traffic("render_fast_kernel"), ncu_traffic("tonemap_kernel")
    flops = _abi.algorithmic_flops(sph_t, pl_t, tri_t)           # 20 / 14 / 46 flop per test
    achieved_tf = flops / world / (kern_ms * 1e-3) / 1e12       # per GPU: per-rank flops / slowest rank's kernel time
    clk = clocks.get("sm_mhz") or pk.get("sm_max_mhz", 1965.0)
    nominal_tf = 148 * 128 * 2 * clk * 1e6 / 1e12
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 filter + f64 confirm", "data": "synthetic",
        "config": {"workload": label, "width": W, "height": H, **arrays.counts(), "seed": 1234 if args.workload != "c5" else 5678,
                   "rays_per_frame": int(rays), "sharding": "single band" if world == 1 else f"interleaved {args.stripe}-row stripes; " + (
                       ("RGB8 stored by the render call's tonemap pass into every rank's frame over NVLink (CUDA IPC peer frames) + " + ("frame fence through NVLink flags written by the same kernel (rte_frame_fence), no collective" if use_flag_fence else "4-byte NCCL all-reduce as frame fence"))
                       if gather == "peer" else "NCCL all-gather of RGB8" + peer_note),
                   "tonemap": ("auto max-white: max-luminance kernel + " + ("NCCL all-reduce MAX + " if world > 1 else "") + "tonemap pass") if auto_white else "exact FP64 ACES pass over the render kernel's FP64 HDR, inside the same rte_render call",
                   "l2": "256 MiB memset between timed steps; outputs (HDR float4 + RGB8 = 19 B/px) exceed L2",
                   "timing": "CUDA events per step on the launching stream, summed, max over ranks"},
        "ms_per_frame": round(ms_per_step, 4),
        "wall_ms_per_step_incl_flush": round(wall / args.steps * 1e3, 3),
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "ms_per_step": round(e2e_ms, 3), "h2d_bytes_per_step": int(scene_bytes),
                "d2h_bytes_per_step": int(my_rows * W * 3), "api": "rte_upload_scene + rte_render_host (pinned host RGB8)"},
        # per step and rank: render kernel + exact tonemap pass (auto-white pipeline: render, max-luminance, tonemap); memsets and NCCL are not ours
        "gpu_launches": args.steps * (launches_per_render + (2 if auto_white else 0)),  # render call (+ max luminance + tonemap for c5); NCCL kernels not counted
        "roofline": {"bound": "fp32", "kernel": "rte::render_fast_kernel (primary-ray launch + shading launch on big frames; + rte::tonemap64_kernel, ~1 % of the call)", "achieved": round(achieved_tf, 3), "peak": round(fp32_peak_tf, 3),
                     "unit": "TFLOP/s", "frac": round(achieved_tf / fp32_peak_tf, 4),
                     "traffic": (tr_render or {}).get("bytes") if args.workload == "c2" and world == 1 else None, "traffic_capture": tr_render,
                     "peak_source": "FFMA micro-benchmark in this process (rte_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
                     "nominal_peak_at_sampled_clock": round(nominal_tf, 2), "ffma2_peak": round(fp32x2_peak_tf, 3),
                     "algorithmic_flop_per_launch": flops / world, "kernel_ms": round(kern_ms, 4),
                     # the view filter runs 3 FMA-pipe lane-ops (6 flop) per sphere test it really executes: the <= 1 companion of `frac`
                     "executed_frac": round(6.0 * sph_exec / world / (kern_ms * 1e-3) / 1e12 / fp32_peak_tf, 4), "sphere_tests_executed": sph_exec,
                     "tests": {"sphere": sph_t, "plane": pl_t, "triangle": tri_t}, "fp64_confirms": confirms, "exact_marches": marches},
        "roofline_tonemap": {"bound": "hbm", "kernel": "rte::tonemap_kernel<all 7 operators>", "achieved": round(n_px * 33 / (tm_med * 1e-3) / 1e9, 1),
                             "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(n_px * 33 / (tm_med * 1e-3) / 1e9 / pk["hbm_gbs"], 4),
                             "traffic": (tr_tm or {}).get("bytes") if args.workload in ("c2", "c3") and world == 1 else None, "traffic_capture": tr_tm,
                             "peak_source": pk_src, "algorithmic_bytes_per_launch": n_px * 33, "kernel_ms": round(tm_med, 4),
                             "l2": "input (16 B/px float4, two rotating buffers) and output (21 B/px) exceed L2; no flush"},
    }
    if not args.no_cpu_baseline and world == 1:
        orc, row_step = cpu_reference_sample(arrays, cam, args.cpu_seconds)
        r = orc.render(arrays, cam, row_step=row_step, threads=host_threads())
        crays = sample_ray_count(arrays, cam, row_step)
        line["cpu_baseline"] = {"value": round(crays / r["render_seconds"] / 1e6, 4), "unit": UNIT, "cores": r["threads_used"], "kind": orc.kind,
                                "sample": f"every {row_step}th row ({r['rows']} of {H} rows, {crays} rays), RenderImage only, {r['render_seconds']:.2f} s"}
    print(json.dumps(line), file=_OUT, flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c5"])
    ap.add_argument("--scale", type=float, default=1.0, help="developer switch: shrink the image (not the named config)")
    ap.add_argument("--stripe", type=int, default=0, help="rows per interleaved stripe at N>1 (0: 8 for c2/c3, 16 for c5 whose streamed-view kernel works on 32x16 pixel batches)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU sample size per pass of the reference leg")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="N>1: how the RGB8 stripes reach every rank")
    ap.add_argument("--fence", default="flags", choices=["flags", "nccl"], help="N>1, --gather peer: frame fence through NVLink flags (rte_frame_fence) or a 4-byte NCCL all-reduce")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fixed-white", action="store_true", help="c5: the reference's constant max_white = 5.0 (fused tonemap) instead of the auto max-white pipeline")
    args = ap.parse_args()
    if args.stripe <= 0:
        args.stripe = 16 if args.workload == "c5" else 8
    # stdout carries exactly ONE line (the JSON); libraries that print to fd 1 (NCCL's version banner) go to stderr
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
