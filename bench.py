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
METRIC = "Mrays/s (primary+shadow) at 3840x2160"  # (the reference arm; the GPU arm names its own frame size)
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
        "impl": "reference", "metric": f"Mrays/s (primary+shadow) at {cam.width}x{cam.height}", "value": round(value, 4), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
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
def frame_crc_golden(name: str):
    """CRC-32 of the reference build's RGB8 frame of a workload at its named size (tools/make_frame_crc.py), or None."""
    try:
        with open(os.path.join(ROOT, "tests", "golden", "frame_crc.json")) as f:
            return json.load(f).get(name)
    except (OSError, ValueError):
        return None


def ncu_kernels():
    """Per-kernel hardware figures of the newest committed ncu capture (profiles/*_kernels.json, tools/profile_summary.py)."""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_kernels.json")), reverse=True):
        try:
            with open(path) as f:
                return json.load(f), os.path.relpath(path, ROOT)
        except (OSError, ValueError):
            continue
    return None, None


class Pipeline:
    """One workload on this rank: engine, row assignment, output buffers, N>1 output exchange, one `step()`."""

    def __init__(self, args, name, torch, dist, dev, rank, world, local_rank, stream):
        import numpy as np
        from raytracingengine_b200 import _abi
        from raytracingengine_b200.engine import Engine
        self.np, self.torch, self.dist, self.abi = np, torch, dist, _abi
        self.args, self.name, self.rank, self.world, self.dev = args, name, rank, world, dev
        self.arrays, self.cam, self.label = workload(name, args.scale)
        W, H = self.cam.width, self.cam.height
        self.W, self.H = W, H
        self.tm_op = _abi.TM_ACES if name != "c5" else _abi.TM_REINHARD_EXT_LUM
        self.stripe = args.stripe if args.stripe > 0 else (16 if name == "c5" else 8)
        self.eng = Engine(local_rank)
        self.eng.set_stream(stream.cuda_stream)  # the library enqueues on torch's current stream: torch events and NCCL see its work
        self.eng.upload(self.arrays, self.cam)
        if world == 1:
            self.ropts = dict(row0=0, row1=H)
        else:
            self.ropts = dict(row0=rank * self.stripe, row1=H, stripe_rows=self.stripe, stripe_period=world * self.stripe)
        self.my_rows = Engine.striped_rows(self.ropts["row0"], H, self.ropts.get("stripe_rows", 0), self.ropts.get("stripe_period", 0)) if self.ropts["row0"] < H else 0
        self.rows_max = self.my_rows
        if dist is not None:
            t = torch.tensor([self.my_rows], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            self.rows_max = int(t.item())
        self.hdr = torch.empty((max(self.my_rows, 1) * W, 4), dtype=torch.float32, device=dev)
        self.rgb_local = torch.zeros((self.rows_max * W * 3,), dtype=torch.uint8, device=dev)
        self.rgb_all = torch.empty((world * self.rows_max * W * 3,), dtype=torch.uint8, device=dev) if world > 1 else None
        self.auto_white = name == "c5" and args.auto_white
        self.gather = "none" if dist is None else ("nccl" if self.auto_white else args.gather)
        self.use_flag_fence, self.peer_note, self.frame_all, self.opened = False, "", 0, []
        if self.gather == "peer":
            self._setup_peers()
        self.rebuild = True
        self.fence = torch.zeros(1, dtype=torch.float32, device=dev)
        self.white = torch.zeros(1, dtype=torch.float32, device=dev)
        self.tm_ptrs = [self.rgb_local.data_ptr() if op == self.tm_op else None for op in range(_abi.TM_COUNT)]

    def _exchange(self, ptr):
        """Every rank's device buffer `ptr` mapped into this process: list of device pointers by rank (CUDA IPC)."""
        torch, dist, _abi = self.torch, self.dist, self.abi
        mine = torch.frombuffer(bytearray(self.eng.ipc_export(ptr)), dtype=torch.uint8).to(self.dev)
        handles = torch.empty(self.world * _abi.IPC_HANDLE_BYTES, dtype=torch.uint8, device=self.dev)
        dist.all_gather_into_tensor(handles, mine)
        handles = handles.cpu().numpy().reshape(self.world, _abi.IPC_HANDLE_BYTES)
        out = []
        for r in range(self.world):
            if r == self.rank:
                out.append(ptr)
            else:
                self.opened.append(self.eng.ipc_open(handles[r].tobytes()))
                out.append(self.opened[-1])
        return out

    def _setup_peers(self):
        # "peer": every rank's render call stores its RGB8 stripes straight into the full-frame buffer of EVERY rank over NVLink
        # (rte_set_rgb8_peers; buffers exchanged as CUDA IPC handles); the frame is fenced by flags written over NVLink by the
        # same call (rte_frame_fence) or by a 4-byte NCCL all-reduce
        torch, dist, _abi = self.torch, self.dist, self.abi
        try:
            self._peer_frames, self._peer_flags = None, None
            self.frame_all = self.eng.malloc(self.W * self.H * 3)
            self._peer_frames = self._exchange(self.frame_all)
            self.eng.set_rgb8_peers(self._peer_frames)
            if self.args.fence == "flags":
                self._peer_flags = self._exchange(self.eng.malloc(4 * _abi.MAX_PEERS))
                self.eng.set_fence_peers(self.rank, self._peer_flags)
                self.use_flag_fence = True
            ok = torch.ones(1, device=self.dev)
        except Exception as exc:  # no peer mapping on this box: say so and use the NCCL all-gather
            self.peer_note = f" (peer frames unavailable: {exc})"
            ok = torch.zeros(1, device=self.dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if ok.item() < 1:
            self.eng.set_rgb8_peers([])
            if self.use_flag_fence:
                self.eng.set_fence_peers(self.rank, [])
            self.use_flag_fence = False
            self.gather = "nccl"
        dist.barrier()  # every rank has zeroed its flags before anybody renders

    def step(self, flags=0):
        eng, dist, W = self.eng, self.dist, self.W
        if self.rebuild:
            # every frame is a new frame: rte_set_camera invalidates the per-frame structures (direction grids / views), so their
            # build launches are inside every timed step (a static scene rendered again would reuse them: "static_scene" leg)
            eng.set_camera(self.cam)
        if self.auto_white:
            # C5's auto max-white pipeline: HDR stripes -> per-rank max luminance (device float) ->
            # NCCL all-reduce MAX -> Reinhard-extended-luminance with that white point -> NCCL all-gather of RGB8
            eng.render_device({"hdr": self.hdr.data_ptr()}, flags=flags, **self.ropts)
            eng.max_luminance_device(self.hdr.data_ptr(), self.my_rows * W, self.white.data_ptr())
            if dist is not None:
                dist.all_reduce(self.white, op=dist.ReduceOp.MAX)
            eng.tonemap_auto_white(self.hdr.data_ptr(), self.my_rows * W, self.tm_ptrs, self.white.data_ptr(), ops=[self.tm_op])
        elif self.gather == "peer":
            eng.render_device({"hdr": self.hdr.data_ptr()}, tonemap_op=self.tm_op, flags=flags, **self.ropts)  # RGB8 goes to every rank's frame
            if self.use_flag_fence:
                eng.frame_fence()            # device-side wait for every rank's "frame delivered" flag (written over NVLink)
            else:
                dist.all_reduce(self.fence)  # frame fence across ranks through NCCL
            return
        else:
            eng.render_device({"hdr": self.hdr.data_ptr(), "rgb8": self.rgb_local.data_ptr()}, tonemap_op=self.tm_op, flags=flags, **self.ropts)
        if dist is not None:  # the RGB8 stripes of every rank, into the PPM writer's buffer (de-interleaved at write time)
            dist.all_gather_into_tensor(self.rgb_all, self.rgb_local)

    def frame_check(self):
        """Untimed: (1) the frame of the EXACT path (RTE_FLAG_EXACT_FP64, fixed max_white) assembled from all ranks has the CRC-32
        of the reference build's frame (tests/golden/frame_crc.json); (2) the production path's frame is within 1 LSB of it;
        (3) N>1: what the peer stores assemble is what an all-gather of the same stripes assembles."""
        import zlib
        np, torch, dist, eng, _abi = self.np, self.torch, self.dist, self.eng, self.abi
        from raytracingengine_b200 import sharding
        W, H = self.W, self.H
        res = {"frame_ok": None, "frame_lsb_max": None}

        def assembled(flags):
            eng.render_device({"rgb8": self.rgb_local.data_ptr()}, tonemap_op=self.tm_op, flags=flags, **self.ropts)
            if dist is None:
                torch.cuda.synchronize()
                return self.rgb_local[:self.my_rows * W * 3].cpu().numpy().reshape(-1, 3)
            dist.all_gather_into_tensor(self.rgb_all, self.rgb_local)
            torch.cuda.synchronize()
            if self.rank != 0:
                return None
            return sharding.deinterleave(self.rgb_all.cpu().numpy().reshape(-1, 3), self.world, H, W, self.stripe, self.rows_max)

        if self.gather == "peer":
            eng.set_rgb8_peers([])  # the checks below want each rank's stripes in its own buffer
            if self.use_flag_fence:
                eng.set_fence_peers(self.rank, [])
        exact = assembled(_abi.FLAG_EXACT_FP64)
        prod = assembled(0)
        if self.gather == "peer":
            self._reregister()
            self.step()
            torch.cuda.synchronize()
            dist.barrier()
        if self.rank == 0:
            gold = frame_crc_golden(self.name) if self.args.scale == 1.0 else None
            if gold and (gold["width"], gold["height"]) == (W, H):
                res["frame_ok"] = bool(zlib.crc32(exact.tobytes()) == gold["crc32"])
                res["frame_crc32"] = zlib.crc32(exact.tobytes())
            res["frame_lsb_max"] = int(np.abs(prod.astype(np.int16) - exact.astype(np.int16)).max())
            if self.gather == "peer":
                got = np.zeros((H * W, 3), np.uint8)
                eng.d2h(got, self.frame_all)
                res["peer_frame_equals_gather"] = bool(np.array_equal(got, prod))
        return res

    def _reregister(self):
        """Peer frames / flags again after frame_check() or the e2e leg unregistered them (pointers are still mapped)."""
        self.eng.set_rgb8_peers(self._peer_frames)
        if self._peer_flags:
            self.eng.set_fence_peers(self.rank, self._peer_flags)
            self.dist.barrier()

    def close(self):
        if self.opened:
            self.eng.set_rgb8_peers([])
            if self.use_flag_fence:
                self.eng.set_fence_peers(self.rank, [])
            for p in self.opened:
                self.eng.ipc_close(p)
        self.eng.close()


def measure(pipe, steps, warmup, flush, stream):
    """W warm-ups, then K timed steps: CUDA events per step on the launching stream (L2 flushed between steps, outside the
    event pair), barrier + synchronize on both sides, max over ranks."""
    torch, dist, eng = pipe.torch, pipe.dist, pipe.eng

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(warmup, 3)):
        flush.zero_()
        pipe.step()
    sync_all()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    kernel_ms = []
    wall0 = time.perf_counter()
    for i in range(steps):
        flush.zero_()
        ev[i][0].record(stream)
        pipe.step()
        ev[i][1].record(stream)
        if pipe.world == 1:
            kernel_ms.append(eng.stats()["kernel_ms"])  # (synchronises) device time of the library's own launches of this step
    sync_all()
    wall = time.perf_counter() - wall0
    launches = eng.last_render_launches()
    st = eng.stats()
    if pipe.world > 1:
        kernel_ms = [st["kernel_ms"]]
    my_ms = sum(a.elapsed_time(b) for a, b in ev)
    k_mean = sum(kernel_ms) / len(kernel_ms)
    counters = torch.tensor([st["primary_rays"] + st["shadow_rays"], st["sphere_tests"], st["plane_tests"], st["triangle_tests"],
                             st["fp64_confirms"], st["exact_marches"], st["sphere_tests_executed"]], dtype=torch.float64, device=pipe.dev)
    tmax = torch.tensor([my_ms, k_mean, -k_mean], dtype=torch.float64, device=pipe.dev)
    if dist is not None:
        dist.all_reduce(counters, op=dist.ReduceOp.SUM)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    rays, sph_t, pl_t, tri_t, confirms, marches, sph_exec = [float(x) for x in counters.tolist()]
    total_ms, k_max, k_min_neg = [float(x) for x in tmax.tolist()]
    ms = total_ms / steps
    return {"ms_per_step": ms, "value": rays / (ms * 1e-3) / 1e6, "rays": rays, "tests": (sph_t, pl_t, tri_t), "fixup_pixels": confirms,
            "marches": marches, "candidates": sph_exec, "kernel_ms_max": k_max, "kernel_ms_min": -k_min_neg, "launches": launches,
            "wall_ms": wall / steps * 1e3}


def breakdown(pipe, m, steps, flush, stream):
    """N > 1 (peer frames): where a step's time goes.  The same band rendered WITHOUT registered peer frames is the render kernel
    alone (RGB8 into the local band buffer, no exchange launch, no fence); the rest follows by difference.  Untimed legs."""
    if pipe.world == 1 or pipe.gather != "peer":
        return None
    torch, dist, eng = pipe.torch, pipe.dist, pipe.eng
    eng.set_rgb8_peers([])
    if pipe.use_flag_fence:
        eng.set_fence_peers(pipe.rank, [])
    ks, ts = [], []
    for i in range(3 + steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        eng.set_camera(pipe.cam)
        eng.render_device({"hdr": pipe.hdr.data_ptr(), "rgb8": pipe.rgb_local.data_ptr()}, tonemap_op=pipe.tm_op, **pipe.ropts)
        b.record(stream)
        k = eng.stats()["kernel_ms"]
        if i >= 3:
            ks.append(k); ts.append(a.elapsed_time(b))
    t = torch.tensor([sum(ks) / len(ks), sum(ts) / len(ts)], dtype=torch.float64, device=pipe.dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    pipe._reregister()
    render_ms, local_ms = [float(x) for x in t.tolist()]
    return {"render_kernel_ms": round(render_ms, 4), "grid_build_ms": round(local_ms - render_ms, 4),
            "exchange_launch_ms": round(m["kernel_ms_max"] - render_ms, 4),
            "fence_and_rank_skew_ms": round(m["ms_per_step"] - m["kernel_ms_max"] - (local_ms - render_ms), 4),
            "note": "max over ranks; render_kernel_ms: the band without peer frames; exchange = (render + exchange launches) - render; "
                    "fence = step - launches - build (waiting for the slowest rank's flags over NVLink)"}


def c1_leg(torch, stream, flush, local_rank, steps):
    """C1 = main()'s scene (5 reflective planes, the box mesh, 2 lights, maxRecursion 10) at 1000x1000: device time of rte_render
    (HDR + RGB8 with ACES) at 1 sample per pixel and at the reference's default 32 (RE/Math.h:94); L2 flushed between frames."""
    import dataclasses
    from raytracingengine_b200 import _abi, scenes
    from raytracingengine_b200.engine import Engine
    out = {"workload": "C1: main()'s scene at 1000x1000 (5 reflective planes, box mesh, 2 lights, maxRecursion 10), ACES"}
    for aa in (1, 32):
        arrays, cam = scenes.config1()
        cam = dataclasses.replace(cam, antiAliasingAmount=aa, jitterSeed=7)
        e = Engine(local_rank)
        e.set_stream(stream.cuda_stream)
        e.upload(arrays, cam)
        n = cam.width * cam.height
        hdr = torch.empty(n * 4, dtype=torch.float32, device="cuda")
        rgb = torch.empty(n * 3 + 16, dtype=torch.uint8, device="cuda")
        k = steps if aa == 1 else max(3, steps // 3)
        ts = []
        for i in range(3 + k):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            e.set_camera(cam)
            e.render_device({"hdr": hdr.data_ptr(), "rgb8": rgb.data_ptr()}, tonemap_op=_abi.TM_ACES)
            b.record(stream)
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b))
        s = e.stats()
        rays = s["primary_rays"] + s["shadow_rays"] + s["secondary_rays"]
        ms = sum(ts) / len(ts)
        out[f"spp{aa}"] = {"ms_per_frame": round(ms, 4), "rays_traced": int(rays), "Mrays_per_s": round(rays / ms / 1e3, 1), "steps": k,
                           "gpu_launches_per_frame": e.last_render_launches(), "pixels_left_to_the_exact_kernel": int(s["fp64_confirms"])}
        e.close()
    out["note"] = ("the reflective FP32 kernel ends a chain of reflections below a path weight of 1e-7, so it traces fewer rays than the "
                   "reference (30.0 M per sample); parity: tests/test_gpu_fullsize.py::test_c1_whole_frame, both arms: tools/c1_bench.py")
    return out


def sharding_text(pipe):
    if pipe.world == 1:
        return "single band"
    s = f"interleaved {pipe.stripe}-row stripes; "
    if pipe.gather == "peer":
        return s + "RGB8 band copied into every rank's frame over NVLink by the exchange launch behind the render kernel (16-byte peer stores, CUDA IPC peer frames) + " + (
            "frame fence through NVLink flags written by the call's last launch (rte_frame_fence), no collective" if pipe.use_flag_fence
            else "4-byte NCCL all-reduce as frame fence")
    return s + "NCCL all-gather of RGB8" + pipe.peer_note


def run_gpu(args) -> None:
    import numpy as np
    import torch

    from raytracingengine_b200 import _abi
    from raytracingengine_b200.engine import PinnedBuffer

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
    stream = torch.cuda.Stream(device=dev)  # a real (non-default) stream: handle 0 would mean "ctx-private" to the library
    torch.cuda.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    pipe = Pipeline(args, args.workload, torch, dist, dev, rank, world, local_rank, stream)
    arrays, cam, W, H, eng = pipe.arrays, pipe.cam, pipe.W, pipe.H, pipe.eng
    check = pipe.frame_check()
    fp32_peak_tf, _ = eng.measure_fp32_peak(15.0)
    sampler = ClockSampler(visible_index(local_rank))
    with sampler:
        m = measure(pipe, args.steps, args.warmup, flush, stream)
        pipe.rebuild = False  # the same frame again without rebuilding the grids (reported beside the headline, never as it)
        m_static = measure(pipe, max(3, args.steps // 2), 3, flush, stream)
        pipe.rebuild = True
        bd = breakdown(pipe, m, max(3, args.steps // 2), flush, stream)

    # ---- e2e: C-ABI with HOST buffers — scene description H2D (rte_upload_scene) + render + RGB8 D2H, wall clock
    pin = PinnedBuffer(max(pipe.my_rows, 1) * W * 3)
    host_rgb = pin.view(np.uint8, (max(pipe.my_rows, 1) * W, 3))
    scene_bytes = sum(a.nbytes for a in (arrays.materials, arrays.spheres, arrays.planes, arrays.triangles, arrays.models,
                                         arrays.model_vertices, arrays.model_indices, arrays.lights))
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    if pipe.gather == "peer":
        eng.set_rgb8_peers([])  # the end-to-end leg returns each rank's stripes to its own host buffer
        if pipe.use_flag_fence:
            eng.set_fence_peers(rank, [])

    def e2e_step():
        eng.upload(arrays, cam)
        eng.render_host(outputs=("rgb8",), into={"rgb8": host_rgb}, tonemap_op=pipe.tm_op, **pipe.ropts)

    for _ in range(3):
        e2e_step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    te = torch.tensor([(time.perf_counter() - e0) * 1e3], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms = float(te.item()) / args.steps
    e2e_value = m["rays"] / (e2e_ms * 1e-3) / 1e6

    # ---- the HBM-bound pass of the path: all seven tonemap operators over the (band) HDR buffer (C3's pass)
    n_px = pipe.my_rows * W
    tm_out = [torch.empty(n_px * 3 + 16, dtype=torch.uint8, device=dev) for _ in range(_abi.TM_COUNT)]
    tm_ms = []
    hdr2 = pipe.hdr.clone()  # two rotating inputs: 2 x 133 MB in + 174 MB out per launch stream through the 126 MB L2 (no flush needed)
    for i in range(3 + 20):
        eng.tonemap_device((pipe.hdr if i % 2 else hdr2).data_ptr(), n_px, [t.data_ptr() for t in tm_out])
        if i >= 3:
            tm_ms.append(eng.stats()["tonemap_ms"])
    tm_med = sorted(tm_ms)[len(tm_ms) // 2]
    del tm_out, hdr2
    pipe.close()

    # ---- the north star's scaling target in the same run: C5 at 7680x4320 (4096 spheres, 8 lights, Reinhard-extended-luminance
    # with the reference's constant max_white = 5.0), fewer steps
    c5 = None
    if args.workload != "c5" and not args.no_c5 and args.scale == 1.0:
        p5 = Pipeline(args, "c5", torch, dist, dev, rank, world, local_rank, stream)
        chk5 = p5.frame_check()
        m5 = measure(p5, max(3, args.steps // 4), 3, flush, stream)
        bd5 = breakdown(p5, m5, 3, flush, stream)
        c5 = {"workload": p5.label, "ms_per_step": round(m5["ms_per_step"], 4), "value": round(m5["value"], 2), "unit": UNIT,
              "rays_per_frame": int(m5["rays"]), "kernel_ms_max": round(m5["kernel_ms_max"], 4), "kernel_ms_min": round(m5["kernel_ms_min"], 4),
              "steps": max(3, args.steps // 4), "sharding": sharding_text(p5), "fixup_pixels": int(m5["fixup_pixels"]), "breakdown": bd5, **chk5}
        p5.close()

    # ---- the reference's own default workload in the same run (BASELINE.json configs[0]): main()'s scene at 1000x1000 with 1 and
    # with the default 32 samples per pixel, one GPU (the frame is 0.7 ms: nothing to shard)
    c1 = None
    if world == 1 and not args.no_c5 and args.scale == 1.0:
        c1 = c1_leg(torch, stream, flush, local_rank, max(3, args.steps // 2))

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    pk, pk_src = peaks()
    clocks = sampler.summary()
    kern, kern_src = ncu_kernels()
    tr_tm = ncu_traffic("tonemap_kernel")
    sph_t, pl_t, tri_t = m["tests"]
    flops = _abi.algorithmic_flops(sph_t, pl_t, tri_t)           # 20 / 14 / 46 flop per test
    achieved_tf = flops / world / (m["kernel_ms_max"] * 1e-3) / 1e12   # per GPU: per-rank flops / slowest rank's device time
    r32 = (kern or {}).get("render32_kernel", {})
    line = {
        "metric": f"Mrays/s (primary+shadow) at {W}x{H}", "value": round(m["value"], 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(m["ms_per_step"], 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 values, decisions from f32 intervals with an exact f64 fix-up", "data": "synthetic",
        "config": {"workload": pipe.label, "width": W, "height": H, **arrays.counts(), "seed": 1234 if args.workload != "c5" else 5678,
                   "rays_per_frame": int(m["rays"]), "sharding": sharding_text(pipe),
                   "tonemap": ("auto max-white: max-luminance kernel + " + ("NCCL all-reduce MAX + " if world > 1 else "") + "tonemap pass") if pipe.auto_white
                   else "operator + toColor inside the render kernel (FP32, +-1 LSB of the reference); exact FP64 for the fix-up pixels",
                   "l2": "256 MiB memset between timed steps; outputs (HDR float4 + RGB8 = 19 B/px) exceed L2",
                   "timing": "CUDA events per step on the launching stream, summed, max over ranks"},
        "ms_per_frame": round(m["ms_per_step"], 4),
        "static_scene": {"ms_per_step": round(m_static["ms_per_step"], 4), "gpu_launches_per_step": m_static["launches"],
                         "note": "same frame, direction grids of the previous frame reused (scene and camera unchanged): NOT the headline"},
        "breakdown": bd,
        "grids": "rebuilt inside every timed step (count + scan + fill launches in front of the render kernel)",
        "wall_ms_per_step_incl_flush": round(m["wall_ms"], 3), "clocks": clocks,
        **check,
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "ms_per_step": round(e2e_ms, 3), "h2d_bytes_per_step": int(scene_bytes),
                "d2h_bytes_per_step": int(pipe.my_rows * W * 3), "api": "rte_upload_scene + rte_render_host (pinned host RGB8)"},
        # per step and rank: the library's launches of one render call (grid count / scan / fill + render kernel, + exchange at N > 1; + max luminance + tonemap for the
        # auto-white pipeline); memsets and NCCL are not ours
        "gpu_launches": args.steps * (m["launches"] + (2 if pipe.auto_white else 0)),
        "roofline": {"bound": "fp32", "kernel": "rte::render32_kernel (incl. fix_batch, the exact FP64 re-render of the flagged pixels, in the same launch)",
                     "achieved": round(achieved_tf, 3), "peak": round(fp32_peak_tf, 3), "unit": "TFLOP/s", "frac": round(achieved_tf / fp32_peak_tf, 4),
                     "frac_note": "ALGORITHMIC: the reference's 20 flop per ray-sphere test over ALL spheres (SURVEY §8(d)) / time / measured FFMA peak; the "
                                  "direction grids test ~3 of 1024 spheres per ray, so this is a speed-up over brute force, not a hardware utilisation — "
                                  "the bounded hardware figures are issue_active_pct / pipe_*_pct below (ncu)",
                     "algorithmic_tflops": round(achieved_tf, 3), "algorithmic_flop_per_launch": flops / world,
                     "traffic": r32.get("dram_bytes"), "kernel_ms": round(m["kernel_ms_max"], 4), "kernel_ms_min_rank": round(m["kernel_ms_min"], 4),
                     "frac_hardware": (round(r32["issue_active_pct"] / 100.0, 4) if r32.get("issue_active_pct") else None),
                     "frac_hardware_note": "bounded figure: share of the SM's issue slots the kernel used (ncu smsp__issue_active, committed capture) — the kernel is issue-bound, not FMA-bound",
                     "issue_active_pct": r32.get("issue_active_pct"), "pipe_fma_pct": r32.get("pipe_fma_pct"), "pipe_fp64_pct": r32.get("pipe_fp64_pct"),
                     "pipe_alu_pct": r32.get("pipe_alu_pct"), "pipe_xu_pct": r32.get("pipe_xu_pct"), "ncu_kernels": kern, "ncu_source": kern_src,
                     "peak_source": "FFMA micro-benchmark in this process (rte_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
                     "candidates_tested": m["candidates"], "candidates_per_ray": round(m["candidates"] / max(1.0, m["rays"]), 3),
                     "fixup_pixels": int(m["fixup_pixels"]), "tests": {"sphere": sph_t, "plane": pl_t, "triangle": tri_t}, "exact_marches": m["marches"]},
        "roofline_tonemap": {"bound": "hbm", "kernel": "rte::tonemap_kernel<all 7 operators>", "achieved": round(n_px * 33 / (tm_med * 1e-3) / 1e9, 1),
                             "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(n_px * 33 / (tm_med * 1e-3) / 1e9 / pk["hbm_gbs"], 4),
                             "traffic": (tr_tm or {}).get("bytes") if args.workload in ("c2", "c3") and world == 1 else None, "traffic_capture": tr_tm,
                             "peak_source": pk_src, "algorithmic_bytes_per_launch": n_px * 33, "kernel_ms": round(tm_med, 4),
                             "frac_moved_bytes": round(n_px * 37 / (tm_med * 1e-3) / 1e9 / pk["hbm_gbs"], 4),
                             "frac_note": "frac counts SURVEY's 33 B/px (12 B of HDR read per pixel); the ABI's HDR is float4, so 37 B/px really move "
                                          "(frac_moved_bytes). The same loads and stores WITHOUT the operators' arithmetic take 64-68 us "
                                          "(tonemap_null_kernel, tools/tm_probe.py): the pass sits on the bandwidth of its access pattern",
                             "l2": "input (16 B/px float4, two rotating buffers) and output (21 B/px) exceed L2; no flush"},
    }
    if c5 is not None:
        line["c5_8k"] = c5
    if c1 is not None:
        line["c1_default_scene"] = c1
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
    ap.add_argument("--auto-white", action="store_true", help="--workload c5: the auto max-white pipeline (max-luminance kernel + NCCL all-reduce MAX + tonemap pass; no reference counterpart, SURVEY D6) instead of the reference's constant max_white = 5.0")
    ap.add_argument("--no-c5", action="store_true", help="skip the c5_8k leg (C5 at 7680x4320, the north star's scaling target)")
    args = ap.parse_args()
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
