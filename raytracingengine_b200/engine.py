"""ctypes binding of librte_b200.so (the C ABI of include/rte.h) — the only way this package computes.

There is deliberately no fallback: if the CUDA library is missing or no B200 is visible, construction
raises.  Everything below is marshalling; intersection, shading and tonemapping live in
raytracingengine_b200/csrc/*.cu.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Iterable, Optional, Sequence

import numpy as np

from . import _abi
from .scene import Camera, SceneArrays

LIB_PATH = os.environ.get("RTE_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "librte_b200.so")  # env: developer A/B builds

# every symbol include/rte.h declares (tests/test_abi.py checks the header against this list)
SYMBOLS = (
    "rte_version", "rte_device_count", "rte_last_error", "rte_create", "rte_destroy", "rte_set_stream",
    "rte_synchronize", "rte_upload_scene", "rte_set_camera", "rte_render", "rte_render_host", "rte_get_stats", "rte_last_render_launches",
    "rte_tonemap", "rte_tonemap_host", "rte_tonemap_exact_host", "rte_tonemap_values_host", "rte_tonemap_auto_white", "rte_max_luminance", "rte_max_luminance_device", "rte_max_luminance_noalloc", "rte_measure_fp32_peak", "rte_measure_fp32x2_peak", "rte_intersect_closest", "rte_intersect_any_before",
    "rte_malloc", "rte_free", "rte_memcpy_d2h", "rte_memcpy_h2d", "rte_malloc_host", "rte_free_host",
    "rte_write_ppm", "rte_write_png", "rte_set_rgb8_peers", "rte_set_fence_peers", "rte_frame_fence", "rte_ipc_export", "rte_ipc_open", "rte_ipc_close",
)

_lib = None


class RteError(RuntimeError):
    """Non-zero rte_status; the reference's convention is std::runtime_error (RE/Image.cpp:15,27)."""

    def __init__(self, status: int, message: str):
        super().__init__(f"[rte status {status}] {message}")
        self.status = status


def load_library() -> C.CDLL:
    """dlopen the in-tree library; fails loudly when it has not been built (no CPU path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C raytracingengine_b200/csrc` "
                          "(or __graft_entry__.build()); this package has no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    lib.rte_last_error.restype = C.c_char_p
    lib.rte_last_error.argtypes = [C.c_void_p]
    lib.rte_destroy.restype = None
    lib.rte_destroy.argtypes = [C.c_void_p]
    lib.rte_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.rte_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    lib.rte_synchronize.argtypes = [C.c_void_p]
    lib.rte_upload_scene.argtypes = [C.c_void_p, C.POINTER(_abi.SceneDesc)]
    lib.rte_set_camera.argtypes = [C.c_void_p, C.POINTER(_abi.Camera)]
    lib.rte_render.argtypes = [C.c_void_p, C.POINTER(_abi.RenderOpts), C.POINTER(_abi.Frame)]
    lib.rte_render_host.argtypes = [C.c_void_p, C.POINTER(_abi.RenderOpts), C.POINTER(_abi.Frame)]
    lib.rte_get_stats.argtypes = [C.c_void_p, C.POINTER(_abi.Stats)]
    lib.rte_last_render_launches.argtypes = [C.c_void_p]
    lib.rte_last_render_launches.restype = C.c_int
    tm_args = [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint32, C.c_double, C.c_double, C.POINTER(C.c_void_p)]
    lib.rte_tonemap.argtypes = tm_args
    lib.rte_tonemap_host.argtypes = tm_args
    lib.rte_tonemap_exact_host.argtypes = tm_args
    lib.rte_tonemap_values_host.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_int64, C.c_void_p]
    lib.rte_max_luminance.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_float)]
    lib.rte_max_luminance_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    lib.rte_max_luminance_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    lib.rte_tonemap_auto_white.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint32, C.c_void_p, C.c_double, C.POINTER(C.c_void_p)]
    lib.rte_measure_fp32_peak.argtypes = [C.c_void_p, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.rte_measure_fp32x2_peak.argtypes = [C.c_void_p, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.rte_intersect_closest.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.rte_intersect_any_before.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.rte_malloc.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
    lib.rte_set_rgb8_peers.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
    lib.rte_set_fence_peers.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    lib.rte_frame_fence.argtypes = [C.c_void_p, C.c_double]
    lib.rte_ipc_export.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p]
    lib.rte_ipc_open.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
    lib.rte_ipc_close.argtypes = [C.c_void_p, C.c_void_p]
    lib.rte_free.argtypes = [C.c_void_p, C.c_void_p]
    lib.rte_memcpy_d2h.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    lib.rte_memcpy_h2d.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    lib.rte_malloc_host.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    lib.rte_free_host.argtypes = [C.c_void_p]
    lib.rte_write_ppm.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint64]
    lib.rte_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint64]
    _lib = lib
    return lib


def _check(lib, rc: int, ctx=None) -> None:
    if rc != 0:
        msg = lib.rte_last_error(ctx)
        raise RteError(rc, msg.decode() if msg else "unknown error")


def write_ppm(path: str, rgb8: np.ndarray, width: int, height: int) -> None:
    """writePPM(filename, pixels, width, height), RE/Image.cpp:11-31 (needs no device)."""
    lib = load_library()
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    if a.size != width * height * 3:
        raise ValueError("rgb8 must hold width*height RGB triplets")
    _check(lib, lib.rte_write_ppm(os.fsencode(path), C.c_void_p(a.ctypes.data), width, height))


def write_png(path: str, rgb8: np.ndarray, width: int, height: int) -> None:
    """The .png main() obtains through ffmpeg (RE/RaytracingEngine.cpp:317-327), written directly (needs no device)."""
    lib = load_library()
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    if a.size != width * height * 3:
        raise ValueError("rgb8 must hold width*height RGB triplets")
    _check(lib, lib.rte_write_png(os.fsencode(path), C.c_void_p(a.ctypes.data), width, height))


class PinnedBuffer:
    """Page-locked host memory from rte_malloc_host, viewed as a numpy array."""

    def __init__(self, nbytes: int):
        self.lib = load_library()
        self.ptr = C.c_void_p()
        _check(self.lib, self.lib.rte_malloc_host(max(nbytes, 1), C.byref(self.ptr)))
        self.nbytes = nbytes
        self._raw = (C.c_uint8 * max(nbytes, 1)).from_address(self.ptr.value)

    def view(self, dtype, shape) -> np.ndarray:
        return np.frombuffer(self._raw, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def close(self):
        if self.ptr:
            self.lib.rte_free_host(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# name -> (numpy dtype, trailing shape, bytes per pixel) of every rte_frame member
FRAME_FIELDS = {
    "hdr": (np.float32, (4,)), "hdr64": (np.float64, (3,)), "rgb8": (np.uint8, (3,)),
    "prim_id": (np.uint32, ()), "depth": (np.float32, ()), "depth64": (np.float64, ()),
    "normal": (np.float32, (4,)), "albedo": (np.float32, (4,)), "light_mask": (np.uint32, ()), "depth_norm": (np.float32, ()),
}


class Engine:
    """One rte_ctx (one GPU)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.ctx = C.c_void_p()
        _check(self.lib, self.lib.rte_create(device, C.byref(self.ctx)))
        self.device = device
        self.camera: Optional[Camera] = None
        self._arrays: Optional[SceneArrays] = None
        self._dev_allocs = []

    # ---- lifetime
    def close(self):
        if getattr(self, "ctx", None):
            for p in self._dev_allocs:
                self.lib.rte_free(self.ctx, p)
            self._dev_allocs = []
            self.lib.rte_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _ck(self, rc):
        _check(self.lib, rc, self.ctx)

    # ---- scene
    def upload(self, arrays: SceneArrays, camera: Optional[Camera] = None) -> None:
        desc = arrays.desc()
        self._ck(self.lib.rte_upload_scene(self.ctx, C.byref(desc)))
        self._arrays = arrays
        if camera is not None:
            self.set_camera(camera)

    def set_camera(self, camera: Camera) -> None:
        c = camera.to_c()
        self._ck(self.lib.rte_set_camera(self.ctx, C.byref(c)))
        self.camera = camera

    def set_stream(self, cuda_stream: int) -> None:
        self._ck(self.lib.rte_set_stream(self.ctx, C.c_void_p(cuda_stream)))

    def synchronize(self) -> None:
        self._ck(self.lib.rte_synchronize(self.ctx))

    # ---- device memory
    def malloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._ck(self.lib.rte_malloc(self.ctx, nbytes, C.byref(p)))
        self._dev_allocs.append(p)
        return p.value

    # ---- peer frames (multi-GPU output fused into the render kernel)
    def set_rgb8_peers(self, frames: Sequence[int]) -> None:
        arr = (C.c_void_p * max(1, len(frames)))(*[C.c_void_p(p) for p in frames])
        self._ck(self.lib.rte_set_rgb8_peers(self.ctx, len(frames), arr))

    def set_fence_peers(self, rank: int, flags: Sequence[int]) -> None:
        """rte_set_fence_peers: flags[r] = device pointer of rank r's flag array (RTE_MAX_PEERS uint32); [] unregisters."""
        arr = (C.c_void_p * max(1, len(flags)))(*[C.c_void_p(p) for p in flags])
        self._ck(self.lib.rte_set_fence_peers(self.ctx, int(rank), len(flags), arr))

    def frame_fence(self, timeout_ms: float = 0.0) -> None:
        """rte_frame_fence: device-side wait (on the context's stream) until every rank delivered the last rendered frame."""
        self._ck(self.lib.rte_frame_fence(self.ctx, float(timeout_ms)))

    def ipc_export(self, dev: int) -> bytes:
        buf = C.create_string_buffer(_abi.IPC_HANDLE_BYTES)
        self._ck(self.lib.rte_ipc_export(self.ctx, C.c_void_p(dev), buf))
        return buf.raw

    def ipc_open(self, handle: bytes) -> int:
        p = C.c_void_p()
        self._ck(self.lib.rte_ipc_open(self.ctx, C.c_char_p(bytes(handle)), C.byref(p)))
        return p.value

    def ipc_close(self, dev: int) -> None:
        self._ck(self.lib.rte_ipc_close(self.ctx, C.c_void_p(dev)))

    def d2h(self, host: np.ndarray, dev: int, nbytes: Optional[int] = None) -> None:
        self._ck(self.lib.rte_memcpy_d2h(self.ctx, C.c_void_p(host.ctypes.data), C.c_void_p(dev), host.nbytes if nbytes is None else nbytes))

    def h2d(self, dev: int, host: np.ndarray) -> None:
        host = np.ascontiguousarray(host)
        self._ck(self.lib.rte_memcpy_h2d(self.ctx, C.c_void_p(dev), C.c_void_p(host.ctypes.data), host.nbytes))

    # ---- render
    @staticmethod
    def _opts(row0=0, row1=0, max_depth=0, tonemap_op=_abi.TM_NONE, bias=0.0, max_white=0.0, jodie_a=0.0, flags=0,
              stripe_rows=0, stripe_period=0):
        return _abi.RenderOpts(row0, row1, max_depth, tonemap_op, bias, max_white, jodie_a, 0, flags, stripe_rows, stripe_period)

    @staticmethod
    def supports_stripes() -> bool:
        return True

    @staticmethod
    def striped_rows(row0: int, row1: int, stripe_rows: int, stripe_period: int) -> int:
        """Rows a striped call renders (mirrors resolve_opts in rte_api.cu)."""
        if stripe_rows <= 0:
            return row1 - row0
        span = row1 - row0
        return (span // stripe_period) * stripe_rows + min(span % stripe_period, stripe_rows)

    def band_pixels(self, row0=0, row1=0, stripe_rows=0, stripe_period=0) -> int:
        H, W = self.camera.height, self.camera.width
        return self.striped_rows(row0, row1 if row1 > 0 else H, stripe_rows, stripe_period) * W

    def render_device(self, frame_ptrs: Dict[str, int], **opts) -> None:
        """rte_render: asynchronous, `frame_ptrs` maps rte_frame member names to device pointers."""
        o = self._opts(**opts)
        f = _abi.Frame()
        for k, v in frame_ptrs.items():
            setattr(f, k, v)
        self._ck(self.lib.rte_render(self.ctx, C.byref(o), C.byref(f)))

    def render_host(self, outputs: Iterable[str] = ("hdr",), into: Optional[Dict[str, np.ndarray]] = None, **opts) -> Dict[str, np.ndarray]:
        """rte_render_host: renders and copies the requested rte_frame members to host arrays."""
        n = self.band_pixels(opts.get("row0", 0), opts.get("row1", 0), opts.get("stripe_rows", 0), opts.get("stripe_period", 0))
        o = self._opts(**opts)
        f = _abi.Frame()
        res = {}
        for name in outputs:
            dt, tail = FRAME_FIELDS[name]
            a = into[name] if into and name in into else np.empty((n,) + tail, dtype=dt)
            assert a.dtype == dt and a.size == n * int(np.prod(tail, dtype=np.int64)) and a.flags.c_contiguous
            res[name] = a
            setattr(f, name, a.ctypes.data)
        self._ck(self.lib.rte_render_host(self.ctx, C.byref(o), C.byref(f)))
        return res

    def stats(self) -> dict:
        s = _abi.Stats()
        self._ck(self.lib.rte_get_stats(self.ctx, C.byref(s)))
        return s.as_dict()

    def last_render_launches(self) -> int:
        """rte_last_render_launches: kernels the last render call enqueued."""
        return int(self.lib.rte_last_render_launches(self.ctx))

    # ---- tonemap
    @staticmethod
    def _mask(ops: Optional[Sequence[int]]) -> int:
        if ops is None:
            return (1 << _abi.TM_COUNT) - 1
        m = 0
        for op in ops:
            m |= 1 << int(op)
        return m

    def tonemap_device(self, hdr_dev: int, n_pixels: int, rgb8_dev: Sequence[Optional[int]], ops=None, max_white=0.0, jodie_a=0.0) -> None:
        ptrs = (C.c_void_p * _abi.TM_COUNT)(*[C.c_void_p(p) if p else None for p in rgb8_dev])
        self._ck(self.lib.rte_tonemap(self.ctx, C.c_void_p(hdr_dev), n_pixels, self._mask(ops), max_white, jodie_a, ptrs))

    def tonemap_host(self, hdr: np.ndarray, ops=None, max_white=0.0, jodie_a=0.0) -> Dict[int, np.ndarray]:
        """tonemapAll / tonemap over a host float4 (n,4) float32 HDR buffer; returns {op: (n,3) uint8}."""
        h = np.ascontiguousarray(hdr, dtype=np.float32).reshape(-1, 4)
        n = len(h)
        mask = self._mask(ops)
        out = {op: np.empty((n, 3), dtype=np.uint8) for op in range(_abi.TM_COUNT) if mask & (1 << op)}
        ptrs = (C.c_void_p * _abi.TM_COUNT)(*[C.c_void_p(out[op].ctypes.data) if op in out else None for op in range(_abi.TM_COUNT)])
        self._ck(self.lib.rte_tonemap_host(self.ctx, C.c_void_p(h.ctypes.data), n, mask, max_white, jodie_a, ptrs))
        return out

    def tonemap_exact_host(self, hdr64: np.ndarray, ops=None, max_white=0.0, jodie_a=0.0) -> Dict[int, np.ndarray]:
        """tonemapAll / tonemap over the FP64 HDR (n,3) float64 — the exact FP64 operators: the reference's own bytes."""
        h = np.ascontiguousarray(hdr64, dtype=np.float64).reshape(-1, 3)
        n = len(h)
        mask = self._mask(ops)
        out = {op: np.empty((n, 3), dtype=np.uint8) for op in range(_abi.TM_COUNT) if mask & (1 << op)}
        ptrs = (C.c_void_p * _abi.TM_COUNT)(*[C.c_void_p(out[op].ctypes.data) if op in out else None for op in range(_abi.TM_COUNT)])
        self._ck(self.lib.rte_tonemap_exact_host(self.ctx, C.c_void_p(h.ctypes.data), n, mask, max_white, jodie_a, ptrs))
        return out

    def tonemap_values_host(self, op: int, colors: np.ndarray, max_white=0.0, jodie_a=0.0) -> np.ndarray:
        """One operator function (simple ... aces_approx) on (n,3) float64 colours -> (n,3) float64."""
        c = np.ascontiguousarray(colors, dtype=np.float64).reshape(-1, 3)
        out = np.empty_like(c)
        self._ck(self.lib.rte_tonemap_values_host(self.ctx, int(op), max_white, jodie_a, C.c_void_p(c.ctypes.data), len(c), C.c_void_p(out.ctypes.data)))
        return out

    def max_luminance(self, hdr_dev: int, n_pixels: int) -> float:
        v = C.c_float()
        self._ck(self.lib.rte_max_luminance(self.ctx, C.c_void_p(hdr_dev), n_pixels, C.byref(v)))
        return float(v.value)

    def max_luminance_device(self, hdr_dev: int, n_pixels: int, out_dev: int) -> None:
        """Stream-ordered: max luminance into a device float (the value a multi-GPU run all-reduces with MAX)."""
        self._ck(self.lib.rte_max_luminance_device(self.ctx, C.c_void_p(hdr_dev), n_pixels, C.c_void_p(out_dev)))

    def max_luminance_noalloc(self, hdr_dev: int, n_pixels: int, out_dev: int) -> None:
        """rte_max_luminance_noalloc: the same value with the scratch in a static device array (no allocation in the call)."""
        self._ck(self.lib.rte_max_luminance_noalloc(self.ctx, C.c_void_p(hdr_dev), n_pixels, C.c_void_p(out_dev)))

    def tonemap_auto_white(self, hdr_dev: int, n_pixels: int, rgb8_dev: Sequence[Optional[int]], max_white_dev: int, ops=None, jodie_a=0.0) -> None:
        """rte_tonemap with max_white read from a device float (auto max-white, after the all-reduce)."""
        ptrs = (C.c_void_p * _abi.TM_COUNT)(*[C.c_void_p(p) if p else None for p in rgb8_dev])
        self._ck(self.lib.rte_tonemap_auto_white(self.ctx, C.c_void_p(hdr_dev), n_pixels, self._mask(ops), C.c_void_p(max_white_dev), jodie_a, ptrs))

    def measure_fp32_peak(self, target_ms: float = 20.0, packed: bool = False):
        """(TFLOP/s, ms) of a register-resident FFMA (or packed FFMA2) kernel on this GPU: the FP32 roofline denominator."""
        tf, ms = C.c_double(), C.c_double()
        fn = self.lib.rte_measure_fp32x2_peak if packed else self.lib.rte_measure_fp32_peak
        self._ck(fn(self.ctx, target_ms, C.byref(tf), C.byref(ms)))
        return float(tf.value), float(ms.value)

    # ---- ray queries
    def intersect_closest(self, origins, directions) -> np.ndarray:
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        hits = np.zeros(len(o), dtype=_abi.hit_dt)
        self._ck(self.lib.rte_intersect_closest(self.ctx, len(o), C.c_void_p(o.ctypes.data), C.c_void_p(d.ctypes.data), C.c_void_p(hits.ctypes.data)))
        return hits

    def intersect_any_before(self, origins, directions, max_dist) -> np.ndarray:
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        m = np.ascontiguousarray(max_dist, dtype=np.float64).reshape(-1)
        out = np.zeros(len(o), dtype=np.uint8)
        self._ck(self.lib.rte_intersect_any_before(self.ctx, len(o), C.c_void_p(o.ctypes.data), C.c_void_p(d.ctypes.data),
                                                   C.c_void_p(m.ctypes.data), C.c_void_p(out.ctypes.data)))
        return out
