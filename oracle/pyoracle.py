"""TEST INFRASTRUCTURE ONLY — ctypes loader for the two CPU checkers (see oracle/oracle_abi.h).

Allowed importers: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline and --impl reference).
The product package raytracingengine_b200 never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

from raytracingengine_b200 import _abi
from raytracingengine_b200.scene import Camera, SceneArrays

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libref_oracle.so")
PORT_LIB = os.path.join(HERE, "liboracle_port.so")


class OrcOpts(C.Structure):
    _fields_ = [("row0", C.c_int32), ("row1", C.c_int32), ("row_step", C.c_int32), ("threads", C.c_int32),
                ("max_depth", C.c_int32), ("want_aovs", C.c_int32), ("bias", C.c_double)]


class OrcOutputs(C.Structure):
    _fields_ = [("hdr", C.c_void_p), ("prim_id", C.c_void_p), ("depth", C.c_void_p), ("normal", C.c_void_p),
                ("albedo", C.c_void_p), ("light_mask", C.c_void_p), ("tie_mask", C.c_void_p),
                ("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("secondary_rays", C.c_uint64),
                ("sphere_tests", C.c_uint64), ("plane_tests", C.c_uint64), ("triangle_tests", C.c_uint64),
                ("closest_queries", C.c_uint64), ("render_seconds", C.c_double),
                ("threads_used", C.c_int32), ("reserved", C.c_int32)]


def build(which: str = "all") -> None:
    """Compile the checkers (make -C oracle); the reference one only where /root/reference exists."""
    subprocess.run(["make", "-C", HERE, which], check=True, capture_output=True)


class Oracle:
    """One loaded checker library; `prefix` is 'ref_' or 'port_'."""

    def __init__(self, path: str, prefix: str):
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.path = path
        f = self._fn("kind")
        f.restype = C.c_char_p
        self.kind = f().decode()

    def _fn(self, name):
        return getattr(self.lib, self.prefix + name)

    def render(self, arrays: SceneArrays, camera: Camera, row0=0, row1=0, row_step=1, threads=0,
               want_aovs=False, hdr=True, max_depth=0, bias=0.0) -> dict:
        W, H = camera.width, camera.height
        row1 = row1 if row1 > 0 else H
        nrows = (row1 - row0 + row_step - 1) // row_step
        n = nrows * W
        out = OrcOutputs()
        res = {}
        if hdr:
            res["hdr"] = np.zeros((n, 3), dtype=np.float64)
            out.hdr = res["hdr"].ctypes.data
        if want_aovs:
            res["prim_id"] = np.zeros(n, dtype=np.uint32)
            res["depth"] = np.zeros(n, dtype=np.float64)
            res["normal"] = np.zeros((n, 3), dtype=np.float64)
            res["albedo"] = np.zeros((n, 3), dtype=np.float64)
            res["light_mask"] = np.zeros(n, dtype=np.uint32)
            res["tie_mask"] = np.zeros(n, dtype=np.uint32)
            for k in ("prim_id", "depth", "normal", "albedo", "light_mask", "tie_mask"):
                setattr(out, k, res[k].ctypes.data)
        opts = OrcOpts(row0, row1, row_step, threads, max_depth, 1 if want_aovs else 0, bias)
        desc = arrays.desc()
        cam = camera.to_c()
        rc = self._fn("render")(C.byref(desc), C.byref(cam), C.byref(opts), C.byref(out))
        if rc != 0:
            raise RuntimeError(f"{self.prefix}render failed: {rc}")
        for k in ("primary_rays", "shadow_rays", "secondary_rays", "sphere_tests", "plane_tests", "triangle_tests",
                  "closest_queries", "render_seconds", "threads_used"):
            res[k] = getattr(out, k)
        res["rows"] = nrows
        return res

    def intersect_closest(self, arrays: SceneArrays, origins, directions) -> np.ndarray:
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        hits = np.zeros(len(o), dtype=_abi.hit_dt)
        desc = arrays.desc()
        rc = self._fn("intersect_closest")(C.byref(desc), C.c_int64(len(o)), C.c_void_p(o.ctypes.data),
                                           C.c_void_p(d.ctypes.data), C.c_void_p(hits.ctypes.data))
        assert rc == 0
        return hits

    def intersect_any_before(self, arrays: SceneArrays, origins, directions, max_dist) -> np.ndarray:
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        m = np.ascontiguousarray(max_dist, dtype=np.float64).reshape(-1)
        out = np.zeros(len(o), dtype=np.uint8)
        desc = arrays.desc()
        rc = self._fn("intersect_any_before")(C.byref(desc), C.c_int64(len(o)), C.c_void_p(o.ctypes.data),
                                              C.c_void_p(d.ctypes.data), C.c_void_p(m.ctypes.data),
                                              C.c_void_p(out.ctypes.data))
        assert rc == 0
        return out

    def camera_ray(self, camera: Camera, px: int, py: int):
        o = (C.c_double * 3)()
        d = (C.c_double * 3)()
        cam = camera.to_c()
        f = self._fn("camera_ray")
        f.restype = None
        f(C.byref(cam), C.c_uint64(px), C.c_uint64(py), o, d)
        return np.array(o[:]), np.array(d[:])

    def tonemap_all(self, hdr, want_float=False):
        h = np.ascontiguousarray(hdr, dtype=np.float64).reshape(-1, 3)
        n = len(h)
        rgb = [np.zeros((n, 3), dtype=np.uint8) for _ in range(_abi.TM_COUNT)]
        ptrs = (C.c_void_p * _abi.TM_COUNT)(*[a.ctypes.data for a in rgb])
        fl, fptrs = None, None
        if want_float:
            fl = [np.zeros((n, 3), dtype=np.float64) for _ in range(_abi.TM_COUNT)]
            fptrs = (C.c_void_p * _abi.TM_COUNT)(*[a.ctypes.data for a in fl])
        rc = self._fn("tonemap_all")(C.c_void_p(h.ctypes.data), C.c_int64(n), ptrs, fptrs)
        assert rc == 0
        return (rgb, fl) if want_float else rgb

    def tonemap_one(self, op, hdr, max_white=5.0, jodie_a=0.18):
        h = np.ascontiguousarray(hdr, dtype=np.float64).reshape(-1, 3)
        out = np.zeros((len(h), 3), dtype=np.uint8)
        rc = self._fn("tonemap_one")(C.c_int(op), C.c_double(max_white), C.c_double(jodie_a),
                                     C.c_void_p(h.ctypes.data), C.c_int64(len(h)), C.c_void_p(out.ctypes.data))
        assert rc == 0
        return out

    def load_obj(self, path):
        """LoadObject(path) of the reference (tinyobjloader, RE/RaytracingEngine.cpp:15-65): (vertices (n,3) float64, indices int32).
        Reference-built checker only."""
        f = self._fn("load_obj")
        nv, ni = C.c_int64(), C.c_int64()
        if f(path.encode(), C.byref(nv), C.byref(ni), None, None) != 0:
            raise RuntimeError("ref_load_obj failed")
        v = np.zeros((nv.value, 3), dtype=np.float64)
        i = np.zeros(ni.value, dtype=np.int32)
        if f(path.encode(), C.byref(nv), C.byref(ni), C.c_void_p(v.ctypes.data), C.c_void_p(i.ctypes.data)) != 0:
            raise RuntimeError("ref_load_obj failed")
        return v, i

    def write_ppm(self, path, rgb8, width, height) -> int:
        a = np.ascontiguousarray(rgb8, dtype=np.uint8)
        return self._fn("write_ppm")(path.encode(), C.c_void_p(a.ctypes.data), C.c_uint64(width), C.c_uint64(height))


def load_ref() -> Optional[Oracle]:
    return Oracle(REF_LIB, "ref_") if os.path.exists(REF_LIB) else None


def load_port() -> Optional[Oracle]:
    return Oracle(PORT_LIB, "port_") if os.path.exists(PORT_LIB) else None


def load_best() -> Oracle:
    """The reference-built checker when present, else the port."""
    o = load_ref() or load_port()
    if o is None:
        raise RuntimeError("no oracle library built: run `make -C oracle`")
    return o
