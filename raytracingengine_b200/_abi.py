"""ctypes / numpy mirror of include/rte.h (struct layouts only, no behaviour).

Every structure cites the reference type it flattens ("RE/" = /root/reference/RaytracingEngine/).
The numpy dtypes have exactly the C layout so that whole primitive arrays can be generated and
handed to the C ABI without per-element marshalling.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

RTE_VERSION = 1

RTE_OK = 0
RTE_ERR_INVALID = -1
RTE_ERR_CUDA = -2
RTE_ERR_NO_DEVICE = -3
RTE_ERR_IO = -4
RTE_ERR_LIMIT = -5
RTE_ERR_TIMEOUT = -6

HIT_NONE, HIT_SPHERE, HIT_PLANE, HIT_TRIANGLE = 0, 1, 2, 3  # RE/Shape.h:21-26
PRIM_MISS = 0xFFFFFFFF

# tonemapAll order, RE/RaytracingEngine.cpp:205-211,303-311
TM_NAMES = (
    "simple",
    "reinhard_simple",
    "reinhard_extended",
    "reinhard_extended_luminance",
    "reinhard_jodie",
    "uncharted2",
    "aces",
)
TM_COUNT = 7
TM_NONE = -1
TM_SIMPLE, TM_REINHARD_SIMPLE, TM_REINHARD_EXT, TM_REINHARD_EXT_LUM, TM_REINHARD_JODIE, TM_UNCHARTED2, TM_ACES = range(7)
FLAG_EXACT_SHADOWS, FLAG_NO_FILTER, FLAG_GENERAL_KERNEL, FLAG_NO_CULL = 1, 2, 4, 8
FLAG_SINGLE_PASS, FLAG_TWO_PASS = 16, 32
FLAG_EXACT_FP64 = 64
FLAG_EXACT_GRID = 128
IPC_HANDLE_BYTES = 64  # RTE_IPC_HANDLE_BYTES
MAX_PEERS = 8          # RTE_MAX_PEERS
MAX_DEPTH = 24         # RTE_MAX_DEPTH
MAX_AA = 4096          # RTE_MAX_AA
NOALLOC_MAX_PIXELS = 1 << 26  # RTE_NOALLOC_MAX_PIXELS

# --- numpy dtypes (C layout) -------------------------------------------------------------------
material_dt = np.dtype(  # Material, RE/Shape.h:13-19
    [("color", "<f8", 3), ("shininess", "<f8"), ("specular", "<f8"), ("transparency", "<f8"),
     ("refractive_index", "<f8")], align=True)
sphere_dt = np.dtype(  # Sphere, RE/Shape.h:59-70
    [("center", "<f8", 3), ("radius", "<f8"), ("material", "<i4"), ("reserved", "<i4")], align=True)
plane_dt = np.dtype(  # Plane, RE/Shape.h:135-147
    [("point", "<f8", 3), ("normal", "<f8", 3), ("material", "<i4"), ("reserved", "<i4")], align=True)
triangle_dt = np.dtype(  # Triangle, RE/Shape.h:188-200
    [("v0", "<f8", 3), ("v1", "<f8", 3), ("v2", "<f8", 3), ("translate", "<f8", 3),
     ("material", "<i4"), ("reserved", "<i4")], align=True)
model_dt = np.dtype(  # Model, RE/Shape.h:248-256
    [("first_vertex", "<i8"), ("vertex_count", "<i8"), ("first_index", "<i8"), ("index_count", "<i8"),
     ("translate", "<f8", 3), ("material", "<i4"), ("reserved", "<i4")], align=True)
light_dt = np.dtype(  # Light, RE/Light.h:6-12
    [("position", "<f8", 3), ("color", "<f8", 3), ("intensity", "<f8")], align=True)
hit_dt = np.dtype(  # HitInfo, RE/Shape.h:28-34
    [("distance", "<f8"), ("normal", "<f8", 3), ("point", "<f8", 3), ("prim_id", "<u4"),
     ("material", "<i4")], align=True)

assert material_dt.itemsize == 56 and sphere_dt.itemsize == 40 and plane_dt.itemsize == 56
assert triangle_dt.itemsize == 104 and model_dt.itemsize == 64 and light_dt.itemsize == 56
assert hit_dt.itemsize == 64


# --- ctypes structures ---------------------------------------------------------------------------
class Camera(C.Structure):  # Camera, RE/Math.h:85-97
    _fields_ = [("position", C.c_double * 3), ("focal", C.c_double), ("width", C.c_uint64),
                ("height", C.c_uint64), ("near_plane", C.c_double), ("far_plane", C.c_double),
                ("anti_aliasing", C.c_int32), ("jitter_seed", C.c_int32)]


class SceneDesc(C.Structure):  # the vectors of Scene, RE/Scene.h:17-23
    _fields_ = [("materials", C.c_void_p), ("n_materials", C.c_int64),
                ("spheres", C.c_void_p), ("n_spheres", C.c_int64),
                ("planes", C.c_void_p), ("n_planes", C.c_int64),
                ("triangles", C.c_void_p), ("n_triangles", C.c_int64),
                ("models", C.c_void_p), ("n_models", C.c_int64),
                ("model_vertices", C.c_void_p), ("n_model_vertices", C.c_int64),
                ("model_indices", C.c_void_p), ("n_model_indices", C.c_int64),
                ("lights", C.c_void_p), ("n_lights", C.c_int64)]


class RenderOpts(C.Structure):
    _fields_ = [("row0", C.c_int32), ("row1", C.c_int32), ("max_depth", C.c_int32),
                ("tonemap_op", C.c_int32), ("bias", C.c_double), ("max_white", C.c_double),
                ("jodie_a", C.c_double), ("spp_seed", C.c_int32), ("flags", C.c_int32),
                ("stripe_rows", C.c_int32), ("stripe_period", C.c_int32)]


class Frame(C.Structure):
    _fields_ = [("hdr", C.c_void_p), ("hdr64", C.c_void_p), ("rgb8", C.c_void_p),
                ("prim_id", C.c_void_p), ("depth", C.c_void_p), ("depth64", C.c_void_p),
                ("normal", C.c_void_p), ("albedo", C.c_void_p), ("light_mask", C.c_void_p), ("depth_norm", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("secondary_rays", C.c_uint64), ("sphere_tests", C.c_uint64),
                ("plane_tests", C.c_uint64), ("triangle_tests", C.c_uint64),
                ("exact_marches", C.c_uint64), ("fp64_confirms", C.c_uint64),
                ("kernel_ms", C.c_double), ("tonemap_ms", C.c_double), ("sphere_tests_executed", C.c_uint64)]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


assert C.sizeof(Camera) == 72 and C.sizeof(SceneDesc) == 128
assert C.sizeof(RenderOpts) == 56 and C.sizeof(Frame) == 80 and C.sizeof(Stats) == 88

# algorithmic flops per ray-primitive test
FLOP_SPHERE, FLOP_PLANE, FLOP_TRIANGLE = 20, 14, 46


def algorithmic_flops(sphere_tests: int, plane_tests: int, triangle_tests: int) -> int:
    return FLOP_SPHERE * sphere_tests + FLOP_PLANE * plane_tests + FLOP_TRIANGLE * triangle_tests
