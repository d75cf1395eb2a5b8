"""Host-side scene setup API mirroring the reference's (Python flavour, used by tests and bench).

Same names, argument order and defaults as the reference's C++ types ("RE/" =
/root/reference/RaytracingEngine/); the C++20 twin lives in include/rte/engine.hpp.  These classes only
*describe* a scene: every intersection / shading / tonemap computation happens in the CUDA library
behind the C ABI (include/rte.h).  There is no CPU implementation in this package.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import _abi


def _v3(v) -> np.ndarray:
    a = np.asarray(v, dtype=np.float64).reshape(3)
    return a


@dataclass
class Material:
    """Material, RE/Shape.h:13-19 (same defaults)."""
    color: Sequence[float] = (0.0, 0.0, 0.0)
    shininess: float = 128.0
    specular: float = 0.0
    transparency: float = 0.0
    refractiveIndex: float = 1.0


@dataclass
class Transform:
    """Transform, RE/Shape.h:7-11. Only `position` is ever applied by the reference (RE/Shape.h:198-200)."""
    position: Sequence[float] = (0.0, 0.0, 0.0)
    rotation: Sequence[float] = (0.0, 0.0, 0.0)
    scale: Sequence[float] = (0.0, 0.0, 0.0)


@dataclass
class Sphere:
    """Sphere(r=1, pos=0, mat), RE/Shape.h:65."""
    radius: float = 1.0
    position: Sequence[float] = (0.0, 0.0, 0.0)
    material: Material = field(default_factory=Material)


@dataclass
class Plane:
    """Plane(pos=(0,1,0), norm=(0,1,0), mat), RE/Shape.h:141 (normal normalised by the library)."""
    position: Sequence[float] = (0.0, 1.0, 0.0)
    normal: Sequence[float] = (0.0, 1.0, 0.0)
    material: Material = field(default_factory=Material)


@dataclass
class Triangle:
    """Triangle(v0, v1, v2, mat, transform), RE/Shape.h:194."""
    v0: Sequence[float]
    v1: Sequence[float]
    v2: Sequence[float]
    material: Material = field(default_factory=Material)
    transform: Transform = field(default_factory=Transform)


@dataclass
class Model:
    """Model(indices, transform, material, vertexPositions), RE/Shape.h:256."""
    vertices: Sequence[int]  # flat index list, 3 per triangle (the reference calls it `vertices`)
    transform: Transform = field(default_factory=Transform)
    material: Material = field(default_factory=Material)
    vertexPositions: Optional[np.ndarray] = None  # (n,3) float64


@dataclass
class Light:
    """Light(position, color, intensity); default (origin, white, 1), RE/Light.h:10-15."""
    position: Sequence[float] = (0.0, 0.0, 0.0)
    color: Sequence[float] = (1.0, 1.0, 1.0)
    intensity: float = 1.0


@dataclass
class Camera:
    """Camera(position, focal=1, width=800, height=600, near=1, far=1000), RE/Math.h:96-97."""
    position: Sequence[float]
    focal: float = 1.0
    width: int = 800
    height: int = 600
    nearPlaneDistance: float = 1.0
    farPlaneDistance: float = 1000.0
    antiAliasingAmount: int = 32  # RE/Math.h:94
    jitterSeed: int = 0           # rte_camera.jitter_seed (upstream seeds its jitter from std::random_device)

    def to_c(self) -> _abi.Camera:
        c = _abi.Camera()
        p = _v3(self.position)
        c.position[0], c.position[1], c.position[2] = float(p[0]), float(p[1]), float(p[2])
        c.focal = float(self.focal)
        c.width, c.height = int(self.width), int(self.height)
        c.near_plane, c.far_plane = float(self.nearPlaneDistance), float(self.farPlaneDistance)
        c.anti_aliasing = int(self.antiAliasingAmount)
        c.jitter_seed = int(self.jitterSeed)
        return c


class SceneArrays:
    """Flat arrays of one scene in the C layout of include/rte.h (what rte_upload_scene consumes)."""

    def __init__(self, materials=None, spheres=None, planes=None, triangles=None, models=None,
                 model_vertices=None, model_indices=None, lights=None):
        def arr(a, dt):
            if a is None:
                return np.zeros(0, dtype=dt)
            a = np.ascontiguousarray(a, dtype=dt)
            return a
        self.materials = arr(materials, _abi.material_dt)
        self.spheres = arr(spheres, _abi.sphere_dt)
        self.planes = arr(planes, _abi.plane_dt)
        self.triangles = arr(triangles, _abi.triangle_dt)
        self.models = arr(models, _abi.model_dt)
        self.model_vertices = np.ascontiguousarray(
            np.zeros((0, 3)) if model_vertices is None else model_vertices, dtype=np.float64).reshape(-1, 3)
        self.model_indices = np.ascontiguousarray(
            np.zeros(0, dtype=np.int32) if model_indices is None else model_indices, dtype=np.int32)
        self.lights = arr(lights, _abi.light_dt)

    def desc(self) -> _abi.SceneDesc:
        """The rte_scene_desc view (pointers into this object's arrays: keep `self` alive)."""
        d = _abi.SceneDesc()

        def ptr(a):
            return a.ctypes.data if a.size else None
        d.materials, d.n_materials = ptr(self.materials), len(self.materials)
        d.spheres, d.n_spheres = ptr(self.spheres), len(self.spheres)
        d.planes, d.n_planes = ptr(self.planes), len(self.planes)
        d.triangles, d.n_triangles = ptr(self.triangles), len(self.triangles)
        d.models, d.n_models = ptr(self.models), len(self.models)
        d.model_vertices, d.n_model_vertices = ptr(self.model_vertices), len(self.model_vertices)
        d.model_indices, d.n_model_indices = ptr(self.model_indices), len(self.model_indices)
        d.lights, d.n_lights = ptr(self.lights), len(self.lights)
        return d

    @property
    def n_model_triangles(self) -> int:
        return int(self.models["index_count"].sum() // 3) if len(self.models) else 0

    def counts(self) -> dict:
        return {"spheres": len(self.spheres), "planes": len(self.planes),
                "triangles": len(self.triangles) + self.n_model_triangles, "models": len(self.models),
                "lights": len(self.lights), "materials": len(self.materials)}


class Scene:
    """Scene(camera) with AddSphere/AddPlane/AddLight/AddTriangle/AddModel, RE/Scene.h:200-212.

    Value semantics like the reference: Add* copies what it is given.  RenderImage() dispatches to
    the CUDA library (engine.Engine); nothing is computed on the host.
    """

    def __init__(self, camera: Camera):
        self.camera = Camera(**vars(camera))  # the reference copies the camera, RE/Scene.h:200
        self._materials: List[Material] = []
        self._spheres: List[tuple] = []
        self._planes: List[tuple] = []
        self._triangles: List[tuple] = []
        self._models: List[tuple] = []
        self._lights: List[Light] = []
        self._engine = None

    def _mat(self, m: Material) -> int:
        self._materials.append(Material(tuple(_v3(m.color)), m.shininess, m.specular, m.transparency, m.refractiveIndex))
        return len(self._materials) - 1

    def AddSphere(self, sphere: Sphere) -> None:
        self._spheres.append((tuple(_v3(sphere.position)), float(sphere.radius), self._mat(sphere.material)))

    def AddPlane(self, plane: Plane) -> None:
        self._planes.append((tuple(_v3(plane.position)), tuple(_v3(plane.normal)), self._mat(plane.material)))

    def AddTriangle(self, t: Triangle) -> None:
        self._triangles.append((tuple(_v3(t.v0)), tuple(_v3(t.v1)), tuple(_v3(t.v2)),
                                tuple(_v3(t.transform.position)), self._mat(t.material)))

    def AddModel(self, m: Model) -> None:
        pos = np.zeros((0, 3)) if m.vertexPositions is None else np.array(m.vertexPositions, dtype=np.float64).reshape(-1, 3)
        idx = np.array(m.vertices, dtype=np.int32).reshape(-1)
        if idx.size % 3:
            raise ValueError("Model index list must hold 3 indices per triangle")
        if idx.size and (idx.min() < 0 or idx.max() >= len(pos)):
            raise IndexError("Model index out of range")  # the reference would read out of bounds
        self._models.append((pos, idx, tuple(_v3(m.transform.position)), self._mat(m.material)))

    def AddLight(self, light: Light) -> None:
        self._lights.append(Light(tuple(_v3(light.position)), tuple(_v3(light.color)), float(light.intensity)))

    def GetPixelIndex(self, x: int, y: int) -> int:
        """RE/Scene.h:214-216."""
        return y * self.camera.width + x

    def arrays(self) -> SceneArrays:
        mats = np.zeros(len(self._materials), dtype=_abi.material_dt)
        for i, m in enumerate(self._materials):
            mats[i] = (m.color, m.shininess, m.specular, m.transparency, m.refractiveIndex)
        sph = np.zeros(len(self._spheres), dtype=_abi.sphere_dt)
        for i, (c, r, mi) in enumerate(self._spheres):
            sph[i] = (c, r, mi, 0)
        pl = np.zeros(len(self._planes), dtype=_abi.plane_dt)
        for i, (p, n, mi) in enumerate(self._planes):
            pl[i] = (p, n, mi, 0)
        tri = np.zeros(len(self._triangles), dtype=_abi.triangle_dt)
        for i, (a, b, c, t, mi) in enumerate(self._triangles):
            tri[i] = (a, b, c, t, mi, 0)
        models = np.zeros(len(self._models), dtype=_abi.model_dt)
        verts, inds = [], []
        nv = ni = 0
        for i, (pos, idx, t, mi) in enumerate(self._models):
            models[i] = (nv, len(pos), ni, len(idx), t, mi, 0)
            verts.append(pos)
            inds.append(idx)
            nv += len(pos)
            ni += len(idx)
        lights = np.zeros(len(self._lights), dtype=_abi.light_dt)
        for i, l in enumerate(self._lights):
            lights[i] = (l.position, l.color, l.intensity)
        return SceneArrays(mats, sph, pl, tri, models,
                           np.concatenate(verts) if verts else None,
                           np.concatenate(inds) if inds else None, lights)

    # --- the hot path: delegated to the CUDA library --------------------------------------------
    def engine(self):
        from .engine import Engine
        if self._engine is None:
            self._engine = Engine()
        self._engine.upload(self.arrays(), self.camera)
        return self._engine

    def RenderImage(self) -> np.ndarray:
        """Scene::RenderImage(), RE/Scene.h:311 — (H*W, 3) float64 linear HDR, rendered on the GPU."""
        return self.engine().render_host(outputs=("hdr64",))["hdr64"].reshape(-1, 3)
