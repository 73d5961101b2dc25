"""Scene constants without the product library: ctypes over libgprt_scene.so (host/scene.cpp +
host/scene_capi.cpp built with plain g++, no CUDA).  bench.py's reference arm and CPU-only tools
use this so that timing the reference's code never maps libgprt.so.
"""
import ctypes as C
import os

import numpy as np

from . import _capi
from ._capi import Config, SceneView
from .renderer import SceneData

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgprt_scene.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(LIB_PATH)
        vp, u32, i32, f32 = C.c_void_p, C.c_uint32, C.c_int, C.c_float
        L.gprt_scene_create.restype = i32
        L.gprt_scene_create.argtypes = [C.POINTER(vp), i32, u32, u32, u32]
        L.gprt_scene_destroy.restype = None
        L.gprt_scene_destroy.argtypes = [vp]
        L.gprt_scene_update.restype = i32
        L.gprt_scene_update.argtypes = [vp, f32, f32, f32]
        L.gprt_scene_get.restype = i32
        L.gprt_scene_get.argtypes = [vp, C.POINTER(SceneView)]
        _lib = L
    return _lib


class SceneOnly:
    """The host half of Renderer (scene_id, size, on_update, scene()) and nothing else."""

    def __init__(self, scene_id=0, width=0, height=0, grid_n=0):
        self._lib = lib()
        self._h = C.c_void_p()
        rc = self._lib.gprt_scene_create(C.byref(self._h), int(scene_id), int(width), int(height), int(grid_n))
        if rc != 0:
            raise RuntimeError(f"gprt_scene_create failed: {rc}")

    def on_update(self, animate_geometry_time=0.0, camera_orbit_deg=0.0, light_orbit_deg=0.0):
        rc = self._lib.gprt_scene_update(self._h, float(animate_geometry_time), float(camera_orbit_deg), float(light_orbit_deg))
        if rc != 0:
            raise RuntimeError(f"gprt_scene_update failed: {rc}")

    def scene(self) -> SceneData:
        v = SceneView()
        rc = self._lib.gprt_scene_get(self._h, C.byref(v))
        if rc != 0:
            raise RuntimeError(f"gprt_scene_get failed: {rc}")
        va = _capi.view_array
        cfg = Config()
        C.memmove(C.byref(cfg), C.byref(v.config), C.sizeof(Config))
        return SceneData(width=v.width, height=v.height, config=cfg,
                         scene_cb=va(v.scene_cb, 1, _capi.SCENE_CB_DTYPE), per_frame=va(v.per_frame, v.n_per_frame, _capi.PER_FRAME_DTYPE),
                         aabbs=va(v.aabbs, v.n_aabbs, _capi.AABB_DTYPE), hit_groups=va(v.hit_groups, v.n_hit_groups, _capi.HIT_GROUP_DTYPE),
                         vertices=va(v.vertices, v.n_vertices, _capi.VERTEX_DTYPE), indices=va(v.indices, v.n_indices, np.dtype("<u2")),
                         instances=va(v.instances, v.n_instances, _capi.INSTANCE_DTYPE))

    def close(self):
        if self._h:
            self._lib.gprt_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
