"""ctypes loader of the CPU oracle.  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Nothing under gpuraytracer_b200/ does.

Two oracles exist, with the same C interface:
  kind "reference": oracle/_ref/liboracle_ref*.so -- the reference's own HLSL shader sources
                    compiled as C++ (oracle/build_ref.py, oracle/ref_harness.cpp)
  kind "port":      oracle/_build/liboracle_port*.so -- the hand restatement in
                    oracle/restated/ (plain C++, no reference sources needed to build)
`load()` prefers "reference" and falls back to "port".
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class OracleConfig(C.Structure):
    _fields_ = [("max_recursion_depth", C.c_uint32), ("sdf_max_steps", C.c_uint32), ("metaball_max_steps", C.c_uint32),
                ("force_brute_force", C.c_uint32), ("fractal_iterations", C.c_uint32)]


class OracleCounters(C.Structure):
    _fields_ = [("radiance_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("isect", C.c_uint64 * 3), ("sdf_evals", C.c_uint64 * 8),
                ("metaball_steps", C.c_uint64), ("aabb_tests", C.c_uint64)]

    def as_dict(self):
        return {"radiance_rays": int(self.radiance_rays), "shadow_rays": int(self.shadow_rays), "isect": [int(x) for x in self.isect],
                "sdf_evals": [int(x) for x in self.sdf_evals], "metaball_steps": int(self.metaball_steps), "aabb_tests": int(self.aabb_tests)}


class Rect(C.Structure):
    _fields_ = [("x", C.c_uint32), ("y", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32)]


def lib_path(kind="reference", metaballs=3, fast=False):
    tag = ("_fast" if fast else "") + ("_mb5" if metaballs == 5 else "")
    if kind == "reference":
        return os.path.join(_HERE, "_ref", f"liboracle_ref{tag}.so")
    return os.path.join(_HERE, "_build", f"liboracle_port{tag}.so")


class Oracle:
    def __init__(self, path):
        self.path = path
        L = C.CDLL(path)
        vp, u32, f32 = C.c_void_p, C.c_uint32, C.c_float
        L.oracle_set_scene.restype = C.c_int
        L.oracle_set_scene.argtypes = [vp, u32, vp, u32, vp, u32, vp, u32, vp, u32, u32, C.POINTER(OracleConfig)]
        L.oracle_render.restype = C.c_int
        L.oracle_render.argtypes = [vp, vp, u32, u32, u32, Rect, vp, vp, vp, C.POINTER(OracleCounters), C.c_int]
        L.oracle_intersect_local.restype = C.c_int
        L.oracle_intersect_local.argtypes = [u32, u32, vp, vp, f32, f32, u32, f32, f32, C.POINTER(f32), vp]
        L.oracle_sdf_distance.restype = f32
        L.oracle_sdf_distance.argtypes = [u32, vp]
        L.oracle_set_fractal_iterations.restype = None
        L.oracle_set_fractal_iterations.argtypes = [u32]
        L.oracle_checkers_box_filter.restype = f32
        L.oracle_checkers_box_filter.argtypes = [vp, vp, vp, u32]
        L.oracle_fresnel_schlick.restype = None
        L.oracle_fresnel_schlick.argtypes = [vp, vp, vp, vp]
        L.oracle_animation_interpolant.restype = f32
        L.oracle_animation_interpolant.argtypes = [f32, f32]
        L.oracle_camera_ray.restype = None
        L.oracle_camera_ray.argtypes = [vp, u32, u32, u32, u32, vp, vp]
        L.oracle_metaballs_count.restype = C.c_int
        L.oracle_kind.restype = C.c_char_p
        L.oracle_max_threads.restype = C.c_int
        L.oracle_last_threads.restype = C.c_int
        self.L = L
        self.kind = L.oracle_kind().decode()
        self.metaballs = L.oracle_metaballs_count()
        self._keep = None

    def max_threads(self):
        return int(self.L.oracle_max_threads())

    def last_threads(self):
        """Threads the last render() really ran on."""
        return int(self.L.oracle_last_threads())

    def set_scene(self, scene, force_brute_force=False, max_recursion_depth=None, sdf_max_steps=None, metaball_max_steps=None,
                  fractal_iterations=None):
        """scene: gpuraytracer_b200.SceneData (numpy arrays with the ABI layouts)."""
        cfg = OracleConfig(max_recursion_depth if max_recursion_depth is not None else scene.config.max_recursion_depth,
                           sdf_max_steps if sdf_max_steps is not None else scene.config.sdf_max_steps,
                           metaball_max_steps if metaball_max_steps is not None else scene.config.metaball_max_steps,
                           1 if force_brute_force else 0,
                           fractal_iterations if fractal_iterations is not None else getattr(scene.config, "fractal_iterations", 4))
        if scene.config.metaballs_count != self.metaballs:
            raise ValueError(f"oracle built for {self.metaballs} metaballs, scene wants {scene.config.metaballs_count}")
        a = {k: np.ascontiguousarray(getattr(scene, k)) for k in ("aabbs", "vertices", "indices", "instances", "hit_groups")}
        self._keep = a
        rc = self.L.oracle_set_scene(a["aabbs"].ctypes.data, len(a["aabbs"]), a["vertices"].ctypes.data, len(a["vertices"]),
                                     a["indices"].ctypes.data, len(a["indices"]), a["instances"].ctypes.data, len(a["instances"]),
                                     a["hit_groups"].ctypes.data, len(a["hit_groups"]), a["hit_groups"].dtype.itemsize, C.byref(cfg))
        if rc != 0:
            raise RuntimeError(f"oracle_set_scene failed: {rc}")

    def render(self, scene, tile=None, threads=0, want_f32=False, want_ids=True, want_rgba8=True):
        """Returns dict(rgba8, rgba_f32, hit_ids, counters)."""
        x, y, w, h = tile if tile is not None else (0, 0, scene.width, scene.height)
        rgba8 = np.zeros((h, w, 4), np.uint8) if want_rgba8 else None
        f32 = np.zeros((h, w, 4), np.float32) if want_f32 else None
        ids = np.zeros((h, w), np.uint32) if want_ids else None
        cnt = OracleCounters()
        cb = np.ascontiguousarray(scene.scene_cb)
        pf = np.ascontiguousarray(scene.per_frame)
        rc = self.L.oracle_render(cb.ctypes.data, pf.ctypes.data, len(pf), scene.width, scene.height, Rect(x, y, w, h),
                                  rgba8.ctypes.data if rgba8 is not None else None, f32.ctypes.data if f32 is not None else None,
                                  ids.ctypes.data if ids is not None else None, C.byref(cnt), int(threads))
        if rc != 0:
            raise RuntimeError(f"oracle_render failed: {rc}")
        return {"rgba8": rgba8, "rgba_f32": f32, "hit_ids": ids, "counters": cnt.as_dict()}

    # ---- function-level probes --------------------------------------------------------------
    def intersect_local(self, hit_group_kind, primitive_type, origin, direction, tmin=0.0, tcurrent=10000.0, ray_flags=0x10,
                        elapsed_time=0.0, step_scale=1.0):
        o = np.asarray(origin, np.float32)
        d = np.asarray(direction, np.float32)
        t = C.c_float(0)
        n = np.zeros(3, np.float32)
        hit = self.L.oracle_intersect_local(hit_group_kind, primitive_type, o.ctypes.data, d.ctypes.data, tmin, tcurrent, ray_flags,
                                            elapsed_time, step_scale, C.byref(t), n.ctypes.data)
        return (bool(hit), float(t.value), n)

    def set_fractal_iterations(self, n):
        self.L.oracle_set_fractal_iterations(int(n))

    def sdf_distance(self, primitive_type, p):
        p = np.asarray(p, np.float32)
        return float(self.L.oracle_sdf_distance(primitive_type, p.ctypes.data))

    def checkers(self, uv, dpdx, dpdy, ratio):
        a, b, c = (np.asarray(v, np.float32) for v in (uv, dpdx, dpdy))
        return float(self.L.oracle_checkers_box_filter(a.ctypes.data, b.ctypes.data, c.ctypes.data, ratio))

    def fresnel(self, i, n, f0):
        a, b, c = (np.asarray(v, np.float32) for v in (i, n, f0))
        out = np.zeros(3, np.float32)
        self.L.oracle_fresnel_schlick(a.ctypes.data, b.ctypes.data, c.ctypes.data, out.ctypes.data)
        return out

    def animation_interpolant(self, elapsed, cycle):
        return float(self.L.oracle_animation_interpolant(elapsed, cycle))

    def camera_ray(self, scene, x, y):
        cb = np.ascontiguousarray(scene.scene_cb)
        o = np.zeros(3, np.float32)
        d = np.zeros(3, np.float32)
        self.L.oracle_camera_ray(cb.ctypes.data, scene.width, scene.height, x, y, o.ctypes.data, d.ctypes.data)
        return o, d


def available(kind="reference", metaballs=3, fast=False):
    return os.path.exists(lib_path(kind, metaballs, fast))


def load(kind=None, metaballs=3, fast=False) -> Oracle:
    """kind None: "reference" if built, else "port"."""
    kinds = [kind] if kind else ["reference", "port"]
    for k in kinds:
        p = lib_path(k, metaballs, fast)
        if os.path.exists(p):
            return Oracle(p)
    raise FileNotFoundError(f"no oracle library for kinds={kinds} metaballs={metaballs} fast={fast}; run __graft_entry__.build()")
