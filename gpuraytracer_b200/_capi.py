"""ctypes binding of libgprt.so (include/gprt.h + include/gprt_host.h).

The library is the product: if it is missing this module raises at import of the symbols --
there is no Python or CPU fallback for the render path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgprt.so")

GPRT_ABI_VERSION = 2
GPRT_MATH_PARITY, GPRT_MATH_FAST, GPRT_MATH_FMA = 0, 1, 2
GPRT_FLAG_BRUTE_FORCE_TRAVERSAL = 1
GPRT_FLAG_PROFILE_KERNELS = 2
GPRT_FLAG_MEGAKERNEL = 4
GPRT_K_TRAVERSE, GPRT_K_INTERSECT0, GPRT_K_RESOLVE, GPRT_K_SHADE_SPAWN, GPRT_K_FINAL, GPRT_K_INTERSECT_ALL, GPRT_K_COUNT = 0, 1, 11, 12, 13, 14, 15
GPRT_HIT_MISS = 0xFFFFFFFF
GPRT_HIT_TRIANGLE = 0x00000000
GPRT_HIT_AABB = 0x10000000
GPRT_HG_TRIANGLE, GPRT_HG_AABB_ANALYTIC, GPRT_HG_AABB_VOLUMETRIC, GPRT_HG_AABB_SDF = 0, 1, 2, 3

GPRT_OK = 0
GPRT_E_INVALID_ARG, GPRT_E_CUDA, GPRT_E_NO_DEVICE, GPRT_E_OUT_OF_MEMORY = -1, -2, -3, -4
GPRT_E_NOT_BUILT, GPRT_E_QUEUE_OVERFLOW, GPRT_E_UNSUPPORTED, GPRT_E_ABI_MISMATCH, GPRT_E_TOO_LARGE = -5, -6, -7, -8, -9
GPRT_NOALLOC_MAX_PRIMS, GPRT_NOALLOC_SCRATCH_BYTES = 65536, 5 << 20


class Config(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("max_recursion_depth", C.c_uint32), ("metaballs_count", C.c_uint32),
                ("metaball_max_steps", C.c_uint32), ("sdf_max_steps", C.c_uint32), ("math_mode", C.c_uint32),
                ("cand_per_ray", C.c_uint32), ("flags", C.c_uint32), ("fractal_iterations", C.c_uint32)]


class Rect(C.Structure):
    _fields_ = [("x", C.c_uint32), ("y", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32)]


class Counters(C.Structure):
    _fields_ = [("radiance_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("candidates", C.c_uint64 * 3),
                ("bvh_node_visits", C.c_uint64), ("queue_overflow", C.c_uint64)]


class KernelProfile(C.Structure):
    _fields_ = [("ms", C.c_float * 15), ("launches", C.c_uint32 * 15), ("work", C.c_uint64 * 10), ("candidates", C.c_uint64 * 10)]


class WaveStats(C.Structure):
    _fields_ = [("n_radiance", C.c_uint32), ("n_shadow", C.c_uint32), ("candidates", C.c_uint32 * 10), ("work", C.c_uint32 * 10),
                ("max_evals", C.c_uint32 * 10), ("evals_hist", (C.c_uint32 * 12) * 10),
                ("node_visits", C.c_uint32), ("max_node_visits", C.c_uint32)]


class IpcHandle(C.Structure):
    _fields_ = [("bytes", C.c_ubyte * 64)]


class Timing(C.Structure):
    _fields_ = [("dispatch_ms", C.c_float), ("build_accel_ms", C.c_float)]


class SceneView(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("config", Config),
                ("scene_cb", C.c_void_p), ("per_frame", C.c_void_p), ("n_per_frame", C.c_uint32),
                ("aabbs", C.c_void_p), ("n_aabbs", C.c_uint32),
                ("hit_groups", C.c_void_p), ("n_hit_groups", C.c_uint32),
                ("vertices", C.c_void_p), ("n_vertices", C.c_uint32),
                ("indices", C.c_void_p), ("n_indices", C.c_uint32),
                ("instances", C.c_void_p), ("n_instances", C.c_uint32)]


# numpy views of the ABI structs (sizes asserted against include/gprt.h in tests/test_abi.py)
KERNEL_NAMES = ["traverse", "isect_box", "isect_spheres", "isect_metaballs", "isect_minispheres", "isect_roundcube", "isect_sqtorus",
                "isect_twtorus", "isect_cog", "isect_cylinder", "isect_pyramid", "resolve", "shade_spawn", "final", "intersect_all"]
TYPE_NAMES = KERNEL_NAMES[1:11]

SCENE_CB_DTYPE = np.dtype([("projection_to_world", "<f4", (4, 4)), ("camera_position", "<f4", 4), ("light_position", "<f4", 4),
                           ("light_ambient_color", "<f4", 4), ("light_diffuse_color", "<f4", 4), ("reflectance", "<f4"),
                           ("elapsed_time", "<f4"), ("_pad", "<f4", 2)])
MATERIAL_DTYPE = np.dtype([("albedo", "<f4", 4), ("reflectance_coefficient", "<f4"), ("diffuse_coefficient", "<f4"),
                           ("specular_coefficient", "<f4"), ("specular_power", "<f4"), ("step_scale", "<f4"), ("padding", "<f4", 3)])
PER_FRAME_DTYPE = np.dtype([("local_space_to_bottom_level_as", "<f4", (4, 4)), ("bottom_level_as_to_local_space", "<f4", (4, 4))])
AABB_DTYPE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3)])
VERTEX_DTYPE = np.dtype([("position", "<f4", 3), ("normal", "<f4", 3)])
INSTANCE_DTYPE = np.dtype([("transform", "<f4", (3, 4)), ("instance_id_and_mask", "<u4"), ("hit_group_contrib_and_flags", "<u4"),
                           ("acceleration_structure", "<u8")])
HIT_GROUP_DTYPE = np.dtype([("shader_id", "<u4", 8), ("material_cb", MATERIAL_DTYPE), ("instance_index", "<u4"),
                            ("primitive_type", "<u4"), ("_pad", "<u4", 2)])

_lib = None


def lib():
    """Load libgprt.so (built by __graft_entry__.build()).  Raises if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`. "
                           "gpuraytracer_b200 has no fallback path without its CUDA library.")
    L = C.CDLL(LIB_PATH)
    vp, u32, i32, f32 = C.c_void_p, C.c_uint32, C.c_int, C.c_float
    sig = {
        "gprt_default_config": (None, [C.POINTER(Config)]),
        "gprt_create": (i32, [C.POINTER(vp), i32, C.POINTER(Config)]),
        "gprt_destroy": (None, [vp]),
        "gprt_set_geometry": (i32, [vp, vp, u32, vp, u32, vp, u32, vp, u32]),
        "gprt_set_hit_group_table": (i32, [vp, vp, u32, u32]),
        "gprt_build_accel": (i32, [vp]),
        "gprt_refit_accel": (i32, [vp, vp, u32]),
        "gprt_dispatch": (i32, [vp, vp, vp, u32, u32, u32, Rect, vp, C.c_size_t, vp, vp]),
        "gprt_dispatch_stripes": (i32, [vp, vp, vp, u32, u32, u32, u32, u32, u32, vp, C.c_size_t, vp]),
        "gprt_stripe_local_rows": (u32, [u32, u32, u32, u32]),
        "gprt_dispatch_host_async": (i32, [vp, vp, vp, u32, u32, u32, Rect, vp, C.c_size_t, vp]),
        "gprt_dispatch_stripes_host_async": (i32, [vp, vp, vp, u32, u32, u32, u32, u32, u32, vp, C.c_size_t]),
        "gprt_get_stream": (vp, [vp]),
        "gprt_dispatch_host": (i32, [vp, vp, vp, u32, u32, u32, Rect, vp, C.c_size_t, vp]),
        "gprt_sync": (i32, [vp]),
        "gprt_poll_status": (i32, [vp]),
        "gprt_get_counters": (i32, [vp, C.POINTER(Counters)]),
        "gprt_last_timing": (i32, [vp, C.POINTER(Timing)]),
        "gprt_get_kernel_profile": (i32, [vp, C.POINTER(KernelProfile)]),
        "gprt_measure_fp32_peak": (i32, [vp, C.POINTER(C.c_float)]),
        "gprt_get_wave_stats": (i32, [vp, u32, C.POINTER(WaveStats)]),
        "gprt_last_cuda_error": (i32, [vp]),
        "gprt_set_stream": (i32, [vp, vp]),
        "gprt_last_launch_count": (i32, [vp]),
        "gprt_alloc_host": (vp, [C.c_size_t]),
        "gprt_free_host": (None, [vp]),
        "gprt_ipc_alloc": (i32, [i32, C.c_size_t, C.POINTER(vp), C.POINTER(IpcHandle)]),
        "gprt_ipc_open": (i32, [i32, C.POINTER(IpcHandle), C.POINTER(vp)]),
        "gprt_ipc_close": (i32, [i32, vp]),
        "gprt_ipc_free": (i32, [i32, vp]),
        "gprt_lbvh_build_noalloc": (i32, [i32, vp, u32, vp, vp, vp, C.POINTER(C.c_int32), vp]),
        "gprt_get_accel": (i32, [vp, u32, vp, vp, vp, C.POINTER(C.c_int32)]),
        "gprt_error_string": (C.c_char_p, [i32]),
        "gprt_abi_version": (u32, []),
        "gprt_renderer_create": (i32, [C.POINTER(vp), i32, u32, u32, u32, i32, C.POINTER(Config)]),
        "gprt_renderer_destroy": (None, [vp]),
        "gprt_renderer_on_init": (i32, [vp]),
        "gprt_renderer_on_update": (i32, [vp, f32, f32, f32]),
        "gprt_renderer_on_render": (i32, [vp, Rect, vp, vp]),
        "gprt_renderer_do_raytracing": (i32, [vp, Rect, vp, vp, vp]),
        "gprt_renderer_do_raytracing_stripes": (i32, [vp, u32, u32, u32, vp, vp]),
        "gprt_renderer_on_render_async": (i32, [vp, vp]),
        "gprt_renderer_do_raytracing_async": (i32, [vp, vp]),
        "gprt_renderer_do_raytracing_stripes_async": (i32, [vp, u32, u32, u32, vp]),
        "gprt_renderer_on_render_stripes_async": (i32, [vp, u32, u32, u32, vp]),
        "gprt_renderer_create_frame_contexts": (i32, [vp]),
        "gprt_renderer_set_frame_count": (i32, [vp, i32]),
        "gprt_renderer_get_frame_count": (i32, [vp]),
        "gprt_renderer_set_device_transforms": (i32, [vp, i32]),
        "gprt_renderer_get_device_transforms": (i32, [vp]),
        "gprt_set_animation": (i32, [vp, vp, u32]),
        "gprt_update_transforms": (i32, [vp, f32]),
        "gprt_get_transforms": (i32, [vp, vp, u32]),
        "gprt_renderer_next_frame_slot": (i32, [vp]),
        "gprt_renderer_wait_for_gpu": (i32, [vp]),
        "gprt_renderer_frame_context": (vp, [vp, i32]),
        "gprt_renderer_on_size_changed": (i32, [vp, u32, u32]),
        "gprt_renderer_scene": (i32, [vp, C.POINTER(SceneView)]),
        "gprt_renderer_get_size": (i32, [vp, C.POINTER(u32), C.POINTER(u32)]),
        "gprt_renderer_context": (vp, [vp]),
        "gprt_renderer_primary_mrays_per_s": (f32, [vp]),
        "gprt_write_ppm": (i32, [C.c_char_p, vp, u32, u32]),
        "gprt_enable_peer_access": (i32, [vp, i32]),
        "gprt_multi_create": (i32, [C.POINTER(vp), i32, u32, u32, u32, C.POINTER(i32), i32, C.POINTER(Config)]),
        "gprt_multi_destroy": (None, [vp]),
        "gprt_multi_on_init": (i32, [vp]),
        "gprt_multi_on_update": (i32, [vp, f32, f32, f32]),
        "gprt_multi_on_render_async": (i32, [vp, vp]),
        "gprt_multi_do_raytracing_async": (i32, [vp, vp]),
        "gprt_multi_wait_for_gpu": (i32, [vp]),
        "gprt_multi_wait_for_frame_slot": (i32, [vp, i32]),
        "gprt_multi_set_frame_count": (i32, [vp, i32]),
        "gprt_multi_get_size": (i32, [vp, C.POINTER(u32), C.POINTER(u32)]),
        "gprt_multi_stripe_rows": (u32, [vp]),
        "gprt_multi_context": (vp, [vp, i32]),
        "gprt_scene_create": (i32, [C.POINTER(vp), i32, u32, u32, u32]),
        "gprt_scene_destroy": (None, [vp]),
        "gprt_scene_update": (i32, [vp, f32, f32, f32]),
        "gprt_scene_get": (i32, [vp, C.POINTER(SceneView)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


EXPORTED_SYMBOLS = [
    "gprt_default_config", "gprt_alloc_host", "gprt_free_host", "gprt_create", "gprt_destroy", "gprt_set_geometry", "gprt_set_hit_group_table", "gprt_build_accel",
    "gprt_refit_accel", "gprt_dispatch", "gprt_dispatch_stripes", "gprt_stripe_local_rows", "gprt_dispatch_host", "gprt_dispatch_host_async", "gprt_dispatch_stripes_host_async", "gprt_get_stream", "gprt_sync", "gprt_poll_status", "gprt_get_counters", "gprt_last_timing",
    "gprt_get_kernel_profile", "gprt_get_wave_stats", "gprt_measure_fp32_peak", "gprt_last_cuda_error", "gprt_set_stream", "gprt_last_launch_count", "gprt_error_string", "gprt_abi_version",
    "gprt_renderer_create", "gprt_renderer_destroy", "gprt_renderer_on_init", "gprt_renderer_on_update", "gprt_renderer_on_render",
    "gprt_renderer_do_raytracing", "gprt_renderer_do_raytracing_stripes", "gprt_renderer_on_render_async", "gprt_renderer_do_raytracing_async",
    "gprt_renderer_do_raytracing_stripes_async", "gprt_renderer_on_render_stripes_async", "gprt_renderer_create_frame_contexts", "gprt_renderer_set_frame_count", "gprt_renderer_get_frame_count", "gprt_renderer_set_device_transforms", "gprt_renderer_get_device_transforms", "gprt_renderer_next_frame_slot",
    "gprt_set_animation", "gprt_update_transforms", "gprt_get_transforms",
    "gprt_renderer_wait_for_gpu", "gprt_renderer_frame_context", "gprt_renderer_on_size_changed", "gprt_renderer_scene", "gprt_renderer_get_size", "gprt_renderer_context",
    "gprt_renderer_primary_mrays_per_s", "gprt_write_ppm",
    "gprt_enable_peer_access", "gprt_multi_create", "gprt_multi_destroy", "gprt_multi_on_init", "gprt_multi_on_update", "gprt_multi_on_render_async",
    "gprt_multi_do_raytracing_async", "gprt_multi_wait_for_gpu", "gprt_multi_wait_for_frame_slot", "gprt_multi_set_frame_count", "gprt_multi_get_size", "gprt_multi_stripe_rows", "gprt_multi_context",
    "gprt_scene_create", "gprt_scene_destroy", "gprt_scene_update", "gprt_scene_get",
    "gprt_ipc_alloc", "gprt_ipc_open", "gprt_ipc_close", "gprt_ipc_free",
    "gprt_lbvh_build_noalloc", "gprt_get_accel",
]


class GprtError(RuntimeError):
    def __init__(self, status, what=""):
        self.status = status
        msg = lib().gprt_error_string(status).decode()
        super().__init__(f"gprt: {what}: {msg} ({status})")


def check(status, what=""):
    if status != GPRT_OK:
        raise GprtError(status, what)


def view_array(ptr, count, dtype):
    """numpy copy of `count` structs of `dtype` at C pointer `ptr`."""
    if count == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count).copy()
