"""Python face of the headless Renderer (thin: every call goes straight to libgprt.so).

Mirrors the reference's Renderer entry points (src/Renderer.h:24-28): on_init, on_update,
on_render, on_size_changed, on_destroy, plus do_raytracing for device-resident output.
torch is used only to hold device buffers; nothing is computed in Python.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _capi
from ._capi import Config, Counters, KernelProfile, Rect, SceneView, Timing, WaveStats, check, lib


@dataclass
class SceneData:
    """Host copies of everything do_raytracing / build_acceleration_structures consume."""
    width: int
    height: int
    config: Config
    scene_cb: np.ndarray    # SCENE_CB_DTYPE, shape (1,)
    per_frame: np.ndarray   # PER_FRAME_DTYPE, shape (n,)
    aabbs: np.ndarray       # AABB_DTYPE, shape (n,)
    hit_groups: np.ndarray  # HIT_GROUP_DTYPE, shape (2 + 2n,)
    vertices: np.ndarray
    indices: np.ndarray     # uint16
    instances: np.ndarray   # INSTANCE_DTYPE, shape (2,)


def default_config() -> Config:
    cfg = Config()
    lib().gprt_default_config(C.byref(cfg))
    return cfg


class Renderer:
    """scene_id 0 = reference scene; 1..5 = BASELINE configs; 6 = config 3 with long march budget.

    device=None builds the scene only (for feeding the oracle in CPU tests); rendering then
    raises GprtError(GPRT_E_NO_DEVICE): there is no CPU render path in this package.
    """

    def __init__(self, scene_id=0, width=0, height=0, grid_n=0, device=0, config: Config = None):
        self._h = C.c_void_p()
        self._lib = lib()
        dev = -1 if device is None else int(device)
        check(self._lib.gprt_renderer_create(C.byref(self._h), int(scene_id), int(width), int(height), int(grid_n), dev,
                                             C.byref(config) if config is not None else None), "renderer_create")
        self.device = device

    def on_init(self):
        check(self._lib.gprt_renderer_on_init(self._h), "on_init")
        return self

    def on_update(self, animate_geometry_time=0.0, camera_orbit_deg=0.0, light_orbit_deg=0.0):
        check(self._lib.gprt_renderer_on_update(self._h, float(animate_geometry_time), float(camera_orbit_deg), float(light_orbit_deg)),
              "on_update")

    def on_size_changed(self, width, height):
        check(self._lib.gprt_renderer_on_size_changed(self._h, int(width), int(height)), "on_size_changed")

    def _tile(self, tile):
        if tile is None:
            return Rect(0, 0, 0, 0), self.width, self.height
        x, y, w, h = tile
        return Rect(x, y, w, h), w, h

    def on_render(self, tile=None, want_hit_ids=False, out=None):
        """Render and copy to HOST memory: returns (rgba8[h,w,4] uint8, hit_ids[h,w] uint32 | None)."""
        rect, w, h = self._tile(tile)
        img = out if out is not None else np.empty((h, w, 4), dtype=np.uint8)
        ids = np.empty((h, w), dtype=np.uint32) if want_hit_ids else None
        check(self._lib.gprt_renderer_on_render(self._h, rect, img.ctypes.data, ids.ctypes.data if ids is not None else None), "on_render")
        return img, ids

    def do_raytracing(self, d_rgba8_ptr, tile=None, d_hit_ids_ptr=None, d_rgba_f32_ptr=None):
        """Asynchronous dispatch into caller-owned DEVICE memory (raw device pointers)."""
        rect, _, _ = self._tile(tile)
        check(self._lib.gprt_renderer_do_raytracing(self._h, rect, d_rgba8_ptr, d_hit_ids_ptr, d_rgba_f32_ptr), "do_raytracing")

    def do_raytracing_striped(self, d_rgba8_ptr, stripe_rows, rank, world, d_hit_ids_ptr=None):
        """Multi-GPU: render this rank's row stripes into a packed DEVICE buffer (asynchronous)."""
        check(self._lib.gprt_renderer_do_raytracing_stripes(self._h, int(stripe_rows), int(rank), int(world), d_rgba8_ptr, d_hit_ids_ptr),
              "do_raytracing_stripes")

    def on_render_async(self, h_rgba8_ptr):
        """on_render with 3 frames in flight: enqueue this frame (host output, pinned memory) and return."""
        check(self._lib.gprt_renderer_on_render_async(self._h, h_rgba8_ptr), "on_render_async")

    def do_raytracing_async(self, d_rgba8_ptr):
        """do_raytracing with 3 frames in flight (device output)."""
        check(self._lib.gprt_renderer_do_raytracing_async(self._h, d_rgba8_ptr), "do_raytracing_async")

    def do_raytracing_striped_async(self, d_rgba8_ptr, stripe_rows, rank, world):
        """Multi-GPU stripes with 3 frames in flight (device output, packed local rows)."""
        check(self._lib.gprt_renderer_do_raytracing_stripes_async(self._h, int(stripe_rows), int(rank), int(world), d_rgba8_ptr),
              "do_raytracing_stripes_async")

    def on_render_striped_async(self, h_frame_ptr, stripe_rows, rank, world):
        """Multi-GPU stripes with host output: this rank's rows land on their global rows of the shared (pinned / registered)
        full-frame host buffer at h_frame_ptr."""
        check(self._lib.gprt_renderer_on_render_stripes_async(self._h, int(stripe_rows), int(rank), int(world), h_frame_ptr),
              "on_render_stripes_async")

    def set_device_transforms(self, on: bool):
        """Per-frame transforms computed on the GPU (default for scenes of >= 4096 primitives) or on the host."""
        check(self._lib.gprt_renderer_set_device_transforms(self._h, 1 if on else 0), "set_device_transforms")

    def uses_device_transforms(self) -> bool:
        return bool(self._lib.gprt_renderer_get_device_transforms(self._h))

    def device_transforms(self) -> np.ndarray:
        """The transforms currently resident on the device of frame slot 0 (tests)."""
        n = len(self.scene().aabbs)
        out = np.zeros(n, dtype=_capi.PER_FRAME_DTYPE)
        check(self._lib.gprt_get_transforms(self.context, out.ctypes.data, n), "get_transforms")
        return out

    def create_frame_contexts(self):
        check(self._lib.gprt_renderer_create_frame_contexts(self._h), "create_frame_contexts")

    def set_frame_count(self, n: int):
        """Frames in flight for the *_async calls (default 3 = the reference's frame_count, src/Renderer.h:92)."""
        check(self._lib.gprt_renderer_set_frame_count(self._h, int(n)), "set_frame_count")

    @property
    def frame_count(self) -> int:
        return int(self._lib.gprt_renderer_get_frame_count(self._h))

    def next_frame_slot(self) -> int:
        return int(self._lib.gprt_renderer_next_frame_slot(self._h))

    def wait_for_gpu(self):
        check(self._lib.gprt_renderer_wait_for_gpu(self._h), "wait_for_gpu")

    def frame_stream(self, slot):
        """cudaStream_t (int) of frame slot `slot`, or 0 if the slot has not been used yet."""
        ctx = self._lib.gprt_renderer_frame_context(self._h, int(slot))
        return int(self._lib.gprt_get_stream(ctx) or 0) if ctx else 0

    def sync(self):
        check(self._lib.gprt_sync(self.context), "sync")

    @property
    def context(self):
        return self._lib.gprt_renderer_context(self._h)

    def set_stream(self, cuda_stream_ptr):
        check(self._lib.gprt_set_stream(self.context, cuda_stream_ptr), "set_stream")

    def counters(self) -> Counters:
        c = Counters()
        check(self._lib.gprt_get_counters(self.context, C.byref(c)), "get_counters")
        return c

    def timing(self) -> Timing:
        t = Timing()
        check(self._lib.gprt_last_timing(self.context, C.byref(t)), "last_timing")
        return t

    def kernel_profile(self) -> KernelProfile:
        p = KernelProfile()
        check(self._lib.gprt_get_kernel_profile(self.context, C.byref(p)), "get_kernel_profile")
        return p

    def wave_stats(self, wave) -> WaveStats:
        w = WaveStats()
        check(self._lib.gprt_get_wave_stats(self.context, int(wave), C.byref(w)), "get_wave_stats")
        return w

    def measure_fp32_peak(self) -> float:
        v = C.c_float(0)
        check(self._lib.gprt_measure_fp32_peak(self.context, C.byref(v)), "measure_fp32_peak")
        return float(v.value)

    def launch_count(self) -> int:
        return int(self._lib.gprt_last_launch_count(self.context))

    def primary_mrays_per_s(self) -> float:
        return float(self._lib.gprt_renderer_primary_mrays_per_s(self._h))

    def _view(self) -> SceneView:
        v = SceneView()
        check(self._lib.gprt_renderer_scene(self._h, C.byref(v)), "scene")
        return v

    def _size(self):
        w, h = C.c_uint32(0), C.c_uint32(0)
        check(self._lib.gprt_renderer_get_size(self._h, C.byref(w), C.byref(h)), "get_size")
        return int(w.value), int(h.value)

    @property
    def width(self):
        return self._size()[0]

    @property
    def height(self):
        return self._size()[1]

    def scene(self) -> SceneData:
        v = self._view()
        va = _capi.view_array
        cfg = Config()
        C.memmove(C.byref(cfg), C.byref(v.config), C.sizeof(Config))
        return SceneData(
            width=v.width, height=v.height, config=cfg,
            scene_cb=va(v.scene_cb, 1, _capi.SCENE_CB_DTYPE),
            per_frame=va(v.per_frame, v.n_per_frame, _capi.PER_FRAME_DTYPE),
            aabbs=va(v.aabbs, v.n_aabbs, _capi.AABB_DTYPE),
            hit_groups=va(v.hit_groups, v.n_hit_groups, _capi.HIT_GROUP_DTYPE),
            vertices=va(v.vertices, v.n_vertices, _capi.VERTEX_DTYPE),
            indices=va(v.indices, v.n_indices, np.dtype("<u2")),
            instances=va(v.instances, v.n_instances, _capi.INSTANCE_DTYPE),
        )

    def on_destroy(self):
        if self._h:
            self._lib.gprt_renderer_destroy(self._h)
            self._h = C.c_void_p()

    close = on_destroy

    def __del__(self):
        try:
            self.on_destroy()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.on_destroy()


class MultiRenderer:
    """One frame cut into interleaved row stripes over several CUDA devices from one process (gprt_multi_*,
    host/multi_renderer.h).  `devices` may name a device more than once: the entries are independent contexts."""

    def __init__(self, scene_id=0, width=0, height=0, grid_n=0, devices=(0,), config: Config = None, frames_in_flight=1):
        self._lib = lib()
        self._h = C.c_void_p()
        arr = (C.c_int * len(devices))(*[int(d) for d in devices])
        check(self._lib.gprt_multi_create(C.byref(self._h), int(scene_id), int(width), int(height), int(grid_n), arr, len(devices),
                                          C.byref(config) if config is not None else None), "multi_create")
        self.devices = list(devices)
        check(self._lib.gprt_multi_on_init(self._h), "multi_on_init")
        check(self._lib.gprt_multi_set_frame_count(self._h, int(frames_in_flight)), "multi_set_frame_count")
        w, h = C.c_uint32(0), C.c_uint32(0)
        check(self._lib.gprt_multi_get_size(self._h, C.byref(w), C.byref(h)), "multi_get_size")
        self.width, self.height = int(w.value), int(h.value)

    @property
    def stripe_rows(self):
        return int(self._lib.gprt_multi_stripe_rows(self._h))

    def on_update(self, animate_geometry_time=0.0, camera_orbit_deg=0.0, light_orbit_deg=0.0):
        check(self._lib.gprt_multi_on_update(self._h, float(animate_geometry_time), float(camera_orbit_deg), float(light_orbit_deg)), "multi_on_update")

    def on_render_async(self, h_frame_ptr):
        """Every device's rows to their global rows of the page-locked host frame at h_frame_ptr."""
        check(self._lib.gprt_multi_on_render_async(self._h, h_frame_ptr), "multi_on_render_async")

    def do_raytracing_async(self, d_frame_ptr):
        """The same into a device frame on devices[0] (peer copies)."""
        check(self._lib.gprt_multi_do_raytracing_async(self._h, d_frame_ptr), "multi_do_raytracing_async")

    def wait_for_gpu(self):
        check(self._lib.gprt_multi_wait_for_gpu(self._h), "multi_wait_for_gpu")

    def stream(self, i):
        """cudaStream_t (int) of entry i's frame slot 0."""
        ctx = self._lib.gprt_multi_context(self._h, int(i))
        return int(self._lib.gprt_get_stream(ctx) or 0) if ctx else 0

    def on_render(self):
        """Synchronous: returns the frame as a numpy array (rendered into page-locked memory)."""
        n = self.width * self.height * 4
        p = self._lib.gprt_alloc_host(n)
        if not p:
            raise MemoryError("gprt_alloc_host")
        try:
            self.on_render_async(p)
            self.wait_for_gpu()
            return np.frombuffer((C.c_char * n).from_address(p), dtype=np.uint8).reshape(self.height, self.width, 4).copy()
        finally:
            self._lib.gprt_free_host(p)

    def close(self):
        if self._h:
            self._lib.gprt_multi_destroy(self._h)
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


def write_ppm(path, rgba8: np.ndarray):
    h, w = rgba8.shape[:2]
    a = np.ascontiguousarray(rgba8)
    check(lib().gprt_write_ppm(str(path).encode(), a.ctypes.data, w, h), "write_ppm")
