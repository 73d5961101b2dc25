"""gpuraytracer_b200 -- B200-native (sm_100a CUDA) implementation of GPURaytracer's DXR hot path.

Only the path named by BASELINE.json's north_star lives here: Renderer::do_raytracing ->
DispatchRays over the procedural-geometry shader library, plus build_acceleration_structures.
The product is libgprt.so (C ABI in include/gprt.h); this package is a thin ctypes layer.
"""
from ._capi import (GPRT_FLAG_BRUTE_FORCE_TRAVERSAL, GPRT_FLAG_MEGAKERNEL, GPRT_HIT_AABB, GPRT_HIT_MISS, GPRT_HIT_TRIANGLE, GPRT_MATH_FAST, GPRT_MATH_FMA,
                    GPRT_MATH_PARITY, Config, Counters, GprtError, Rect, Timing)
from .renderer import MultiRenderer, Renderer, SceneData, default_config, write_ppm

__all__ = ["Renderer", "MultiRenderer", "SceneData", "Config", "Counters", "Timing", "Rect", "GprtError", "default_config", "write_ppm",
           "GPRT_MATH_PARITY", "GPRT_MATH_FAST", "GPRT_FLAG_BRUTE_FORCE_TRAVERSAL", "GPRT_HIT_MISS", "GPRT_HIT_TRIANGLE",
           "GPRT_HIT_AABB"]
