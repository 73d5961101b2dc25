#!/usr/bin/env python3
"""Per-kernel device times of one frame (CUDA events around every launch) plus parity against the
CPU oracle.  Development / profiling aid; run on a GPU box:

    python tools/profile_frame.py --scene 2 [--fast] [--no-oracle] [--width W --height H]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from gpuraytracer_b200 import Renderer, default_config  # noqa: E402
from gpuraytracer_b200._capi import GPRT_FLAG_PROFILE_KERNELS  # noqa: E402

from gpuraytracer_b200._capi import KERNEL_NAMES as KNAMES  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", type=int, default=2)
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--grid-n", type=int, default=0)
    ap.add_argument("--fast", action="store_true")
    ap.add_argument("--no-oracle", action="store_true")
    ap.add_argument("--time", type=float, default=0.0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--save", default="")
    ap.add_argument("--per-type", action="store_true", help="one intersection launch per primitive type (GPRT_PER_TYPE_LAUNCHES=1)")
    a = ap.parse_args()
    if a.per_type:
        os.environ["GPRT_PER_TYPE_LAUNCHES"] = "1"

    cfg = default_config()
    cfg.math_mode = 1 if a.fast else 0
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = 0  # keep the scene's
    out = {"scene": a.scene, "fast": a.fast}
    for flags, label in ((0, "plain"), (GPRT_FLAG_PROFILE_KERNELS, "profiled")):
        cfg.flags = flags
        with Renderer(scene_id=a.scene, width=a.width, height=a.height, grid_n=a.grid_n, device=0, config=cfg) as r:
            r.on_init()
            r.on_update(a.time)
            ms = []
            for _ in range(a.reps):
                img, ids = r.on_render(want_hit_ids=True)
                ms.append(r.timing().dispatch_ms)
            out[label + "_dispatch_ms"] = [round(m, 4) for m in ms]
            c = r.counters()
            out["rays"] = [int(c.radiance_rays), int(c.shadow_rays)]
            out["launches"] = r.launch_count()
            out["build_ms"] = r.timing().build_accel_ms
            if flags:
                p = r.kernel_profile()
                out["kernel_ms"] = {KNAMES[i]: round(p.ms[i], 4) for i in range(len(KNAMES)) if p.launches[i]}
                out["work"] = {KNAMES[1 + t]: int(p.work[t]) for t in range(10) if p.candidates[t]}
                out["candidates"] = {KNAMES[1 + t]: int(p.candidates[t]) for t in range(10) if p.candidates[t]}
                out["fp32_peak_tflops"] = round(r.measure_fp32_peak(), 2)
            scene = r.scene()
    if a.save:
        np.save(a.save, img)
    if not a.no_oracle:
        from oracle import oracle
        from tests.parity import report
        orc = oracle.load(metaballs=scene.config.metaballs_count)
        orc.set_scene(scene)
        t0 = time.time()
        ref = orc.render(scene)
        out["oracle_s"] = round(time.time() - t0, 3)
        out["oracle_threads"] = orc.max_threads()
        out["parity"] = report(ref["rgba8"], img, ref["hit_ids"], ids)
        out["oracle_counters"] = ref["counters"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
