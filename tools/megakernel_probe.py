#!/usr/bin/env python3
"""Wavefront form vs the one-kernel-per-frame comparison point (GPRT_FLAG_MEGAKERNEL) on the small-scene configs: device
time of one frame (gprt_last_timing, median of a few), same image.  GPU box:  python tools/megakernel_probe.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gpuraytracer_b200 import Renderer, default_config
from gpuraytracer_b200._capi import GPRT_FLAG_MEGAKERNEL

out = {}
for scene in (1, 2, 3, 4):
    row = {}
    imgs = []
    for name, flags in (("wavefront", 0), ("megakernel", GPRT_FLAG_MEGAKERNEL)):
        cfg = default_config()
        cfg.flags = flags
        cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0
        with Renderer(scene_id=scene, device=0, config=cfg) as r:
            r.on_init()
            ms = []
            for i in range(6):
                r.on_update(i / 60.0)
                img, _ = r.on_render()
                ms.append(r.timing().dispatch_ms)
            c = r.counters()
            imgs.append(img.copy())
            row[name + "_ms"] = round(sorted(ms[1:])[len(ms[1:]) // 2], 3)
            row["rays"] = int(c.radiance_rays) + int(c.shadow_rays)
            row["size"] = [r.width, r.height]
    row["same_image"] = bool((imgs[0] == imgs[1]).all())
    row["megakernel_over_wavefront"] = round(row["megakernel_ms"] / row["wavefront_ms"], 2)
    out[f"config{scene}"] = row
    print(f"config{scene}", row, flush=True)
json.dump(out, open(os.path.join("gpurun_out", "r02_megakernel_vs_wavefront.json"), "w"), indent=1)
