#!/usr/bin/env python3
"""Mismatch fraction of the CUDA path against the CPU oracle for every BASELINE config (GPU box; a few seconds each).
    python tools/parity_report.py [--fast] > profiles/rNN_parity_report.json"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpuraytracer_b200 import Renderer, default_config
from oracle import oracle
from tests.parity import report

CASES = [("config1 640x480", 1, 0, 0, 0, 0.0), ("config2 1920x1080 t=0", 2, 0, 0, 0, 0.0), ("config2 1920x1080 t=3", 2, 0, 0, 0, 3.0),
         ("config3 960x540", 3, 960, 540, 0, 0.0), ("config3 long budget 480x270", 6, 480, 270, 0, 0.0),
         ("config4 1920x1080 t=2.25", 4, 0, 0, 0, 2.25), ("config4 1920x1080 t=9", 4, 0, 0, 0, 9.0),
         ("config5 grid 64 1920x1080", 5, 1920, 1080, 64, 1.0), ("config5 grid 256 (65536 primitives, spill rounds on) 1920x1080", 5, 1920, 1080, 256, 1.0),
         ("instanced scene 7 (3 AABB + 2 triangle instances) 1920x1080 t=1", 7, 1920, 1080, 0, 1.0)]
out = {"math_mode": "fast" if "--fast" in sys.argv else "parity", "rule": ">= 99.9 % of pixels within 2/255 per channel; interior hit ids equal", "cases": {}}
for name, sid, w, h, g, t in CASES:
    cfg = default_config()
    cfg.math_mode = 1 if "--fast" in sys.argv else 0
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = 0
    with Renderer(scene_id=sid, width=w, height=h, grid_n=g, device=0, config=cfg) as r:
        r.on_init(); r.on_update(t)
        img, ids = r.on_render(want_hit_ids=True)
        scene = r.scene()
    orc = oracle.load(metaballs=scene.config.metaballs_count)
    orc.set_scene(scene)
    ref = orc.render(scene)
    out["cases"][name] = report(ref["rgba8"], img, ref["hit_ids"], ids)
print(json.dumps(out, indent=1))
