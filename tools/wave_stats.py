#!/usr/bin/env python3
"""Per-wave, per-type candidate cost statistics of one frame (development aid, GPU box):
how many field evaluations the candidates of each wave need (histogram by log2, maximum).
Runs the one-candidate-per-thread kernels (GPRT_ISECT_VARIANT=-1, per-type launches).

    GPRT_ISECT_VARIANT=-1 python tools/wave_stats.py --scene 2
"""
import argparse
import json
import os
import sys

if "--production" not in sys.argv:  # default: the one-candidate-per-thread statistics kernels; --production: the shipped kernels (tail counts)
    os.environ.setdefault("GPRT_ISECT_VARIANT", "-1")
    os.environ.setdefault("GPRT_PER_TYPE_LAUNCHES", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from gpuraytracer_b200 import Renderer, default_config  # noqa: E402
from gpuraytracer_b200._capi import GPRT_FLAG_PROFILE_KERNELS  # noqa: E402

TNAMES = ["box", "spheres", "metaballs", "minispheres", "roundcube", "sqtorus", "twtorus", "cog", "cylinder", "pyramid"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", type=int, default=2)
    ap.add_argument("--fast", action="store_true")
    ap.add_argument("--time", type=float, default=0.0)
    ap.add_argument("--production", action="store_true")
    a = ap.parse_args()
    cfg = default_config()
    cfg.math_mode = 1 if a.fast else 0
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = 0
    cfg.flags = GPRT_FLAG_PROFILE_KERNELS
    out = {"scene": a.scene}
    with Renderer(scene_id=a.scene, device=0, config=cfg) as r:
        r.on_init()
        r.on_update(a.time)
        r.on_render()
        depth = r.scene().config.max_recursion_depth
        for w in range(1, depth + 1):
            s = r.wave_stats(w)
            out[f"wave{w}"] = {"radiance": s.n_radiance, "shadow": s.n_shadow, "node_visits": s.node_visits, "max_node_visits": s.max_node_visits,
                               "types": {TNAMES[t]: {"cand": s.candidates[t], "work": s.work[t], "max": s.max_evals[t], "hist": list(s.evals_hist[t])}
                                         for t in range(10) if s.candidates[t]}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
