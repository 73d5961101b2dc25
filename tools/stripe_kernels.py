#!/usr/bin/env python3
"""Where the strong-scaling loss of a striped frame comes from (development aid, ONE GPU, no NCCL):
renders every rank's stripes of `world` one after the other with per-kernel events and compares the SUM over the
ranks with the whole frame rendered at once -- device time per kernel kind, candidates and field evaluations.

    python tools/stripe_kernels.py [--scene 5] [--world 8]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gpuraytracer_b200 import Renderer, default_config, multigpu  # noqa: E402
from gpuraytracer_b200._capi import GPRT_FLAG_PROFILE_KERNELS  # noqa: E402

KINDS = {0: "traverse", 11: "resolve", 12: "shade_spawn", 13: "final", 14: "intersect_all"}


def profile(r, buf, srows, rank, world, t):
    for _ in range(2):
        r.on_update(t)
        if world == 1:
            r.do_raytracing(buf.data_ptr())
        else:
            r.do_raytracing_striped(buf.data_ptr(), srows, rank, world)
        r.sync()
    p = r.kernel_profile()
    out = {name: float(p.ms[k]) for k, name in KINDS.items()}
    out["dispatch"] = float(r.timing().dispatch_ms)
    out["candidates"] = int(sum(p.candidates))
    out["work"] = int(sum(p.work))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", type=int, default=5)
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--stripe-rows", type=int, default=0)
    a = ap.parse_args()
    cfg = default_config()
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0
    cfg.flags = GPRT_FLAG_PROFILE_KERNELS
    with Renderer(scene_id=a.scene, device=0, config=cfg) as r:
        r.on_init()
        W, H = r.width, r.height
        buf = torch.empty((H, W, 4), dtype=torch.uint8, device="cuda")
        whole = profile(r, buf, H, 0, 1, 1.0)
        srows = a.stripe_rows or multigpu.stripe_rows_for(H, a.world)
        ranks = [profile(r, buf, srows, k, a.world, 1.0) for k in range(a.world)]
    total = {k: sum(x[k] for x in ranks) for k in whole}
    print(json.dumps({"scene": a.scene, "world": a.world, "stripe_rows": srows, "whole": whole, "sum_over_ranks": total,
                      "ratio": {k: round(total[k] / whole[k], 4) if whole[k] else None for k in whole},
                      "slowest_rank_dispatch": max(x["dispatch"] for x in ranks)}, indent=1))


if __name__ == "__main__":
    main()
