#!/usr/bin/env python3
"""Mismatch fraction of the CUDA path against the CPU oracle for every BASELINE config AT ITS BASELINE SIZE (GPU box).

    python tools/parity_report.py [--math parity,fma,fast] [--quick] [--out-dir profiles] [--tag r02]

writes <out-dir>/<tag>_parity_report_<mode>.json per math mode.  The oracle image of a case is rendered once and compared
with every mode.  Cases: config 1 640x480; config 2 1920x1080 at t = 0 and 3; config 3 3840x2160 and its
long-budget variant; config 4 1920x1080 at all 16 times k * 0.75 s; config 5 7680x4320 (1 M primitives) on 12 tiles of
640x160 spread over the frame (the oracle walks its own host BVH there); the two-level-instancing scene 7.
--quick shrinks config 3 and the tile/time counts (development runs)."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpuraytracer_b200 import Renderer, default_config  # noqa: E402
from oracle import oracle  # noqa: E402
from tests.parity import report  # noqa: E402

MODES = {"parity": 0, "fast": 1, "fma": 2}


def cases(quick):
    out = [("config1 640x480", 1, 0, 0, 0, 0.0, None), ("config2 1920x1080 t=0", 2, 0, 0, 0, 0.0, None), ("config2 1920x1080 t=3", 2, 0, 0, 0, 3.0, None)]
    if quick:
        out += [("config3 960x540", 3, 960, 540, 0, 0.0, None), ("config3 long budget 480x270", 6, 480, 270, 0, 0.0, None)]
    else:
        out += [("config3 3840x2160", 3, 0, 0, 0, 0.0, None), ("config3 long budget 1920x1080", 6, 1920, 1080, 0, 0.0, None)]
    for k in (range(0, 16, 5) if quick else range(16)):
        out.append((f"config4 1920x1080 t={k * 0.75:g}", 4, 0, 0, 0, k * 0.75, None))
    # config 5 at 7680x4320: tiles on a 4 x 3 lattice from the horizon rows to the bottom edge
    xs, ys = (600, 2600, 4600, 6600), (900, 2100, 3700)
    tiles = [(x, y, 640, 160) for y in ys for x in xs]
    for tile in (tiles[::4] if quick else tiles):
        out.append((f"config5 7680x4320 1M primitives tile x={tile[0]} y={tile[1]} {tile[2]}x{tile[3]}", 5, 0, 0, 0, 1.0, tile))
    out.append(("config5 grid 256 (65536 primitives) 1920x1080", 5, 1920, 1080, 256, 1.0, None))
    out.append(("instanced scene 7 (3 AABB + 2 triangle instances) 1920x1080 t=1", 7, 1920, 1080, 0, 1.0, None))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--math", default="parity")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--out-dir", default="gpurun_out")
    ap.add_argument("--tag", default="r02")
    a = ap.parse_args()
    modes = [m for m in a.math.split(",") if m]
    outs = {m: {"math_mode": m, "rule": ">= 99.9 % of pixels within 2/255 per channel; interior hit ids equal", "cases": {}} for m in modes}
    renderers = {}
    oracles = {}
    last_key = None
    for name, sid, w, h, g, t, tile in cases(a.quick):
        key = (sid, w, h, g)
        if key != last_key:  # a new scene: drop the previous renderers (the 1 M primitive scene holds a few GB each)
            for r in renderers.values():
                r.on_destroy()
            renderers = {}
            last_key = key
        ref = None
        for m in modes:
            if m not in renderers:
                cfg = default_config()
                cfg.math_mode = MODES[m]
                cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0
                renderers[m] = Renderer(scene_id=sid, width=w, height=h, grid_n=g, device=0, config=cfg)
                renderers[m].on_init()
            r = renderers[m]
            r.on_update(t)
            img, ids = r.on_render(tile=tile, want_hit_ids=True)
            if ref is None:
                scene = r.scene()
                mb = scene.config.metaballs_count
                if key not in oracles:
                    oracles.clear()
                    oracles[key] = oracle.load(metaballs=mb)
                    oracles[key].set_scene(scene)
                t0 = time.time()
                ref = oracles[key].render(scene, tile=tile)
                cpu_s = time.time() - t0
            rep = report(ref["rgba8"], img, ref["hit_ids"], ids)
            rep["pixels"] = int(img.shape[0] * img.shape[1])
            rep["oracle_seconds"] = round(cpu_s, 2)
            outs[m]["cases"][name] = rep
            print(m, name, rep, file=sys.stderr, flush=True)
    for r in renderers.values():
        r.on_destroy()
    os.makedirs(a.out_dir, exist_ok=True)
    for m in modes:
        worst = max(c["mismatch_frac_gt2"] for c in outs[m]["cases"].values())
        outs[m]["worst_mismatch_frac_gt2"] = worst
        outs[m]["all_within_rule"] = bool(worst <= 1e-3 and all(c["id_mismatch_interior"] == 0 for c in outs[m]["cases"].values()))
        path = os.path.join(a.out_dir, f"{a.tag}_parity_report_{m}.json")
        json.dump(outs[m], open(path, "w"), indent=1)
        print(path, "worst", worst, "ok", outs[m]["all_within_rule"])


if __name__ == "__main__":
    main()
