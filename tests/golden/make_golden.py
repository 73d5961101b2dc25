#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the reference-compiled oracle (oracle/_ref, i.e. the
reference's own HLSL shader sources compiled as C++ in THIS container, where /root/reference
exists).  The reference ships no golden vectors of its own, so these
fixtures are the pin: the GPU path, the oracle port and any later rebuild of oracle/_ref are
compared against them.

    python tests/golden/make_golden.py [fixture names: default all]

Each fixture: the exact scene inputs are rebuilt from (scene_id, width, height, grid_n, time) by
the host scene builder; stored are the RGBA8 image, the primary hit ids and the work counters.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from gpuraytracer_b200 import Renderer  # noqa: E402
from oracle import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# name: (scene_id, width, height, grid_n, animation time, tile or None)
FIXTURES = {
    "ref_scene_t0": (0, 128, 72, 0, 0.0, None),
    "ref_scene_t3": (0, 128, 72, 0, 3.0, None),
    "config1": (1, 96, 72, 0, 0.0, None),
    "config2_crop": (2, 1920, 1080, 0, 0.5, (860, 420, 96, 64)),
    "config3": (3, 128, 72, 0, 0.0, None),
    "config4_t0": (4, 96, 54, 0, 0.0, None),
    "config4_t225": (4, 96, 54, 0, 2.25, None),
    "config5_grid24": (5, 128, 72, 24, 1.0, None),
    "instanced_t1": (7, 160, 90, 0, 1.0, None),  # two-level instancing: 3 AABB instances, 2 triangle instances
}


def render_fixture(name):
    scene_id, w, h, grid_n, t, tile = FIXTURES[name]
    r = Renderer(scene_id=scene_id, width=w, height=h, grid_n=grid_n, device=None)
    r.on_update(t)
    scene = r.scene()
    orc = oracle.load(kind="reference", metaballs=scene.config.metaballs_count)
    orc.set_scene(scene)
    res = orc.render(scene, tile=tile)
    return scene, res


def main():
    names = sys.argv[1:] or list(FIXTURES)
    meta_path = os.path.join(HERE, "golden_meta.json")
    meta = json.load(open(meta_path)) if (sys.argv[1:] and os.path.exists(meta_path)) else {}
    for name in names:
        scene, res = render_fixture(name)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), rgba8=res["rgba8"], hit_ids=res["hit_ids"])
        meta[name] = {"params": list(FIXTURES[name]), "counters": res["counters"]}
        print(name, res["rgba8"].shape, res["counters"]["radiance_rays"], res["counters"]["shadow_rays"])
    with open(meta_path, "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
