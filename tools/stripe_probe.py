#!/usr/bin/env python3
"""What bounds one rank of an N-GPU striped render (development aid, ONE GPU, no NCCL):
renders rank 0's stripes of `world` with `slots` frames in flight and reports ms per frame
(wall clock over the whole run) and the host time spent submitting a frame.

    python tools/stripe_probe.py --scene 2
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gpuraytracer_b200 import Renderer, default_config, multigpu  # noqa: E402


def run(scene, world, slots, frames):
    cfg = default_config()
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = 0
    with Renderer(scene_id=scene, device=0, config=cfg) as r:
        r.on_init()
        r.set_frame_count(slots)
        r.create_frame_contexts()
        W, H = r.width, r.height
        srows = multigpu.stripe_rows_for(H, world)
        rows = multigpu.max_local_rows(H, srows, world) if world > 1 else H
        bufs = [torch.empty((rows, W, 4), dtype=torch.uint8, device="cuda") for _ in range(slots)]

        def go(k, first):
            host = 0.0
            for i in range(k):
                t0 = time.perf_counter()
                j = r.next_frame_slot()
                r.on_update((first + i) / 60.0)
                if world == 1:
                    r.do_raytracing_async(bufs[j].data_ptr())
                else:
                    r.do_raytracing_striped_async(bufs[j].data_ptr(), srows, 0, world)
                host += time.perf_counter() - t0
            return host

        go(2 * slots, 0)
        r.wait_for_gpu()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        host = go(frames, 2 * slots)
        r.wait_for_gpu()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        del bufs
    return {"world": world, "slots": slots, "ms_per_frame": round(dt / frames * 1e3, 4), "host_submit_ms": round(host / frames * 1e3, 4)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", type=int, default=2)
    ap.add_argument("--frames", type=int, default=60)
    a = ap.parse_args()
    for world in (1, 2, 4, 8):
        for slots in (1, 3, 6, 8):
            print(json.dumps(run(a.scene, world, slots, a.frames)), flush=True)


if __name__ == "__main__":
    main()
