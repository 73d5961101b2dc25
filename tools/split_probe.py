#!/usr/bin/env python3
"""One frame as S concurrent stripe dispatches on ONE GPU (MultiRenderer with the device listed S times): wall-clock latency
of a frame, device output, L2 flushed, against the single-dispatch latency.  GPU box:  python tools/split_probe.py [--scene 2]"""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gpuraytracer_b200 import MultiRenderer, default_config

ap = argparse.ArgumentParser()
ap.add_argument("--scene", type=int, default=2)
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
cfg = default_config()
cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0
flush = torch.empty(192 << 20, dtype=torch.uint8, device="cuda")
for s in (1, 2, 3, 4, 6, 8):
    with MultiRenderer(scene_id=a.scene, devices=[0] * s, config=cfg, frames_in_flight=1) as m:
        frame = torch.zeros((m.height, m.width, 4), dtype=torch.uint8, device="cuda")
        ts = []
        for i in range(3 + a.reps):
            m.on_update(i / 60.0)
            flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            m.do_raytracing_async(frame.data_ptr())
            m.wait_for_gpu()
            ts.append((time.perf_counter() - t0) * 1e3)
        ts = sorted(ts[3:])
        print(f"contexts {s}: median {ts[len(ts)//2]:.3f} ms  min {ts[0]:.3f}  max {ts[-1]:.3f}", flush=True)
