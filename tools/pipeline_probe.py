#!/usr/bin/env python3
"""Throughput of N frames in flight (N independent contexts on N streams) vs one (development aid, GPU box)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gpuraytracer_b200 import Renderer, default_config

def run(n_inflight, frames=30, scene=2, host=False):
    cfg = default_config()
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = 0
    rs, streams, bufs, hosts = [], [], [], []
    for i in range(n_inflight):
        r = Renderer(scene_id=scene, device=0, config=cfg); r.on_init()
        s = torch.cuda.Stream(); r.set_stream(s.cuda_stream)
        rs.append(r); streams.append(s)
        bufs.append(torch.empty((r.height, r.width, 4), dtype=torch.uint8, device="cuda"))
        hosts.append(torch.empty((r.height, r.width, 4), dtype=torch.uint8, pin_memory=True))
    def go(k):
        for i in range(k):
            j = i % n_inflight
            rs[j].on_update(i / 60.0)
            rs[j].do_raytracing(bufs[j].data_ptr())
            if host:
                with torch.cuda.stream(streams[j]):
                    hosts[j].copy_(bufs[j], non_blocking=True)
    go(2 * n_inflight); torch.cuda.synchronize()
    t0 = time.perf_counter(); go(frames); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    for r in rs: r.on_destroy()
    return dt / frames * 1e3

for host in (False, True):
    print("host copies" if host else "device only", {n: round(run(n, host=host), 3) for n in (1, 2, 3, 4)})
