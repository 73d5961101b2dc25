import os, sys
sys.path.insert(0, "/root/repo")
os.environ["GPRT_DUMP_INTERVALS"] = "1"
import torch
from gpuraytracer_b200 import Renderer, default_config, multigpu
from gpuraytracer_b200._capi import GPRT_FLAG_PROFILE_KERNELS
cfg = default_config()
cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0
cfg.flags = GPRT_FLAG_PROFILE_KERNELS
world = int(sys.argv[1])
with Renderer(scene_id=5, device=0, config=cfg) as r:
    r.on_init()
    W, H = r.width, r.height
    srows = multigpu.stripe_rows_for(H, world)
    rows = multigpu.max_local_rows(H, srows, world) if world > 1 else H
    buf = torch.empty((rows, W, 4), dtype=torch.uint8, device="cuda")
    for i in range(3):
        r.on_update(1.0 + i / 60)
        if world == 1: r.do_raytracing(buf.data_ptr())
        else: r.do_raytracing_striped(buf.data_ptr(), srows, 0, world)
        r.sync()
    sys.stderr.write("=== last frame\n")
    p = r.kernel_profile()
    print("dispatch ms", r.timing().dispatch_ms)
