#!/usr/bin/env python3
"""bench.py -- the driver's benchmark contract for the DXR hot path (BASELINE.json metric:
Mrays/s and 1080p frame ms, primary + shadow + reflection rays, recursion 3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config ID]

One "step" = one DispatchRays pass over one frame of the workload (config 2: the reference's full
procedural scene, 1920x1080, depth 3; elapsed_time advances 1/60 s per step so every step has new
per-frame transforms).  `value` = rays traced (radiance + shadow, counted on the device) per
second with the scene resident in HBM and the output left in HBM; `e2e` = the same through the
host-facing call (host scene constants in, host RGBA8 frame out, copies inside the timed region).
N > 1: one process per GPU (torchrun).  The 8K config is sharded by interleaved row stripes: every rank
renders its stripes of every frame and rank 0 gathers the RGBA8 rows over NCCL (strong scaling: total
work fixed).  The 1080p configs are sharded by frames (frame g on rank g % N, gathered to rank 0; a step
is N frames, weak scaling): a 2 ms frame has a 1.5 ms serial floor and cannot be split 8 ways.
`--impl reference` times the reference's own shader code on the host cores (oracle/_ref).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mrays_per_s"
UNIT = "Mrays/s (radiance + shadow rays)"

CONFIGS = {
    1: {"scene_id": 1, "workload": "config1: analytic sphere cluster + ground quad, 640x480, depth 2 (primary + shadow)"},
    2: {"scene_id": 2, "workload": "config2: full reference procedural scene, 1920x1080, recursion 3"},
    3: {"scene_id": 3, "workload": "config3: SDF/fractal-heavy 4x4 grid, 3840x2160, recursion 3"},
    4: {"scene_id": 4, "workload": "config4: 3x3 metaball AABBs with 5 animated metaballs, 1920x1080, recursion 3"},
    5: {"scene_id": 5, "workload": "config5: 1000x1000 instanced AABB grid (1M primitives), 7680x4320, recursion 3"},
}

# Algorithmic FP32 work per counted unit (FMA = 2, everything else = 1),
# per kernel type id: analytic box / spheres per test, metaball per march step (3 balls outside
# their radius + position update), SDF per distance evaluation incl. the 9-flop march overhead.
FLOPS_PER_UNIT = [30 + 42, 3 * 35 + 42, 3 * 11 + 6, 48 + 9, 34 + 9, 16 + 9, 20 + 9, 71 + 9, 66 + 9, 228 + 9]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        for row in self.rows:
            f = [x.strip() for x in row.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        # median of the upper half: the samples taken while kernels were running
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": load[len(load) // 2] if load else None, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm)}


def run_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from gpuraytracer_b200 import GPRT_MATH_FAST, GPRT_MATH_PARITY, Renderer, default_config, multigpu

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    conf = CONFIGS[args.config]
    cfg = default_config()
    cfg.math_mode = GPRT_MATH_PARITY if args.math == "parity" else GPRT_MATH_FAST
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = 0  # the scene's own
    r = Renderer(scene_id=conf["scene_id"], device=local_rank, config=cfg)
    r.on_init()
    W, H = r.width, r.height
    # N > 1: "frames" = every rank renders whole frames (frame g on rank g % N), "stripes" = every rank renders its row
    # stripes of every frame.  A frame has a serial floor (3 waves x the longest ray's dependent chain, ~1.5 ms at any
    # size), so a 2 ms frame cannot be split 8 ways, while a 50 ms one can: frames for the small configs, stripes for 8K.
    mode = "single" if world == 1 else (args.multi_gpu if args.multi_gpu != "auto" else ("frames" if W * H <= (1 << 22) else "stripes"))
    n_slots = args.frames_in_flight or (min(8, 2 + world) if mode == "stripes" else 3)
    r.set_frame_count(n_slots)  # the reference has frame_count = 3 (src/Renderer.h:92); a GPU that renders only a stripe needs more to hide a frame's serial tail
    r.create_frame_contexts()
    frames_per_step = world if mode == "frames" else 1

    def frame_time(i):
        """elapsed_time of this rank's frame of step i"""
        return ((i * world + rank) if mode == "frames" else i) / 60.0
    n_prims = len(r.scene().aabbs)
    main = torch.cuda.current_stream()
    ext = [torch.cuda.ExternalStream(r.frame_stream(j), device=dev) for j in range(n_slots)]

    srows = multigpu.stripe_rows_for(H, world) if mode == "stripes" else H
    pad_rows = multigpu.max_local_rows(H, srows, world) if mode == "stripes" else H
    local = [torch.empty((pad_rows, W, 4), dtype=torch.uint8, device=dev) for _ in range(n_slots)]
    # rank 0: one contiguous [world, rows, W, 4] block per slot, so that the frames of a step leave in one D2H copy
    blocks = [torch.empty((world, pad_rows, W, 4), dtype=torch.uint8, device=dev) if (world > 1 and rank == 0) else None for _ in range(n_slots)]
    gathered = [list(b.unbind(0)) if b is not None else None for b in blocks]
    full = [torch.empty((H, W, 4), dtype=torch.uint8, device=dev) if (mode == "stripes" and rank == 0) else local[j] for j in range(n_slots)]
    host = [torch.empty((frames_per_step, H, W, 4), dtype=torch.uint8, pin_memory=True) for _ in range(n_slots)] if rank == 0 else [None] * n_slots
    flush = [torch.empty(192 << 20, dtype=torch.uint8, device=dev) for _ in range(n_slots)]  # each > the 126 MB L2

    def step(i, to_host, flush_l2):
        """One step = one frame: update (host), [L2 flush], dispatch, [gather + compose on rank 0], [D2H].  Asynchronous:
        frame i runs on frame slot i % n_slots (own context, own stream); the *_async call first waits for that slot's previous frame."""
        j = r.next_frame_slot()
        r.on_update(frame_time(i))
        with torch.cuda.stream(ext[j]):
            if flush_l2:
                flush[j].zero_()
            if mode == "single":
                if to_host:
                    r.on_render_async(host[j].data_ptr())  # host scene constants in, pinned host frame out
                else:
                    r.do_raytracing_async(local[j].data_ptr())
            elif mode == "frames":
                r.do_raytracing_async(local[j].data_ptr())
                multigpu.gather_frames(local[j], gathered[j], dst=0)  # frame i * N + k from rank k
                if rank == 0 and to_host:
                    host[j].copy_(blocks[j], non_blocking=True)
            else:
                r.do_raytracing_striped_async(local[j].data_ptr(), srows, rank, world)
                multigpu.gather_stripes(local[j], gathered[j], dst=0)
                if rank == 0:
                    multigpu.compose(full[j], gathered[j], H, srows, world)
                    if to_host:
                        host[j][0].copy_(full[j], non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def timed(steps, first_step, to_host, flush_l2):
        """EXACTLY `steps` steps between two barriers; device time by CUDA events that every frame stream is ordered after /
        before (max over ranks); for the host-output path also the host wall clock, whichever is larger."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(main)
        for s in ext:
            s.wait_event(e0)
        for i in range(steps):
            step(first_step + i, to_host, flush_l2)
        for s in ext:
            main.wait_stream(s)
        e1.record(main)
        e1.synchronize()
        r.wait_for_gpu()
        wall_ms = (time.perf_counter() - t0) * 1e3
        ms = e0.elapsed_time(e1)
        if to_host:
            ms = max(ms, wall_ms)
        barrier()
        return max_over_ranks(ms)

    def single_frame_latency(steps, first_step):
        """Strictly serial: one blocking frame at a time (no frames in flight), L2 flushed before each, per-frame CUDA events."""
        total = 0.0
        for i in range(steps):
            r.on_update(frame_time(first_step + i))
            flush[0].zero_()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            with torch.cuda.stream(ext[0]):
                ext[0].wait_event(e0)
                if mode == "single":
                    r.do_raytracing(local[0].data_ptr())
                elif mode == "frames":
                    r.do_raytracing(local[0].data_ptr())
                    multigpu.gather_frames(local[0], gathered[0], dst=0)
                else:
                    r.do_raytracing_striped(local[0].data_ptr(), srows, rank, world)
                    multigpu.gather_stripes(local[0], gathered[0], dst=0)
                    if rank == 0:
                        multigpu.compose(full[0], gathered[0], H, srows, world)
            main.wait_stream(ext[0])
            e1.record(main)
            e1.synchronize()
            total += max_over_ranks(e0.elapsed_time(e1))
        return total / steps

    def count_rays(steps, first_step):
        """Rays traced (device counters) by exactly the frames of the timed passes (untimed, synchronous)."""
        rays = 0
        for i in range(steps):
            r.on_update(frame_time(first_step + i))
            if mode != "stripes":
                r.do_raytracing(local[0].data_ptr())
            else:
                r.do_raytracing_striped(local[0].data_ptr(), srows, rank, world)
            c = r.counters()
            rays += int(c.radiance_rays) + int(c.shadow_rays)
        if world > 1:
            t = torch.tensor([rays], dtype=torch.int64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            rays = int(t.item())
        return rays

    # every frame slot allocates its wavefront queues on first use: prime all of them before the W warm-up steps
    for i in range(n_slots):
        step(i, True, True)
    for i in range(args.warmup):
        step(i, False, True)
    for i in range(args.warmup):
        step(i, True, False)
    r.wait_for_gpu()
    barrier()
    launches_per_step = r.launch_count()
    rays = count_rays(args.steps, args.warmup)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    dev_ms = timed(args.steps, args.warmup, False, True)
    e2e_ms = timed(args.steps, args.warmup, True, False)
    lat_ms = single_frame_latency(min(args.steps, 10), args.warmup)
    clocks = sampler.stop() if rank == 0 else None

    result = None
    if rank == 0:
        value = rays / (dev_ms * 1e-3) / 1e6
        par = {"single": "1 GPU",
               "frames": f"{world} GPUs, whole frames round-robin (frame g on rank g % {world}), replicated scene, NCCL gather of the RGBA8 frames to rank 0; "
                         f"a step is {world} frames, one per rank",
               "stripes": f"{world} GPUs, {srows}-row stripes interleaved, replicated scene, NCCL gather of the RGBA8 rows to rank 0"}[mode]
        result = {
            "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(dev_ms / args.steps, 4), "higher_is_better": True, "scaling": "weak" if mode == "frames" else "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": conf["workload"], "width": W, "height": H, "math_mode": args.math,
                       "frames_in_flight": n_slots, "multi_gpu": mode, "frames_per_step": frames_per_step,
                       "l2_flush": "192 MiB memset on the frame's stream before every timed step (and the 0.6 GB wavefront state of a frame exceeds L2)",
                       "timing": "CUDA events ordered around all frame streams, exactly K steps between barriers, max over ranks",
                       "parallelism": par},
            "frame_ms": round(dev_ms / args.steps / frames_per_step, 4),
            "primary_mrays_per_s": round(W * H * frames_per_step * args.steps / (dev_ms * 1e-3) / 1e6, 3),
            "rays_per_step": rays // args.steps,
            "single_frame_ms": round(lat_ms, 4),
            "single_frame_how": "one frame at a time (nothing in flight), L2 flushed, CUDA events per frame: the latency of a 1080p frame",
            "e2e": {"value": round(rays / (e2e_ms * 1e-3) / 1e6, 3), "unit": UNIT, "ms_per_step": round(e2e_ms / args.steps, 4),
                    "h2d_bytes_per_step": 136 + (0 if r.uses_device_transforms() else 128 * n_prims), "d2h_bytes_per_step": W * H * 4 * frames_per_step,
                    "transforms": "computed on the GPU (gprt_update_transforms)" if r.uses_device_transforms() else "computed on the host, uploaded every step",
                    "how": f"Renderer.on_update + on_render_async per step (public API, {n_slots} frames in flight): host scene constants in, pinned host "
                           "RGBA8 frame out every step; max(CUDA events, host wall clock) around all steps + wait_for_gpu"
                           + {"single": "", "frames": "; every step's frames gathered to rank 0 over NCCL, then copied to the host in one block",
                              "stripes": "; stripes gathered to rank 0 over NCCL, composed, then copied to the host"}[mode]},
            "gpu_launches": launches_per_step * (2 * args.steps + min(args.steps, 10)),  # the two timed passes + the single-frame latency pass
            "clocks": clocks,
        }
        result["roofline"], result["roofline_hbm"], result["kernel_ms"] = roofline(args, conf, local_rank, cfg, world)
        if world == 1 and not args.no_cpu:
            result["cpu_baseline"] = cpu_baseline(conf, steps=2)
    # free every tensor that was used on the frame streams BEFORE the contexts (and their streams) are destroyed:
    # the caching allocator records an event on those streams when such a block is released
    barrier()
    del local, gathered, blocks, full, host, flush
    import gc
    gc.collect()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    del ext
    r.on_destroy()
    return result


def roofline(args, conf, device, cfg, world):
    """The dominant kernel of the step is the intersection kernel (all primitive types in one
    persistent launch per wave).  Its launches are timed live with CUDA events on the launching
    stream (GPRT_FLAG_PROFILE_KERNELS) over full single-GPU frames of the same workload, L2 flushed
    between frames; achieved = algorithmic FP32 work of those launches (device work counters x the
    hand-counted flops per unit of DESIGN.md section 5) / their duration, against the FP32-FMA
    issue-rate peak measured in this run.  A second object does the same for the most
    memory-bound kernel (the final fold) against the measured HBM copy bandwidth."""
    import torch
    from gpuraytracer_b200 import Renderer
    from gpuraytracer_b200._capi import GPRT_FLAG_PROFILE_KERNELS, GPRT_K_FINAL, GPRT_K_INTERSECT_ALL, KERNEL_NAMES
    cfg.flags = GPRT_FLAG_PROFILE_KERNELS
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=torch.device("cuda", device))
    with Renderer(scene_id=conf["scene_id"], device=device, config=cfg) as r:
        r.on_init()
        W, H = r.width, r.height
        nk = len(KERNEL_NAMES)
        ms = [0.0] * nk
        work = [0] * 10
        launches = [0] * nk
        reps = 5
        for i in range(2 + reps):
            r.on_update((args.warmup + i) / 60.0)
            flush.zero_()
            torch.cuda.synchronize()
            r.on_render()
            if i < 2:
                continue
            p = r.kernel_profile()
            for k in range(nk):
                ms[k] += p.ms[k]
                launches[k] += p.launches[k]
            for t in range(10):
                work[t] += p.work[t]
        c = r.counters()
        depth = int(r.scene().config.max_recursion_depth)
        fp32_peak = r.measure_fp32_peak()
    cfg.flags = 0
    ms = [m / reps for m in ms]
    work = [w / reps for w in work]
    launches = [n // reps for n in launches]
    kernel_ms = {KERNEL_NAMES[k]: round(ms[k], 4) for k in range(nk) if launches[k]}
    dom = GPRT_K_INTERSECT_ALL
    flops = sum(work[t] * FLOPS_PER_UNIT[t] for t in range(10))
    dur = ms[dom] * 1e-3
    n_l = max(launches[dom], 1)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    achieved = flops / dur / 1e12 if dur > 0 else 0.0
    peak_src = "measured in this run: FFMA issue-rate probe (gprt_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure"
    if not fp32_peak or fp32_peak <= 0:
        fp32_peak, peak_src = 74.4, "nominal fallback: 148 SMs x 128 lanes x 2 x 1.965 GHz"
    # DRAM traffic per launch of the same kernel from the committed ncu --set full capture (profiles/), mean over the
    # three launches (waves) of one frame of this workload; null when no capture is committed for this config
    traffic, traffic_src = None, None
    try:
        if conf["scene_id"] == 2:
            cap = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_full_k_intersect_coop.json")))
            vals = [l["dram_bytes_total"] for l in cap["launches"] if "dram_bytes_total" in l]
            traffic, traffic_src = round(sum(vals) / len(vals)), "profiles/r01_ncu_full_k_intersect_coop.json (dram__bytes_read.sum + dram__bytes_write.sum)"
    except Exception:
        pass
    roof = {"bound": "fp32", "kernel": "k_intersect_coop (all primitive types, one persistent launch per wave)",
            "achieved": round(achieved, 3), "peak": round(fp32_peak, 2), "unit": "TFLOP/s",
            "frac": round(achieved / fp32_peak, 4), "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
            "flops_per_launch": round(flops / n_l), "launches_per_step": n_l, "avg_launch_ms": round(ms[dom] / n_l, 4),
            "share_of_step": round(ms[dom] / max(sum(ms), 1e-9), 3),
            "units_per_step": {KERNEL_NAMES[1 + t]: round(work[t]) for t in range(10) if work[t]},
            "flops_per_unit": {KERNEL_NAMES[1 + t]: FLOPS_PER_UNIT[t] for t in range(10) if work[t]},
            "note": "not a tensor-core path: the march kernels are FP32-issue bound (SURVEY.md 8d); the peak counts FMA as 2 flops while "
                    "the IEEE parity build issues separate FMUL/FADD, so its ceiling is half of this peak"}
    # memory-bound kernel: final fold = per pixel 4 B colour out + 4 B level code per level, plus 49 B of hit record per radiance hit
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    fin_bytes = W * H * (4 + depth * 4) + int(c.radiance_rays) * 49
    fin_dur = ms[GPRT_K_FINAL] * 1e-3
    fin_traffic, fin_traffic_src = None, None
    try:
        if conf["scene_id"] == 2:
            cap = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_full_k_final_color.json")))
            fin_traffic = round(cap["launches"][0]["dram_bytes_total"])
            fin_traffic_src = "profiles/r01_ncu_full_k_final_color.json (dram__bytes_read.sum + dram__bytes_write.sum; the hit records shade_spawn just wrote are partly still in L2)"
    except Exception:
        pass
    hbm = {"bound": "hbm", "kernel": "k_final_color", "achieved": round(fin_bytes / fin_dur / 1e9, 1) if fin_dur > 0 else 0.0, "peak": hbm_peak,
           "unit": "GB/s", "frac": round(fin_bytes / fin_dur / 1e9 / hbm_peak, 4) if fin_dur > 0 else 0.0, "traffic": fin_traffic,
           "traffic_source": fin_traffic_src, "peak_source": hbm_src, "bytes_per_launch": fin_bytes}
    return roof, hbm, kernel_ms


def oracle_for(conf, fast=True):
    from gpuraytracer_b200 import Renderer
    from oracle import oracle
    r = Renderer(scene_id=conf["scene_id"], device=None)
    scene = r.scene()
    mb = scene.config.metaballs_count
    orc = oracle.load(metaballs=mb, fast=fast) if oracle.available("reference", mb, fast) else oracle.load(metaballs=mb)
    return r, orc


def cpu_frames(conf, steps, first_step=0, tile_rows=None):
    """Time `steps` frames (or row bands) of the workload on the host cores with the oracle."""
    r, orc = oracle_for(conf)
    total_s, rays, pixels = 0.0, 0, 0
    for i in range(steps):
        r.on_update((first_step + i) / 60.0)
        scene = r.scene()
        orc.set_scene(scene)
        tile = None if tile_rows is None else (0, (scene.height - tile_rows) // 2, scene.width, tile_rows)
        t0 = time.perf_counter()
        res = orc.render(scene, tile=tile, want_ids=False)
        total_s += time.perf_counter() - t0
        rays += res["counters"]["radiance_rays"] + res["counters"]["shadow_rays"]
        pixels += (scene.width * (tile_rows or scene.height))
    return total_s, rays, pixels, orc


def cpu_baseline(conf, steps=2):
    total_s, rays, pixels, orc = cpu_frames(conf, steps)
    return {"value": round(rays / total_s / 1e6, 3), "unit": UNIT, "cores": orc.max_threads(), "kind": orc.kind,
            "sample": f"{steps} full frames of the same workload ({pixels} pixels), OpenMP over rows, {os.path.basename(orc.path)}",
            "ms_per_step": round(total_s / steps * 1e3, 2)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    conf = CONFIGS[args.config]
    for _ in range(min(args.warmup, 1)):
        cpu_frames(conf, 1)
    total_s, rays, pixels, orc = cpu_frames(conf, args.steps, first_step=args.warmup)
    value = rays / total_s / 1e6
    return {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_s / args.steps * 1e3, 3), "higher_is_better": True,
            # the same label the GPU arm prints for this N and config (frames are dealt out up to 4 Mpixel, stripes above)
            "scaling": "weak" if (args.gpus > 1 and args.multi_gpu != "stripes" and (args.multi_gpu == "frames" or args.config in (1, 2, 4))) else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": conf["workload"]},
            "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": orc.max_threads(), "kind": orc.kind,
                             "sample": f"{args.steps} full frames ({pixels} pixels), OpenMP over rows, {os.path.basename(orc.path)}"},
            "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gprt", choices=["gprt", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS))
    ap.add_argument("--frames-in-flight", type=int, default=0, help="frame slots (contexts + streams); 0 = 3 on one GPU, min(8, 2 + N) on N")
    ap.add_argument("--multi-gpu", default="auto", choices=["auto", "frames", "stripes"],
                    help="N > 1: whole frames round-robin over the ranks, or row stripes of every frame; auto = frames up to 4 Mpixel, stripes above")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (development runs of the large configs)")
    ap.add_argument("--math", default=os.environ.get("GPRT_BENCH_MATH", "parity"), choices=["parity", "fast"])
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "gprt":
        args.warmup = 3
    # stdout carries exactly ONE JSON line: anything libraries print there (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        res = run_reference(args) if args.impl == "reference" else run_gpu(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if res is not None:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
