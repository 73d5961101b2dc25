#!/usr/bin/env python3
"""bench.py -- the driver's benchmark contract for the DXR hot path (BASELINE.json metric:
Mrays/s and 1080p frame ms, primary + shadow + reflection rays, recursion 3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config ID]

One "step" = one DispatchRays pass over one frame of the workload (config 2: the reference's full
procedural scene, 1920x1080, depth 3; elapsed_time advances 1/60 s per step so every step has new
per-frame transforms).  `value` = rays traced (radiance + shadow, counted on the device) per
second with the scene resident in HBM and the output left in HBM; `e2e` = the same through the
host-facing call (host scene constants in, host RGBA8 frame out, copies inside the timed region).

N > 1 (one process per GPU, torchrun):
  * the headline line stays config 2.  A 2 ms frame has a ~1 ms serial floor (the dependent chain of its slowest ray,
    per wave) and cannot be cut 8 ways, so the 1080p configs are dealt out as whole frames (frame g on rank g % N, a
    step is N frames): "scaling": "weak", at every N including 1.
  * the north-star split -- ONE image tile-sharded over the GPUs -- is measured in the same run on the configuration
    the north star names for it (config 5: 7680x4320, 1 M primitives) and reported in the same JSON line as
    `strong_scaling`: interleaved row stripes of every frame, replicated scene, every rank's rows written straight to
    their place in rank 0's device frame over NVLink (peer memory; the ranks' frame streams meet in a 4-byte NCCL
    all-reduce; --composite gather: NCCL gather of packed stripes + an un-interleave pass instead),
    with the one-GPU figure it is relative to measured in the same run (rank 0 alone, the other ranks idle).
`e2e` at N > 1: every rank copies its frames / rows straight into one host frame shared by all ranks (shared memory,
page-locked in each process), so the pixels leave over N PCIe links and nothing is gathered through rank 0's GPU.
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
SIZES = {1: (640, 480), 2: (1920, 1080), 3: (3840, 2160), 4: (1920, 1080), 5: (7680, 4320)}

# Algorithmic FP32 work per counted unit (FMA = 2, everything else = 1),
# per kernel type id: analytic box / spheres per test, metaball per march step (3 balls outside
# their radius + position update), SDF per distance evaluation incl. the 9-flop march overhead.
FLOPS_PER_UNIT = [30 + 42, 3 * 35 + 42, 3 * 11 + 6, 48 + 9, 34 + 9, 16 + 9, 20 + 9, 71 + 9, 66 + 9, 228 + 9]

L2_FLUSH_BYTES = 192 << 20  # > the 126 MB L2


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


# ---------------------------------------------------------------------------------------------------------------------
# What both arms must agree on without a GPU: how a workload is dealt over N ranks, and the `config` object of the line.
# ---------------------------------------------------------------------------------------------------------------------
def stripe_rows_for(height, world, max_rows=8):
    if world <= 1:
        return max(int(height), 1)
    for s in range(max_rows, 0, -1):
        if height % (s * world) == 0:
            return s
    return 1


def plan_for(config_id, world, args, force_mode=None):
    """mode: "single" (one GPU), "frames" (whole frames round-robin: frame g on rank g % N; a step is N frames) or
    "stripes" (every rank renders its row stripes of every frame).  A frame has a serial floor per wave (the dependent
    chain of its slowest ray), so a 2 ms frame cannot be split 8 ways while a 40 ms one can: frames up to 4 Mpixel,
    stripes above."""
    W, H = SIZES[config_id]
    if force_mode:
        mode = force_mode
    elif world == 1:
        mode = "single"
    elif args.multi_gpu != "auto":
        mode = args.multi_gpu
    else:
        mode = "frames" if W * H <= (1 << 22) else "stripes"
    # whole frames: 4 in flight (measured on one B200, config 2: 2 535 / 2 840 / 2 962 / 2 995 Mrays/s at 2 / 3 / 4 / 6); 3 for
    # the 8K frame, whose queues are 44 GB per frame context
    n_slots = args.frames_in_flight or (min(8, 2 + world) if mode == "stripes" else (4 if W * H <= (1 << 22) else 3))
    return {"mode": mode, "n_slots": n_slots, "frames_per_step": world if mode == "frames" else 1,
            "stripe_rows": stripe_rows_for(H, world) if mode == "stripes" else H, "width": W, "height": H}


STRIPE_COMPOSITE = {"peer": "each rank's RGBA8 rows copied straight to their global rows of rank 0's device frame over NVLink (CUDA IPC peer memory, "
                            "gprt_dispatch_stripes_host_async), one 4-byte NCCL all-reduce per frame as the barrier of the ranks' frame streams",
                    "gather": "NCCL gather of the packed RGBA8 stripe buffers to rank 0 + an un-interleave pass there"}


FRAME_COMPOSITE = {"peer": "every rank's final fold stores its RGBA8 frame straight into rank 0's device memory over NVLink (CUDA IPC peer memory), "
                           "one 4-byte NCCL all-reduce per step as the barrier of the ranks' frame streams",
                   "gather": "NCCL gather of the RGBA8 frames to rank 0"}


def config_object(config_id, world, args, plan):
    """The `config` of the JSON line: a function of the command line only, so that the reference arm prints the same object."""
    mode, srows = plan["mode"], plan["stripe_rows"]
    par = {"single": "1 GPU",
           "frames": f"{world} GPUs, whole frames round-robin (frame g on rank g % {world}), replicated scene, " + FRAME_COMPOSITE[args.composite] + "; "
                     f"a step is {world} frames, one per rank",
           "stripes": f"{world} GPUs, {srows}-row stripes interleaved, replicated scene, " + STRIPE_COMPOSITE[args.composite]}[mode]
    return {"workload": CONFIGS[config_id]["workload"], "width": plan["width"], "height": plan["height"], "math_mode": args.math,
            "frames_in_flight": plan["n_slots"], "multi_gpu": mode, "frames_per_step": plan["frames_per_step"],
            "l2_flush": f"{L2_FLUSH_BYTES >> 20} MiB memset on the frame's stream before every timed step, device-output and host-output passes alike "
                        "(and the wavefront state of a frame exceeds L2)",
            "timing": "CUDA events ordered around all frame streams, exactly K steps between barriers, max over ranks",
            "parallelism": par}


def scaling_label(plan):
    # frames / single: per-GPU work fixed as N grows (a step is N frames); stripes: one frame cut N ways
    return "strong" if plan["mode"] == "stripes" else "weak"


# ---------------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------------
class Job:
    """rank / world of the measurement at hand.  solo = rank 0 measures alone while the other ranks wait."""

    def __init__(self, rank, world, local_rank, dev):
        self.rank, self.world, self.local_rank, self.dev = rank, world, local_rank, dev

    def barrier(self):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def sum_over_ranks(self, v):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            t = torch.tensor([v], dtype=torch.int64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            v = int(t.item())
        return v


def measure(job, args, config_id, plan, steps, warmup, math, latency_steps=10, flush_l2=True):
    """Times `steps` steps of one workload on job.world ranks.  Returns the measurements (every rank; rank 0's are used)."""
    import torch

    from gpuraytracer_b200 import GPRT_MATH_FAST, GPRT_MATH_FMA, GPRT_MATH_PARITY, Renderer, default_config, multigpu

    rank, world, dev = job.rank, job.world, job.dev
    conf = CONFIGS[config_id]
    cfg = default_config()
    cfg.math_mode = {"parity": GPRT_MATH_PARITY, "fma": GPRT_MATH_FMA, "fast": GPRT_MATH_FAST}[math]
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0  # the scene's own
    r = Renderer(scene_id=conf["scene_id"], device=job.local_rank, config=cfg)
    r.on_init()
    W, H = r.width, r.height
    assert (W, H) == (plan["width"], plan["height"])
    mode, n_slots, frames_per_step, srows = plan["mode"], plan["n_slots"], plan["frames_per_step"], plan["stripe_rows"]
    r.set_frame_count(n_slots)  # the reference has frame_count = 3 (src/Renderer.h:92); a GPU that renders only a stripe needs more to hide a frame's serial tail
    r.create_frame_contexts()

    def frame_time(i):
        """elapsed_time of this rank's frame of step i"""
        return ((i * world + rank) if mode == "frames" else i) / 60.0
    n_prims = len(r.scene().aabbs)
    main = torch.cuda.current_stream()
    ext = [torch.cuda.ExternalStream(r.frame_stream(j), device=dev) for j in range(n_slots)]

    pad_rows = multigpu.max_local_rows(H, srows, world) if mode == "stripes" else H
    local = [torch.empty((pad_rows, W, 4), dtype=torch.uint8, device=dev) for _ in range(n_slots)]
    # Striped frames, device output: every rank copies its rows straight to their global rows of RANK 0's frame (peer memory
    # over NVLink, multigpu.PeerFrames) and the ranks' frame streams meet in a 4-byte all-reduce; --composite gather is the
    # NCCL gather of the packed stripe buffers + an un-interleave pass on rank 0 it replaced.  Two frames per slot: a rank
    # may start writing the slot's next frame while rank 0 still reads this one.
    peer = mode in ("stripes", "frames") and world > 1 and args.composite == "peer"
    n_peer = 2 * n_slots * (world if mode == "frames" else 1)  # whole frames: one frame per rank and step, side by side in rank 0's memory
    peer_frames = multigpu.PeerFrames(n_peer, H * W * 4, rank, world, job.local_rank, barrier=job.barrier,
                                      exchange=lambda blob: multigpu.broadcast_bytes(blob, 64 * n_peer, dev)) if peer else None
    composite = ("peer" if peer else "gather") if world > 1 else "none (one GPU)"
    if peer and job.max_over_ranks(1.0 if peer_frames.error else 0.0) > 0.0:
        # some rank could not allocate or map the frames (no CUDA IPC in this environment, no peer access): every rank takes
        # the NCCL gather instead, and the line says so
        if rank == 0:
            print(f"bench: peer-memory composite unavailable ({peer_frames.error or 'another rank failed'}); using the NCCL gather", file=sys.stderr)
        peer_frames.close()
        peer_frames, peer, composite = None, False, "gather (peer memory unavailable)"
    met = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(n_slots)] if peer else None
    slot_uses = [0] * n_slots
    # rank 0: one contiguous [world, rows, W, 4] block per slot for the gathered frames / stripe buffers
    blocks = [torch.empty((world, pad_rows, W, 4), dtype=torch.uint8, device=dev) if (world > 1 and rank == 0 and not peer) else None for _ in range(n_slots)]
    gathered = [list(b.unbind(0)) if b is not None else None for b in blocks]
    full = [torch.empty((H, W, 4), dtype=torch.uint8, device=dev) if (mode == "stripes" and rank == 0 and not peer) else local[j] for j in range(n_slots)]

    def frame_to_rank0(j):
        """This rank's whole frame of the step, written by the final fold itself into rank 0's memory (stores over NVLink)."""
        k = (j + n_slots * (slot_uses[j] & 1)) * world + rank
        slot_uses[j] += 1
        r.do_raytracing_async(peer_frames.ptr(k))
        multigpu.stream_barrier(met[j])
        return k

    def striped_to_rank0(j):
        """This rank's stripes of the current frame into rank 0's device frame (slot j); returns the frame's index in peer_frames."""
        k = j + n_slots * (slot_uses[j] & 1)
        slot_uses[j] += 1
        r.on_render_striped_async(peer_frames.ptr(k), srows, rank, world)
        multigpu.stream_barrier(met[j])
        return k
    # host side of the e2e pass: per slot one [frames_per_step, H, W, 4] block every rank can write (see SharedHostFrames)
    frame_bytes = H * W * 4
    # the same name on every rank of this measurement (ranks may have run different numbers of solo measurements before it)
    host = multigpu.SharedHostFrames(f"{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}_c{config_id}_{mode}{world}", n_slots, frames_per_step * frame_bytes,
                                     rank, world, barrier=job.barrier)
    flush = [torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev) for _ in range(n_slots)]

    def step(i, to_host, flush_l2):
        """One step: update (host), [L2 flush], dispatch; device output: gather + compose on rank 0; host output: every rank copies
        its frame / rows into the shared host block.  Asynchronous: frame i runs on frame slot i % n_slots (own context, own stream);
        the *_async call first waits for that slot's previous frame."""
        j = r.next_frame_slot()
        r.on_update(frame_time(i))
        with torch.cuda.stream(ext[j]):
            if flush_l2:
                flush[j].zero_()
            if mode == "single":
                if to_host:
                    r.on_render_async(host.ptr(j))  # host scene constants in, pinned host frame out
                else:
                    r.do_raytracing_async(local[j].data_ptr())
            elif mode == "frames":
                if to_host:
                    r.on_render_async(host.ptr(j, rank * frame_bytes))  # frame i * N + rank, over this rank's own PCIe link
                elif peer:
                    frame_to_rank0(j)
                else:
                    r.do_raytracing_async(local[j].data_ptr())
                    multigpu.gather_frames(local[j], gathered[j], dst=0)  # frame i * N + k from rank k
            else:
                if to_host:
                    r.on_render_striped_async(host.ptr(j), srows, rank, world)  # this rank's rows to their global rows of the shared frame
                elif peer:
                    striped_to_rank0(j)
                else:
                    r.do_raytracing_striped_async(local[j].data_ptr(), srows, rank, world)
                    multigpu.gather_stripes(local[j], gathered[j], dst=0)
                    if rank == 0:
                        multigpu.compose(full[j], gathered[j], H, srows, world)

    def timed(n, first_step, to_host, flush_l2):
        """EXACTLY n steps between two barriers; device time by CUDA events that every frame stream is ordered after /
        before (max over ranks); for the host-output path also the host wall clock, whichever is larger."""
        job.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(main)
        for s in ext:
            s.wait_event(e0)
        for i in range(n):
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
        job.barrier()
        return job.max_over_ranks(ms)

    def single_frame_latency(n, first_step):
        """Strictly serial: one blocking frame at a time (no frames in flight), L2 flushed before each, per-frame CUDA events."""
        total = 0.0
        for i in range(n):
            r.on_update(frame_time(first_step + i))
            if flush_l2:
                flush[0].zero_()
            job.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            j = r.next_frame_slot() if peer else 0  # the *_async call runs on the next frame slot's stream; the blocking calls on slot 0's
            with torch.cuda.stream(ext[j]):
                ext[j].wait_event(e0)
                if mode == "single":
                    r.do_raytracing(local[0].data_ptr())
                elif mode == "frames" and peer:
                    frame_to_rank0(j)
                elif mode == "frames":
                    r.do_raytracing(local[0].data_ptr())
                    multigpu.gather_frames(local[0], gathered[0], dst=0)
                elif peer:
                    striped_to_rank0(j)
                else:
                    r.do_raytracing_striped(local[0].data_ptr(), srows, rank, world)
                    multigpu.gather_stripes(local[0], gathered[0], dst=0)
                    if rank == 0:
                        multigpu.compose(full[0], gathered[0], H, srows, world)
            main.wait_stream(ext[j])
            e1.record(main)
            e1.synchronize()
            total += job.max_over_ranks(e0.elapsed_time(e1))
        return total / max(n, 1)

    def count_rays(n, first_step):
        """Rays traced (device counters) by exactly the frames of the timed passes (untimed, synchronous)."""
        rays = 0
        for i in range(n):
            r.on_update(frame_time(first_step + i))
            if mode != "stripes":
                r.do_raytracing(local[0].data_ptr())
            else:
                r.do_raytracing_striped(local[0].data_ptr(), srows, rank, world)
            c = r.counters()
            rays += int(c.radiance_rays) + int(c.shadow_rays)
        return job.sum_over_ranks(rays)

    # every frame slot allocates its wavefront queues on first use: prime all of them before the W warm-up steps
    for i in range(n_slots):
        step(i, True, flush_l2)
    for i in range(warmup):
        step(i, False, flush_l2)
    for i in range(warmup):
        step(i, True, flush_l2)
    r.wait_for_gpu()
    job.barrier()
    launches_per_step = r.launch_count()
    rays = count_rays(steps, warmup)
    dev_ms = timed(steps, warmup, False, flush_l2)
    e2e_ms = timed(steps, warmup, True, flush_l2)
    lat_n = min(steps, latency_steps)
    lat_ms = single_frame_latency(lat_n, warmup) if lat_n else None
    out = {
        "value": rays / (dev_ms * 1e-3) / 1e6, "ms_per_step": dev_ms / steps, "frame_ms": dev_ms / steps / frames_per_step,
        "primary_mrays_per_s": W * H * frames_per_step * steps / (dev_ms * 1e-3) / 1e6, "rays_per_step": rays // steps,
        "single_frame_ms": lat_ms, "e2e_value": rays / (e2e_ms * 1e-3) / 1e6, "e2e_ms_per_step": e2e_ms / steps,
        "h2d_bytes_per_step": (136 + (0 if r.uses_device_transforms() else 128 * n_prims)) * frames_per_step,
        "d2h_bytes_per_step": W * H * 4 * frames_per_step,
        "device_transforms": r.uses_device_transforms(), "launches_per_step": launches_per_step,
        "gpu_launches": launches_per_step * (2 * steps + lat_n), "n_prims": n_prims, "composite": composite,
    }
    # free every tensor that was used on the frame streams BEFORE the contexts (and their streams) are destroyed:
    # the caching allocator records an event on those streams when such a block is released
    job.barrier()
    host.close()
    if peer_frames is not None:
        peer_frames.close()
    del local, gathered, blocks, full, flush, host, met, peer_frames
    import gc
    gc.collect()
    torch.cuda.synchronize()
    job.barrier()
    del ext
    r.on_destroy()
    torch.cuda.empty_cache()
    return out


def strong_scaling(job_all, args, math):
    """The north-star split in the driver's record: config 5 (7680x4320, 1 M primitives) cut into interleaved row stripes over
    the N ranks of this run, and the one-GPU figure it is relative to measured in the same run (rank 0 alone)."""
    import torch.distributed as dist
    steps = max(3, min(args.steps, args.strong_steps))
    warmup = 3
    one = Job(0, 1, job_all.local_rank, job_all.dev)
    n1 = None
    if job_all.rank == 0:
        n1 = measure(one, args, 5, plan_for(5, 1, args), steps, warmup, math, latency_steps=3, flush_l2=False)
    if job_all.world == 1:
        mN, plan = n1, plan_for(5, 1, args)
    else:
        dist.barrier()
        plan = plan_for(5, job_all.world, args, force_mode="stripes")
        # N times the frames at N GPUs: every rank then renders as many rays inside the timed window as rank 0 did alone, and the
        # fill / drain of the frames in flight at the two ends of the window weighs the same in both figures
        mN = measure(job_all, args, 5, plan, steps * job_all.world, warmup, math, latency_steps=3, flush_l2=False)
    if job_all.rank != 0:
        return None
    W, H = SIZES[5]
    return {
        "workload": CONFIGS[5]["workload"], "width": W, "height": H, "n_gpus": job_all.world, "scaling": "strong", "steps": steps * job_all.world, "warmup": warmup,
        "sharding": "one GPU, whole frame" if job_all.world == 1 else f"{plan['stripe_rows']}-row stripes interleaved over {job_all.world} ranks "
                    "(gprt_dispatch_stripes), replicated scene and LBVH, " + STRIPE_COMPOSITE[args.composite],
        "frames_in_flight": plan["n_slots"], "ms_per_frame": round(mN["frame_ms"], 4), "value": round(mN["value"], 3), "unit": UNIT,
        "single_frame_ms": round(mN["single_frame_ms"], 4), "e2e_value": round(mN["e2e_value"], 3), "e2e_ms_per_frame": round(mN["e2e_ms_per_step"], 4),
        "rays_per_frame": mN["rays_per_step"], "composite": mN["composite"],
        "n1": {"ms_per_frame": round(n1["frame_ms"], 4), "value": round(n1["value"], 3), "e2e_value": round(n1["e2e_value"], 3),
               "single_frame_ms": round(n1["single_frame_ms"], 4), "frames_in_flight": plan_for(5, 1, args)["n_slots"],
               "steps": steps, "how": "same run, rank 0 alone renders whole frames while the other ranks wait; `steps` frames here, N * `steps` "
                                      "frames at N GPUs (the same rays per GPU inside either timed window)"},
        "efficiency": round(n1["frame_ms"] / (job_all.world * mN["frame_ms"]), 4),
        "e2e_efficiency": round(mN["e2e_value"] / (job_all.world * n1["e2e_value"]), 4),
        "l2": "no flush in this leg, at N GPUs and in the one-GPU figure alike: the inputs are larger than the 126 MB L2 (245 MB of scene and LBVH read by every "
              "frame, > 1 GB of ray / candidate queues written and read per frame and rank even at N = 8), so nothing survives from one frame to the "
              "next; a 192 MiB memset per step is a fixed cost per GPU that does not shrink with the rank's share and would be charged N times over",
        "efficiency_how": "T(1 GPU) / (N * T(N GPUs)) per frame, both with frames in flight, CUDA events, max over ranks",
    }


def run_gpu(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    job = Job(rank, world, local_rank, dev)
    plan = plan_for(args.config, world, args)
    conf = CONFIGS[args.config]

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    m = measure(job, args, args.config, plan, args.steps, args.warmup, args.math)
    clocks = sampler.stop() if rank == 0 else None

    strong = None
    if args.config == 2 and not args.no_strong:
        strong = strong_scaling(job, args, args.math)

    result = None
    if rank == 0:
        mode, n_slots = plan["mode"], plan["n_slots"]
        result = {
            "metric": METRIC, "value": round(m["value"], 3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(m["ms_per_step"], 4), "higher_is_better": True, "scaling": scaling_label(plan), "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_object(args.config, world, args, plan),
            "frame_ms": round(m["frame_ms"], 4),
            "primary_mrays_per_s": round(m["primary_mrays_per_s"], 3),
            "rays_per_step": m["rays_per_step"],
            "single_frame_ms": round(m["single_frame_ms"], 4),
            "single_frame_how": "one frame at a time (nothing in flight), L2 flushed, CUDA events per frame: the latency of one frame",
            "e2e": {"value": round(m["e2e_value"], 3), "unit": UNIT, "ms_per_step": round(m["e2e_ms_per_step"], 4),
                    "h2d_bytes_per_step": m["h2d_bytes_per_step"], "d2h_bytes_per_step": m["d2h_bytes_per_step"],
                    "transforms": "computed on the GPU (gprt_update_transforms)" if m["device_transforms"] else "computed on the host, uploaded every step",
                    "how": f"Renderer.on_update + on_render_async per step (public API, {n_slots} frames in flight): host scene constants in, page-locked host "
                           "RGBA8 frame out every step, L2 flushed before every step; max(CUDA events, host wall clock) around all steps + wait_for_gpu"
                           + {"single": "", "frames": "; every rank copies its frame into the host block shared by all ranks (its own PCIe link)",
                              "stripes": "; every rank copies its rows to their global rows of the host frame shared by all ranks (its own PCIe link)"}[mode]},
            "gpu_launches": m["gpu_launches"],
            "composite": m["composite"],  # how the device-resident frame(s) of a step reached rank 0 (what `config.parallelism` announces, or its fallback)
            "clocks": clocks,
        }
        if strong is not None:
            result["strong_scaling"] = strong
    if world > 1:
        dist.barrier()
    if rank == 0:
        result["roofline"], result["roofline_hbm"], result["kernel_ms"] = roofline(args, conf, local_rank, world, m)
        if world == 1 and not args.no_cpu:
            result["cpu_baseline"] = cpu_baseline(args.config, steps=2)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return result


def roofline(args, conf, device, world, m):
    """The dominant kernel of the step is the intersection kernel (all primitive types in one persistent launch per wave).  Its launches are timed live with CUDA events on the launching stream
    (GPRT_FLAG_PROFILE_KERNELS) over full single-GPU frames of the same workload, L2 flushed between frames; achieved =
    algorithmic FP32 work of those launches (device work counters x the hand-counted flops per unit of DESIGN.md section 5) /
    their duration, against the FP32-FMA issue-rate peak measured in this run.  `frac_of_timed_step` divides the same work by
    the timed region's ms per frame instead (frames in flight overlap the kernels of neighbouring frames).  A second object
    does the same for the most memory-bound kernel (the final fold) against the measured HBM copy bandwidth."""
    import torch
    from gpuraytracer_b200 import GPRT_MATH_FAST, GPRT_MATH_FMA, GPRT_MATH_PARITY, Renderer, default_config
    from gpuraytracer_b200._capi import GPRT_FLAG_PROFILE_KERNELS, GPRT_K_FINAL, GPRT_K_INTERSECT_ALL, KERNEL_NAMES
    cfg = default_config()
    cfg.math_mode = {"parity": GPRT_MATH_PARITY, "fma": GPRT_MATH_FMA, "fast": GPRT_MATH_FAST}[args.math]
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0
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
    ms = [x / reps for x in ms]
    work = [w / reps for w in work]
    launches = [n // reps for n in launches]
    kernel_ms = {KERNEL_NAMES[k]: round(ms[k], 4) for k in range(nk) if launches[k]}
    flops = sum(work[t] * FLOPS_PER_UNIT[t] for t in range(10))
    isect_ms = ms[GPRT_K_INTERSECT_ALL]
    dur = isect_ms * 1e-3
    n_l = max(launches[GPRT_K_INTERSECT_ALL], 1)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    achieved = flops / dur / 1e12 if dur > 0 else 0.0
    peak_src = "measured in this run: FFMA issue-rate probe (gprt_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure"
    if not fp32_peak or fp32_peak <= 0:
        fp32_peak, peak_src = 74.4, "nominal fallback: 148 SMs x 128 lanes x 2 x 1.965 GHz"
    # DRAM traffic per launch of the same kernel: NOT measured in this run (no profiler here) -- read from the committed ncu --set full
    # capture of the same workload under profiles/ (mean over the launches of one frame); null when none is committed
    traffic, traffic_src = _committed_traffic(conf, "k_intersect_coop", ("r02_ncu_full_frame_config2.json", "r01_ncu_full_k_intersect_coop.json"))
    plan = plan_for(args.config, world, args)
    # FP32 rate of the whole timed pipeline per GPU: the step's frames x flops per frame / the step's time / N GPUs
    pipeline_tflops = flops * plan["frames_per_step"] / (m["ms_per_step"] * 1e-3) / 1e12 / world if m["ms_per_step"] else None
    roof = {"bound": "fp32", "kernel": "k_intersect_coop (all primitive types, one persistent launch per wave)",
            "achieved": round(achieved, 3), "peak": round(fp32_peak, 2), "unit": "TFLOP/s",
            "frac": round(achieved / fp32_peak, 4),
            "frac_of_timed_step": round(pipeline_tflops / fp32_peak, 4) if pipeline_tflops else None,
            "frac_of_timed_step_how": "the same algorithmic flops per frame x frames per step / ms_per_step of the timed region / N GPUs (frames in flight: "
                                      "kernels of neighbouring frames overlap, so this is the whole pipeline's FP32 rate per GPU, not the kernel's alone)",
            "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
            "flops_per_launch": round(flops / n_l), "flops_per_step": round(flops), "launches_per_step": n_l, "avg_launch_ms": round(isect_ms / n_l, 4),
            "share_of_step": round(isect_ms / max(sum(ms), 1e-9), 3),
            "share_of_step_how": "of the summed kernel times of one frame rendered alone (no frames in flight), CUDA events per launch",
            "units_per_step": {KERNEL_NAMES[1 + t]: round(work[t]) for t in range(10) if work[t]},
            "flops_per_unit": {KERNEL_NAMES[1 + t]: FLOPS_PER_UNIT[t] for t in range(10) if work[t]},
            "note": "not a tensor-core path: the march kernels are FP32-issue bound (SURVEY.md 8d); the peak counts FMA as 2 flops"
                    + ("; the IEEE parity build issues separate FMUL/FADD, so its ceiling is half of this peak" if args.math == "parity" else "")}
    # memory-bound kernel: final fold = per pixel 4 B colour out + 4 B level code per level, plus 49 B of hit record per radiance hit
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    fin_bytes = W * H * (4 + depth * 4) + int(c.radiance_rays) * 49
    fin_dur = ms[GPRT_K_FINAL] * 1e-3
    fin_traffic, fin_traffic_src = _committed_traffic(conf, "k_final_color", ("r02_ncu_full_frame_config2.json", "r01_ncu_full_k_final_color.json"))
    hbm = {"bound": "hbm", "kernel": "k_final_color", "achieved": round(fin_bytes / fin_dur / 1e9, 1) if fin_dur > 0 else 0.0, "peak": hbm_peak,
           "unit": "GB/s", "frac": round(fin_bytes / fin_dur / 1e9 / hbm_peak, 4) if fin_dur > 0 else 0.0, "traffic": fin_traffic,
           "traffic_source": fin_traffic_src, "peak_source": hbm_src, "bytes_per_launch": fin_bytes}
    return roof, hbm, kernel_ms


def _committed_traffic(conf, kernel, names):
    if conf["scene_id"] != 2:
        return None, None
    for name in names:
        try:
            cap = json.load(open(os.path.join(ROOT, "profiles", name)))
            vals = [l["dram_bytes_total"] for l in cap["launches"] if "dram_bytes_total" in l and kernel in l.get("kernel", "")]
            if vals:
                return round(sum(vals) / len(vals)), f"from the committed capture profiles/{name} (dram__bytes_read.sum + dram__bytes_write.sum, mean per launch), not measured in this run"
        except Exception:
            continue
    return None, None


# ---------------------------------------------------------------------------------------------------------------------
# CPU legs: the reference's own shader code (oracle/_ref) on the host cores.  Scene constants come from libgprt_scene.so
# (CUDA-free); libgprt.so is never mapped here.
# ---------------------------------------------------------------------------------------------------------------------
def host_threads():
    """Every core this process may run on.  torchrun exports OMP_NUM_THREADS=1 when it is unset; the CPU legs ignore it
    and pass this count to the oracle explicitly."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_frames(config_id, steps, first_step=0):
    """Time `steps` frames of the workload on the host cores with the oracle; returns (seconds, rays, pixels, oracle, threads)."""
    from gpuraytracer_b200.scene_only import SceneOnly
    from oracle import oracle
    so = SceneOnly(CONFIGS[config_id]["scene_id"])
    mb = so.scene().config.metaballs_count
    orc = oracle.load(metaballs=mb, fast=True) if oracle.available("reference", mb, True) else oracle.load(metaballs=mb)
    n_thr = host_threads()
    total_s, rays, pixels, used = 0.0, 0, 0, 0
    for i in range(steps):
        so.on_update((first_step + i) / 60.0)
        scene = so.scene()
        orc.set_scene(scene)
        t0 = time.perf_counter()
        res = orc.render(scene, want_ids=False, threads=n_thr)
        total_s += time.perf_counter() - t0
        used = max(used, orc.last_threads())
        rays += res["counters"]["radiance_rays"] + res["counters"]["shadow_rays"]
        pixels += scene.width * scene.height
    if used < n_thr:
        raise RuntimeError(f"the oracle ran on {used} threads, {n_thr} cores are available: refusing to report a CPU baseline")
    return total_s, rays, pixels, orc, used


def cpu_baseline(config_id, steps=2):
    total_s, rays, pixels, orc, used = cpu_frames(config_id, steps)
    return {"value": round(rays / total_s / 1e6, 3), "unit": UNIT, "cores": used, "kind": orc.kind,
            "sample": f"{steps} full frames of the same workload ({pixels} pixels), OpenMP over rows on {used} threads, {os.path.basename(orc.path)}",
            "ms_per_step": round(total_s / steps * 1e3, 2)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    world = max(int(os.environ.get("WORLD_SIZE", "1")), args.gpus)
    plan = plan_for(args.config, world, args)
    for _ in range(min(args.warmup, 1)):
        cpu_frames(args.config, 1)
    total_s, rays, pixels, orc, used = cpu_frames(args.config, args.steps, first_step=args.warmup)
    value = rays / total_s / 1e6
    return {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_s / args.steps * 1e3, 3), "higher_is_better": True,
            "scaling": scaling_label(plan), "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_object(args.config, world, args, plan),
            "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": used, "kind": orc.kind,
                             "sample": f"{args.steps} full frames ({pixels} pixels), OpenMP over rows on {used} threads (OMP_NUM_THREADS ignored), "
                                       f"{os.path.basename(orc.path)}"},
            "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gprt", choices=["gprt", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS))
    ap.add_argument("--frames-in-flight", type=int, default=0, help="frame slots (contexts + streams); 0 = 4 for whole frames (measured on one B200: 2 535 / 2 840 / 2 962 / 2 995 Mrays/s at 2 / 3 / 4 / 6), min(8, 2 + N) for stripes on N")
    ap.add_argument("--multi-gpu", default="auto", choices=["auto", "frames", "stripes"],
                    help="N > 1: whole frames round-robin over the ranks, or row stripes of every frame; auto = frames up to 4 Mpixel, stripes above")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (development runs of the large configs)")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong_scaling leg (config 5 stripes) of the default config-2 run")
    ap.add_argument("--composite", choices=("peer", "gather"), default="peer",
                    help="striped frames, device output: rows written straight into rank 0's frame over NVLink (peer) or NCCL gather + un-interleave (gather)")
    ap.add_argument("--strong-steps", type=int, default=6, help="timed frames of the strong_scaling leg")
    ap.add_argument("--math", default=os.environ.get("GPRT_BENCH_MATH", "parity"), choices=["parity", "fma", "fast"])
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
