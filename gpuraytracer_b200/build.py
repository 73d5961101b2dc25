"""Build libgprt.so (the product) in-tree with nvcc for sm_100a.

    python -m gpuraytracer_b200.build [-v]

kernels.cu is compiled twice: the parity build (--fmad=false, IEEE div/sqrt) and the fast build
(--use_fast_math).  Everything else once.  No torch, no JIT cache: the .so travels with the tree.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libgprt.so")
CLI = os.path.join(HERE, "gprt_render")
SCENE_LIB = os.path.join(HERE, "libgprt_scene.so")  # the CUDA-free scene builder on its own (g++ only)

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default", "-Xcompiler", "-fno-fast-math",
          "-Xcompiler", "-ffp-contract=off"]
PARITY = ["--fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false", "-DGPRT_KERNELS_PARITY"]
FAST = ["--use_fast_math", "-DGPRT_KERNELS_FAST", "-DGPRT_FAST_MATH"]
FMA = ["--fmad=true", "-prec-div=true", "-prec-sqrt=true", "-ftz=false", "-DGPRT_KERNELS_FMA"]

UNITS = [
    (os.path.join(CSRC, "kernels.cu"), "kernels_parity.o", PARITY),
    (os.path.join(CSRC, "kernels.cu"), "kernels_fast.o", FAST),
    (os.path.join(CSRC, "kernels.cu"), "kernels_fma.o", FMA),
    (os.path.join(CSRC, "lbvh.cu"), "lbvh.o", []),
    (os.path.join(CSRC, "fp32_probe.cu"), "fp32_probe.o", []),
    (os.path.join(CSRC, "transforms.cu"), "transforms.o", ["--fmad=false"]),
    (os.path.join(CSRC, "noalloc.cu"), "noalloc.o", []),
    (os.path.join(CSRC, "gprt_api.cu"), "gprt_api.o", []),
    (os.path.join(CSRC, "gprt_common.cpp"), "gprt_common.o", []),
    (os.path.join(HOST, "scene.cpp"), "scene.o", []),
    (os.path.join(HOST, "renderer.cpp"), "renderer.o", []),
    (os.path.join(HOST, "scene_capi.cpp"), "scene_capi.o", []),
    (os.path.join(HOST, "multi_renderer.cpp"), "multi_renderer.o", []),
]

HEADERS = [os.path.join(CSRC, f) for f in ("vecmath.cuh", "prims.cuh", "shade.cuh", "wavefront.cuh", "march.cuh", "march_coop.cuh", "march_lane.cuh", "walk.cuh", "megakernel.cuh", "lbvh.h")] + \
          [os.path.join(HOST, f) for f in ("scene.h", "renderer.h", "multi_renderer.h", "mathx.h")] + \
          [os.path.join(HERE, "..", "include", f) for f in ("gprt.h", "gprt_host.h")]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(verbose=False, force=False, ptxas_info=False):
    os.makedirs(OBJ, exist_ok=True)
    objs = []
    procs = []
    for src, obj, flags in UNITS:
        out = os.path.join(OBJ, obj)
        objs.append(out)
        if not force and not _newer(out, [src, __file__] + HEADERS):
            continue
        cmd = ["nvcc", *ARCH, *COMMON, *flags, "-c", src, "-o", out]
        if ptxas_info and src.endswith(".cu"):
            cmd.insert(1, "-Xptxas=-v")
        if verbose:
            print(" ".join(cmd), flush=True)
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose or ptxas_info:
            sys.stdout.write(out)
        if p.returncode != 0:
            failed = True
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or _newer(LIB, objs):
        cmd = ["nvcc", *ARCH, "-shared", "-o", LIB, *objs, "-cudart", "static"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    # the scene builder alone: plain g++, no CUDA runtime, nothing of the render path
    scene_srcs = [os.path.join(HOST, "scene.cpp"), os.path.join(HOST, "scene_capi.cpp"), os.path.join(CSRC, "gprt_common.cpp")]
    if force or _newer(SCENE_LIB, scene_srcs + HEADERS):
        cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-fno-fast-math", "-ffp-contract=off", *scene_srcs, "-o", SCENE_LIB]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    # the headless CLI (host/gprt_render.cpp): plain C++ against the C ABI, finds libgprt.so next to itself
    cli_src = os.path.join(HOST, "gprt_render.cpp")
    if force or _newer(CLI, [cli_src, LIB] + HEADERS):
        cmd = ["g++", "-O2", "-std=c++17", cli_src, "-o", CLI, "-L" + HERE, "-lgprt", "-Wl,-rpath,$ORIGIN"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    build(verbose="-v" in sys.argv, force="-f" in sys.argv, ptxas_info="--ptxas" in sys.argv)
    print(LIB)
