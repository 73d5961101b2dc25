#!/usr/bin/env python3
"""Build oracle/_ref/liboracle_ref*.so from the reference's own shader sources.

TEST INFRASTRUCTURE.  The reference tree (/root/reference, read-only) holds no CPU path; its
hot path is HLSL.  The HLSL compiles as C++17 after a few mechanical rewrites, which this
script applies to copies made in a scratch directory OUTSIDE the repository (never committed):

  1. `inout T name[` / `out T name[`      -> `T name[`      (arrays are by reference already)
  2. `inout T name`  / `out T name`       -> `T& name`
  3. `in T`                               -> `T`
  4. `1.#INF`                             -> `__builtin_inff()`
  5. `(ProceduralPrimitiveAttributes)0`   -> `ProceduralPrimitiveAttributes{}`
  6. Raytracing.hlsl only: drop the `register(...)` resource declarations and `[shader("...")]`
     attributes (oracle/ref_harness.cpp re-declares the resources)
  7. compile-time knobs become run-time variables of the harness:
       #define MAX_RAY_RECURSION_DEPTH 3 -> g_max_recursion_depth
       const UINT MaxSteps = 512;        -> g_sdf_max_steps
       UINT MAX_STEPS = 128;             -> g_metaball_max_steps
       #define FRACTAL_ITERATIONS_COUNT 4 -> g_fractal_iterations
     and METABALLS_COUNT is set to 3 or 5 per variant (the reference ships both code paths,
     src/ConstantBuffers.h:12-25, src/VolumetricPrimitives.hlsli:91-111)
  8. two counting hooks (one line each): SDF distance evaluations and metaball march steps.

Outputs only into oracle/_ref/ (git-ignored, travels to the GPU box):
  liboracle_ref.so          parity build  (-O2 -ffp-contract=off, counters, METABALLS_COUNT 3)
  liboracle_ref_mb5.so      parity build, METABALLS_COUNT 5
  liboracle_ref_fast.so     timing build  (-O3 -march=x86-64-v3, no counters), METABALLS_COUNT 3
  liboracle_ref_fast_mb5.so timing build, METABALLS_COUNT 5
"""
import os
import re
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("GPRT_REF_SRC", "/root/reference/src")
OUT_DIR = os.path.join(HERE, "_ref")

FILES = [
    "ConstantBuffers.h", "HlslCompat.h", "RaytracingShaderHelper.hlsli",
    "ProceduralPrimitivesLibrary.hlsli", "AnalyticPrimitives.hlsli", "VolumetricPrimitives.hlsli",
    "SignedDistancePrimitives.hlsli", "SignedDistanceFractals.hlsli", "Raytracing.hlsl",
]

TYPE = r"[A-Za-z_][\w:]*"
NAME = r"[A-Za-z_]\w*"


def rewrite_code(code: str) -> str:
    code = re.sub(rf"\b(?:inout|out)\s+({TYPE})\s+({NAME})\s*\[", r"\1 \2[", code)
    code = re.sub(rf"\b(?:inout|out)\s+({TYPE})\s+({NAME})", r"\1& \2", code)
    code = re.sub(rf"\bin\s+(?={TYPE}\s+{NAME})", "", code)
    code = code.replace("1.#INF", "__builtin_inff()")
    code = code.replace("(ProceduralPrimitiveAttributes)0", "ProceduralPrimitiveAttributes{}")
    return code


def rewrite_file(name: str, text: str, metaballs: int) -> str:
    out = []
    for line in text.splitlines():
        # keep line numbers identical to the reference file; rewrite only the code part
        code, sep, comment = line.partition("//")
        if name == "Raytracing.hlsl" and ("register(" in code or code.strip().startswith("[shader(")):
            out.append("// (resource declaration / shader attribute removed by build_ref.py)")
            continue
        code = rewrite_code(code)
        if name == "ConstantBuffers.h":
            code = re.sub(r"#define\s+MAX_RAY_RECURSION_DEPTH\s+\d+", "#define MAX_RAY_RECURSION_DEPTH g_max_recursion_depth", code)
            code = re.sub(r"#define\s+METABALLS_COUNT\s+\d+", f"#define METABALLS_COUNT {metaballs}", code)
            code = re.sub(r"#define\s+FRACTAL_ITERATIONS_COUNT\s+\d+", "#define FRACTAL_ITERATIONS_COUNT ((int)g_fractal_iterations)", code)
        if name == "SignedDistancePrimitives.hlsli":
            code = re.sub(r"const\s+UINT\s+MaxSteps\s*=\s*512\s*;", "const UINT MaxSteps = g_sdf_max_steps;", code)
        if name == "VolumetricPrimitives.hlsli":
            code = re.sub(r"UINT\s+MAX_STEPS\s*=\s*128\s*;", "UINT MAX_STEPS = g_metaball_max_steps;", code)
            if "float3 position = ray.origin + t * ray.direction;" in code:
                code += " ORACLE_COUNT_MB_STEP();"
        out.append(code + sep + comment)
    text = "\n".join(out) + "\n"
    if name == "ProceduralPrimitivesLibrary.hlsli":
        text, n = re.subn(r"(float\s+GetDistanceFromSignedDistancePrimitive\s*\([^)]*\)\s*\{)",
                          r"\1 ORACLE_COUNT_SDF(signedDistancePrimitive);", text, count=1)
        assert n == 1, "SDF counting hook not placed"
    return text


def check_rewrites(scratch: str):
    blob = "".join(open(os.path.join(scratch, f)).read() for f in FILES)
    for needle in ("g_max_recursion_depth", "g_sdf_max_steps", "g_metaball_max_steps", "g_fractal_iterations", "ORACLE_COUNT_MB_STEP", "ORACLE_COUNT_SDF"):
        assert needle in blob, f"rewrite for {needle} did not apply (reference changed?)"


def build_variant(tag: str, metaballs: int, flags, counters: int, verbose=False):
    scratch = tempfile.mkdtemp(prefix="gprt_oracle_ref_")
    try:
        for f in FILES:
            with open(os.path.join(REF_SRC, f), encoding="utf-8", errors="replace") as fh:
                text = fh.read()
            with open(os.path.join(scratch, f), "w") as fh:
                fh.write(rewrite_file(f, text, metaballs))
        check_rewrites(scratch)
        os.makedirs(OUT_DIR, exist_ok=True)
        out = os.path.join(OUT_DIR, f"liboracle_ref{tag}.so")
        cmd = ["g++", "-std=c++17", "-shared", "-fPIC", "-fopenmp", "-fsingle-precision-constant", "-fpermissive", "-w",
               f"-DORACLE_COUNTERS={counters}", "-I", scratch, "-I", HERE, *flags,
               os.path.join(HERE, "ref_harness.cpp"), "-o", out]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
        return out
    finally:
        shutil.rmtree(scratch, ignore_errors=True)


PARITY = ["-O2", "-ffp-contract=off", "-fno-fast-math"]
FAST = ["-O3", "-march=x86-64-v3", "-fno-fast-math"]


def main(argv):
    if not os.path.isdir(REF_SRC):
        print(f"build_ref: {REF_SRC} not present; keeping prebuilt oracle/_ref", file=sys.stderr)
        return 0
    verbose = "-v" in argv
    only = [a for a in argv if not a.startswith("-")]
    variants = [("", 3, PARITY, 1), ("_mb5", 5, PARITY, 1), ("_fast", 3, FAST, 0), ("_fast_mb5", 5, FAST, 0)]
    for tag, mb, flags, cnt in variants:
        if only and tag not in only and (tag or "base") not in only:
            continue
        print("built", build_variant(tag, mb, flags, cnt, verbose))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
