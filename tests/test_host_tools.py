"""CPU tests of the host-side pieces around the path: the CUDA-free scene library, the oracle's run-time knobs and the
pieces of bench.py both arms must agree on."""
import argparse
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_scene_library_matches_the_renderer_scene_builder():
    """libgprt_scene.so (plain g++, no CUDA) and the scene half of libgprt.so's Renderer hand out the same bytes."""
    from gpuraytracer_b200 import Renderer
    from gpuraytracer_b200.scene_only import SceneOnly
    for sid, grid in ((0, 0), (2, 0), (3, 0), (4, 0), (5, 24), (7, 0)):
        r = Renderer(scene_id=sid, grid_n=grid, device=None)
        s = SceneOnly(sid, grid_n=grid)
        for t in (0.0, 1.3):
            r.on_update(t)
            s.on_update(t)
            a, b = r.scene(), s.scene()
            for k in ("scene_cb", "per_frame", "aabbs", "hit_groups", "vertices", "indices", "instances"):
                assert getattr(a, k).tobytes() == getattr(b, k).tobytes(), (sid, t, k)
            assert (a.width, a.height) == (b.width, b.height)
            assert bytes(a.config) == bytes(b.config)


def test_scene_library_links_no_cuda():
    out = subprocess.run(["ldd", os.path.join(ROOT, "gpuraytracer_b200", "libgprt_scene.so")], capture_output=True, text=True).stdout
    assert "cuda" not in out.lower() and "libgprt.so" not in out


def test_oracle_fractal_iterations_knob(oracle3):
    """FRACTAL_ITERATIONS_COUNT (src/ConstantBuffers.h:27) as a run-time knob of the oracle: n folds, then sdPyramid * 2^-n
    (src/SignedDistanceFractals.hlsli:34-63), restated here in float32 numpy."""
    f = np.float32

    def pyramid(p, n):
        h = (f(0.894), f(0.447), f(2.0))
        p = np.array([p[0], p[1] + f(1.0), p[2]], dtype=np.float32)
        a = f(f(h[2] * h[1]) / h[0])
        verts = [np.array(v, dtype=np.float32) for v in ((0, h[2], 0), (-a, 0, a), (a, 0, -a), (a, 0, a), (-a, 0, -a))]
        for _ in range(n):
            best, bd = None, None
            for v in verts:
                q = p - v
                d = f(f(f(q[0] * q[0]) + f(q[1] * q[1])) + f(q[2] * q[2]))
                if bd is None or d < bd:
                    best, bd = v, d
            p = (f(2.0) * p - best * f(1.0)).astype(np.float32)
        d = f(f(max(abs(p[0]), abs(p[2])) * h[0]) + f(abs(p[1]) * h[1]))
        octa = f(d - f(h[1] * h[2]))
        return f(max(octa, -p[1])) * f(2.0 ** -n)

    pts = [(0.3, 0.2, 0.1), (-0.8, -0.5, 0.4), (0.05, 0.9, -0.02)]
    try:
        for n in (1, 2, 4, 6):
            oracle3.set_fractal_iterations(n)
            for p in pts:
                got = oracle3.sdf_distance(6, p)
                want = float(pyramid(np.array(p, dtype=np.float32), n))
                assert got == pytest.approx(want, rel=1e-5, abs=1e-7), (n, p)
    finally:
        oracle3.set_fractal_iterations(4)


def test_oracle_render_respects_fractal_iterations(oracle3):
    from gpuraytracer_b200.scene_only import SceneOnly
    s = SceneOnly(3, 96, 54).scene()
    imgs = {}
    for n in (2, 4):
        oracle3.set_scene(s, fractal_iterations=n)
        imgs[n] = oracle3.render(s)["rgba8"]
    oracle3.set_scene(s)
    assert (imgs[2] != imgs[4]).any()
    np.testing.assert_array_equal(oracle3.render(s)["rgba8"], imgs[4])  # the default is the reference's 4


def _args(**kw):
    d = dict(gpus=1, steps=2, warmup=3, config=2, frames_in_flight=0, multi_gpu="auto", math="parity", composite="peer")
    d.update(kw)
    return argparse.Namespace(**d)


def test_bench_plan_and_config_object():
    sys.path.insert(0, ROOT)
    import bench
    a = _args()
    p1 = bench.plan_for(2, 1, a)
    assert (p1["mode"], p1["frames_per_step"], p1["n_slots"]) == ("single", 1, 4)
    p8 = bench.plan_for(2, 8, a)
    assert (p8["mode"], p8["frames_per_step"]) == ("frames", 8)
    assert bench.scaling_label(p1) == bench.scaling_label(p8) == "weak"   # one curve, one label
    s8 = bench.plan_for(5, 8, a)
    assert s8["mode"] == "stripes" and 4320 % (s8["stripe_rows"] * 8) == 0 and bench.scaling_label(s8) == "strong"
    # both arms print the same config object: it is a function of the command line only
    assert bench.config_object(2, 8, a, p8) == bench.config_object(2, 8, _args(gpus=8), bench.plan_for(2, 8, _args(gpus=8)))
    assert bench.config_object(2, 1, a, p1)["width"] == 1920


def test_reference_arm_ignores_omp_num_threads_and_never_maps_the_product_library():
    """torchrun exports OMP_NUM_THREADS=1: the reference arm must still use every core, print the same config object as the
    GPU arm would, and time the reference's code without libgprt.so in its address space."""
    import json
    env = dict(os.environ, OMP_NUM_THREADS="1")
    code = ("import sys, json; sys.argv=['bench.py','--impl','reference','--config','1','--steps','1','--warmup','0','--gpus','4'];"
            "import bench; bench.main(); maps=open('/proc/self/maps').read();"
            "print(json.dumps({'gprt': 'libgprt.so' in maps, 'scene': 'libgprt_scene.so' in maps, 'oracle': 'liboracle_ref' in maps}))")
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    line, maps = json.loads(lines[0]), json.loads(lines[1])
    assert line["impl"] == "reference" and line["n_gpus"] == 4
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert line["config"]["multi_gpu"] == "frames" and line["config"]["width"] == 640
    assert maps == {"gprt": False, "scene": True, "oracle": True}


_INVERSE_PIN = r"""
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include "gpuraytracer_b200/host/mathx.h"
using gprt_host::Mat4;
// the textbook loop mat_inverse restates with compile-time indices: Gauss-Jordan, partial pivoting by a run-time row index
static Mat4 textbook(const Mat4& a) {
    double w[4][8];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { w[i][j] = a.m[i][j]; w[i][j + 4] = (i == j) ? 1.0 : 0.0; }
    for (int c = 0; c < 4; c++) {
        int p = c;
        for (int r = c + 1; r < 4; r++) if (std::fabs(w[r][c]) > std::fabs(w[p][c])) p = r;
        if (p != c) for (int j = 0; j < 8; j++) { double t = w[c][j]; w[c][j] = w[p][j]; w[p][j] = t; }
        const double d = w[c][c];
        for (int j = 0; j < 8; j++) w[c][j] /= d;
        for (int r = 0; r < 4; r++) if (r != c) { const double f = w[r][c]; if (f != 0.0) for (int j = 0; j < 8; j++) w[r][j] -= f * w[c][j]; }
    }
    Mat4 r;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) r.m[i][j] = w[i][j + 4];
    return r;
}
int main() {
    std::mt19937_64 rng(0);
    std::uniform_real_distribution<double> u(-1.0, 1.0);
    int bad = 0, n = 0;
    for (int k = 0; k < 20000; k++) {
        Mat4 m;
        if (k % 2) { // the shape the renderer inverts: scale * rotationY * translation, translations up to +-2000
            const double s = 0.2 + 1.5 * std::fabs(u(rng));
            m = gprt_host::mat_mul(gprt_host::mat_mul(gprt_host::mat_scaling(s, s * (1.0 + 0.3 * u(rng)), s), gprt_host::mat_rotation_y(7.0 * u(rng))),
                                   gprt_host::mat_translation(2000.0 * u(rng), 3.0 * u(rng), 2000.0 * u(rng)));
        } else {
            for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) m.m[i][j] = (k % 7 == 0 && i != j && (i + j) % 3 == 0) ? 0.0 : 10.0 * u(rng);
        }
        const Mat4 a = gprt_host::mat_inverse(m), b = textbook(m);
        n++;
        if (std::memcmp(a.m, b.m, sizeof a.m) != 0) bad++;
    }
    std::printf("%d %d\n", n, bad);
    return bad != 0;
}
"""


def test_matrix_inverse_with_static_indices_is_the_textbook_loop_bit_for_bit(tmp_path):
    """host/mathx.h mat_inverse (XMMatrixInverse restated; also what the device-side transform update compiles) keeps every
    index a compile-time constant; it must give the bits of the plain Gauss-Jordan loop with a run-time pivot row."""
    src = tmp_path / "inv.cpp"
    src.write_text(_INVERSE_PIN)
    exe = tmp_path / "inv"
    out = subprocess.run(["g++", "-O2", "-ffp-contract=off", "-I", ROOT, "-o", str(exe), str(src)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    assert out.stdout.split() == ["20000", "0"]
