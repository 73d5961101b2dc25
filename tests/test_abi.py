"""CPU tests of the drop-in boundary: struct layouts, exported symbols, error behaviour without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from gpuraytracer_b200 import GprtError, Renderer, _capi, default_config
from tests.conftest import ROOT, _has_gpu


def test_struct_sizes_match_reference_layout():
    # src/ConstantBuffers.h:49-91, src/Renderer.cpp:499-503,732-786,938-979 (see include/gprt.h)
    assert _capi.SCENE_CB_DTYPE.itemsize == 144
    assert _capi.SCENE_CB_DTYPE.fields["camera_position"][1] == 64
    assert _capi.SCENE_CB_DTYPE.fields["light_position"][1] == 80
    assert _capi.SCENE_CB_DTYPE.fields["light_ambient_color"][1] == 96
    assert _capi.SCENE_CB_DTYPE.fields["light_diffuse_color"][1] == 112
    assert _capi.SCENE_CB_DTYPE.fields["elapsed_time"][1] == 132
    assert _capi.MATERIAL_DTYPE.itemsize == 48
    assert _capi.MATERIAL_DTYPE.fields["step_scale"][1] == 32
    assert _capi.PER_FRAME_DTYPE.itemsize == 128
    assert _capi.AABB_DTYPE.itemsize == 24
    assert _capi.VERTEX_DTYPE.itemsize == 24
    assert _capi.INSTANCE_DTYPE.itemsize == 64
    assert _capi.HIT_GROUP_DTYPE.itemsize == 96
    assert _capi.HIT_GROUP_DTYPE.fields["material_cb"][1] == 32
    assert _capi.HIT_GROUP_DTYPE.fields["instance_index"][1] == 80
    assert C.sizeof(_capi.Config) == 36
    assert C.sizeof(_capi.Rect) == 16


def _declared_symbols():
    names = set()
    for h in ("gprt.h", "gprt_host.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for m in re.finditer(r"\b(gprt_[a-z0-9_]+)\s*\(", text):
            names.add(m.group(1))
    return names


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(_capi.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 25
    for name in sorted(declared):
        assert hasattr(lib, name), f"libgprt.so does not export {name} declared in include/*.h"
    for name in _capi.EXPORTED_SYMBOLS:
        assert name in declared, f"{name} bound in _capi.py but not declared in include/*.h"


def test_default_config_is_the_reference_compile_time_values():
    cfg = default_config()
    assert (cfg.abi_version, cfg.max_recursion_depth, cfg.metaballs_count) == (2, 3, 3)
    assert (cfg.metaball_max_steps, cfg.sdf_max_steps) == (128, 512)
    assert _capi.lib().gprt_abi_version() == 2


def test_error_strings():
    lib = _capi.lib()
    assert lib.gprt_error_string(0) == b"ok"
    for code in range(-8, 0):
        assert len(lib.gprt_error_string(code)) > 3
    assert b"no CPU fallback" in lib.gprt_error_string(_capi.GPRT_E_NO_DEVICE)


def test_null_and_bad_arguments():
    lib = _capi.lib()
    assert lib.gprt_create(None, 0, None) == _capi.GPRT_E_INVALID_ARG
    assert lib.gprt_sync(None) == _capi.GPRT_E_INVALID_ARG
    assert lib.gprt_build_accel(None) == _capi.GPRT_E_INVALID_ARG
    h = C.c_void_p()
    bad = default_config()
    bad.abi_version = 99
    assert lib.gprt_create(C.byref(h), 0, C.byref(bad)) == _capi.GPRT_E_ABI_MISMATCH
    bad = default_config()
    bad.metaballs_count = 4
    assert lib.gprt_create(C.byref(h), 0, C.byref(bad)) == _capi.GPRT_E_UNSUPPORTED
    bad = default_config()
    bad.max_recursion_depth = 0
    assert lib.gprt_create(C.byref(h), 0, C.byref(bad)) == _capi.GPRT_E_UNSUPPORTED
    with pytest.raises(GprtError):
        Renderer(scene_id=99, device=None)


@pytest.mark.skipif(_has_gpu(), reason="checks the no-device behaviour")
def test_no_device_fails_loudly_no_cpu_fallback():
    h = C.c_void_p()
    cfg = default_config()
    assert _capi.lib().gprt_create(C.byref(h), 0, C.byref(cfg)) == _capi.GPRT_E_NO_DEVICE
    r = Renderer(scene_id=0, width=64, height=36, device=0)
    with pytest.raises(GprtError) as e:
        r.on_init()
    assert e.value.status == _capi.GPRT_E_NO_DEVICE
    # a scene-only renderer cannot render either
    r = Renderer(scene_id=0, width=64, height=36, device=None)
    with pytest.raises(GprtError) as e:
        r.on_render()
    assert e.value.status == _capi.GPRT_E_NO_DEVICE


def test_product_does_not_touch_the_oracle():
    """The product path may not import / link / open anything under oracle/."""
    pkg = os.path.join(ROOT, "gpuraytracer_b200")
    for base, _, files in os.walk(pkg):
        if os.path.basename(base) == "build":
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(base, f), errors="replace").read()
                for needle in ("liboracle", "oracle/", "from oracle", "import oracle", "hlsl_shim", "ref_harness", "oracle_render"):
                    assert needle not in text, f"{f} references {needle}"
    out = os.popen(f"ldd {_capi.LIB_PATH}").read()
    assert "oracle" not in out


def test_write_ppm(tmp_path):
    from gpuraytracer_b200 import write_ppm
    img = np.zeros((4, 6, 4), np.uint8)
    img[..., 0] = 255
    p = tmp_path / "x.ppm"
    write_ppm(p, img)
    data = open(p, "rb").read()
    assert data.startswith(b"P6\n6 4\n255\n") and len(data) == 11 + 6 * 4 * 3


def test_headless_cli_exists_and_fails_loudly_without_a_gpu():
    """gprt_render (host/gprt_render.cpp) is the stand-in for the reference's main.cpp + Window loop.  It links the
    C ABI only; without a GPU it must fail with the library's "no device" error, not fall back to anything."""
    import os
    import subprocess
    import torch
    cli = os.path.join(os.path.dirname(_capi.__file__), "gprt_render")
    assert os.path.exists(cli), "run __graft_entry__.build()"
    out = subprocess.run([cli, "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "--scene" in out.stdout
    assert subprocess.run([cli, "--bogus"], capture_output=True).returncode == 2
    if not torch.cuda.is_available():
        out = subprocess.run([cli, "--scene", "0", "--width", "64", "--height", "36"], capture_output=True, text=True)
        assert out.returncode == 1 and "no usable CUDA device" in out.stderr
