"""gprt_lbvh_build_noalloc (include/gprt.h): the LBVH build with its scratch in a fixed device array of the library, for
callers that may not allocate.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from gpuraytracer_b200 import Renderer, _capi, default_config

pytestmark = pytest.mark.gpu


def _context_tree(grid_n):
    """Scene 5 with grid_n x grid_n primitives built by gprt_build_accel: its AABBs, tag bytes, nodes, sorted leaves, root."""
    lib = _capi.lib()
    cfg = default_config()
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = cfg.fractal_iterations = 0
    with Renderer(scene_id=5, width=64, height=64, grid_n=grid_n, device=0, config=cfg) as r:
        r.on_init()
        aabbs = np.ascontiguousarray(r.scene().aabbs).view(np.float32).reshape(-1, 6).copy()
        n = aabbs.shape[0]
        nodes = np.zeros((max(n - 1, 1), 16), dtype=np.uint32)
        order = np.zeros(n, dtype=np.uint32)
        tags = np.zeros(n, dtype=np.uint8)
        root = C.c_int32(0)
        _capi.check(lib.gprt_get_accel(r.context, n, nodes.ctypes.data, order.ctypes.data, tags.ctypes.data, C.byref(root)), "get_accel")
    return aabbs, tags, nodes[:n - 1], order, root.value


def test_lbvh_build_without_allocation_equals_build_accel():
    """Two calls in a row through the C ABI (a 4 096-primitive grid, then a 900-primitive one over the same scratch, then the
    first again): nodes, leaf order and root are those of gprt_build_accel bit for bit (integer / index work and min / max
    of the same floats: no tolerance)."""
    import torch
    lib = _capi.lib()
    trees = [_context_tree(64), _context_tree(30)]
    dev = torch.device("cuda", 0)
    # every buffer of every call exists before the first call: the entry point itself allocates nothing
    bufs = []
    for aabbs, tags, nodes, order, root in trees:
        n = aabbs.shape[0]
        bufs.append((torch.from_numpy(aabbs).to(dev), torch.from_numpy(tags).to(dev), torch.full((n - 1, 16), 0xCD, dtype=torch.int32, device=dev),
                     torch.full((n,), -1, dtype=torch.int32, device=dev)))
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    for call, k in enumerate((0, 1, 0)):
        aabbs, tags, nodes, order, root = trees[k]
        d_aabbs, d_tags, d_nodes, d_order = bufs[k]
        d_nodes.fill_(0x5A5A5A5A)
        d_order.fill_(-1)
        n = aabbs.shape[0]
        before = torch.cuda.memory_allocated(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        got_root = C.c_int32(123)
        rc = lib.gprt_lbvh_build_noalloc(0, d_aabbs.data_ptr(), n, d_tags.data_ptr(), d_nodes.data_ptr(), d_order.data_ptr(), C.byref(got_root), stream)
        assert rc == _capi.GPRT_OK, lib.gprt_error_string(rc)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        if call > 0:  # (the first call on a device may load the module that holds the scratch array)
            assert torch.cuda.mem_get_info(dev)[0] >= free_before  # nothing new was taken from the device
        assert got_root.value == root
        np.testing.assert_array_equal(d_order.cpu().numpy().view(np.uint32), order)
        np.testing.assert_array_equal(d_nodes.cpu().numpy().view(np.uint32), nodes)


def test_lbvh_build_without_allocation_rejects_sizes_beyond_its_scratch():
    lib = _capi.lib()
    root = C.c_int32(0)
    rc = lib.gprt_lbvh_build_noalloc(0, None, _capi.GPRT_NOALLOC_MAX_PRIMS + 1, None, None, None, C.byref(root), None)
    assert rc == _capi.GPRT_E_TOO_LARGE
    assert b"GPRT_NOALLOC_MAX_PRIMS" in lib.gprt_error_string(rc)
    assert lib.gprt_lbvh_build_noalloc(0, None, 0, None, None, None, C.byref(root), None) == _capi.GPRT_OK and root.value == -2 ** 31
    assert lib.gprt_lbvh_build_noalloc(0, None, 16, None, None, None, C.byref(root), None) == _capi.GPRT_E_INVALID_ARG
