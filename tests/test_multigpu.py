"""Multi-GPU sharding: host logic on CPU (gloo, world_size 2 and 3) and the striped dispatch on the GPU."""
import os
import socket
import sys

import numpy as np
import pytest

from gpuraytracer_b200 import multigpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


@pytest.mark.parametrize("height,srows,world", [(1080, 5, 8), (1080, 8, 8), (17, 3, 2), (5, 2, 4), (4320, 6, 8), (480, 8, 1), (7, 16, 3)])
def test_stripes_partition_the_frame(height, srows, world):
    rows = [multigpu.global_rows(height, srows, r, world) for r in range(world)]
    assert [len(x) for x in rows] == [multigpu.local_rows(height, srows, r, world) for r in range(world)]
    allr = np.sort(np.concatenate(rows))
    np.testing.assert_array_equal(allr, np.arange(height))  # every row exactly once
    for r, x in enumerate(rows):
        assert ((x // srows) % world == r).all()


def test_stripe_rows_for_gives_equal_shares():
    for h, w in [(1080, 2), (1080, 4), (1080, 8), (4320, 8), (2160, 8), (480, 3), (17, 2)]:
        s = multigpu.stripe_rows_for(h, w)
        assert 1 <= s <= 8
        if h % w == 0:
            assert len({multigpu.local_rows(h, s, r, w) for r in range(w)}) == 1
    assert multigpu.stripe_rows_for(1080, 1) == 1080


def test_compose_irregular_numpy():
    h, w, srows, world = 23, 5, 3, 4
    full = np.random.default_rng(0).integers(0, 255, (h, w, 4), dtype=np.uint8)
    pad = multigpu.max_local_rows(h, srows, world)
    gathered = []
    for r in range(world):
        buf = np.zeros((pad, w, 4), np.uint8)
        rows = multigpu.global_rows(h, srows, r, world)
        buf[:len(rows)] = full[rows]
        gathered.append(buf)
    out = multigpu.compose(np.zeros_like(full), gathered, h, srows, world)
    np.testing.assert_array_equal(out, full)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, height_crop, srows, out_path):
    """One rank of the sharded frame on CPU: its stripes come from a committed golden image (the
    oracle's output for config 1), the exchange + composition are the package's own."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from gpuraytracer_b200 import multigpu as mg
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    golden = np.load(os.path.join(GOLDEN, "config1.npz"))["rgba8"][:height_crop]
    h, w = golden.shape[:2]
    pad = mg.max_local_rows(h, srows, world)
    local = torch.zeros((pad, w, 4), dtype=torch.uint8)
    rows = mg.global_rows(h, srows, rank, world)
    local[:len(rows)] = torch.from_numpy(golden[rows])  # stands in for gprt_dispatch_stripes
    gathered = [torch.zeros_like(local) for _ in range(world)] if rank == 0 else None
    mg.gather_stripes(local, gathered, dst=0)
    if rank == 0:
        full = mg.compose(torch.zeros((h, w, 4), dtype=torch.uint8), gathered, h, srows, world)
        np.save(out_path, full.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,height_crop,srows", [(2, 480, 8), (2, 477, 5), (3, 100, 4)])
def test_gather_and_compose_over_gloo(tmp_path, world, height_crop, srows):
    import torch.multiprocessing as mp
    out = str(tmp_path / "full.npy")
    mp.spawn(_worker, args=(world, _free_port(), height_crop, srows, out), nprocs=world, join=True)
    golden = np.load(os.path.join(GOLDEN, "config1.npz"))["rgba8"][:height_crop]
    np.testing.assert_array_equal(np.load(out), golden)


def _frames_worker(rank, world, port, steps, out_path):
    import torch
    import torch.distributed as dist
    from gpuraytracer_b200 import multigpu as mg
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    order = []
    for step in range(steps):
        g = step * world + rank  # the global frame this rank renders in this step
        assert mg.frame_owner(g, world) == rank
        local = torch.full((4, 6, 4), g, dtype=torch.uint8)  # stands in for the rendered frame
        gathered = [torch.zeros_like(local) for _ in range(world)] if rank == 0 else None
        mg.gather_frames(local, gathered, dst=0)
        if rank == 0:
            order += [int(f[0, 0, 0]) for f in gathered]
            assert all(bool((f == f[0, 0, 0]).all()) for f in gathered)
    if rank == 0:
        np.save(out_path, np.array(order))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_frame_sharding_delivers_frames_in_order_over_gloo(tmp_path, world):
    """Frame sharding (bench.py, 1080p configs at N > 1): frame g is rendered by rank g % N and rank 0 receives
    the frames of every step in global order."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "order.npy")
    mp.spawn(_frames_worker, args=(world, _free_port(), 4, out), nprocs=world, join=True)
    np.testing.assert_array_equal(np.load(out), np.arange(4 * world))


def _direct_worker(rank, world, port, height_crop, srows, out_path):
    """The default composite of bench.py on CPU: rank 0 owns the frame and hands its handle to the others
    (multigpu.broadcast_bytes -- here the name of a shared-memory file stands in for the 64-byte CUDA IPC handle),
    every rank writes its stripes straight to their GLOBAL rows, and multigpu.stream_barrier (a 4-byte all-reduce) is the
    point after which rank 0 may read the frame: no gather buffers, no un-interleave pass."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from gpuraytracer_b200 import multigpu as mg
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    golden = np.load(os.path.join(GOLDEN, "config1.npz"))["rgba8"][:height_crop]
    h, w = golden.shape[:2]
    handle = None
    if rank == 0:
        name = f"/dev/shm/gprt_test_{os.getpid()}_{port}".encode().ljust(64, b"\0")
        np.lib.format.open_memmap(name.rstrip(b"\0").decode(), mode="w+", dtype=np.uint8, shape=(h, w, 4))[:] = 0
        handle = name
    handle = mg.broadcast_bytes(handle, 64, "cpu")
    assert len(handle) == 64
    path = handle.rstrip(b"\0").decode()
    frame = np.load(path, mmap_mode="r+")
    rows = mg.global_rows(h, srows, rank, world)
    frame[rows] = golden[rows]  # stands in for gprt_dispatch_stripes_host_async on the mapped frame
    frame.flush()
    flag = torch.zeros(1, dtype=torch.int32)
    mg.stream_barrier(flag)
    if rank == 0:
        np.save(out_path, np.array(frame))
    dist.barrier()
    if rank == 0:
        os.unlink(path)
    dist.destroy_process_group()


@pytest.mark.parametrize("world,height_crop,srows", [(2, 480, 8), (3, 100, 4), (2, 477, 5)])
def test_ranks_write_their_rows_straight_into_one_frame_over_gloo(tmp_path, world, height_crop, srows):
    import torch.multiprocessing as mp
    out = str(tmp_path / "full.npy")
    mp.spawn(_direct_worker, args=(world, _free_port(), height_crop, srows, out), nprocs=world, join=True)
    golden = np.load(os.path.join(GOLDEN, "config1.npz"))["rgba8"][:height_crop]
    np.testing.assert_array_equal(np.load(out), golden)


class _FakeIpcLib:
    """Stands in for libgprt.so's gprt_ipc_* on CPU: hands out fake pointers; `fail` names the call that returns an error."""

    def __init__(self, fail=None):
        self.fail, self.live, self.next = fail, set(), 0x1000

    def _new(self, p):
        self.next += 0x1000
        self.live.add(self.next)
        p._obj.value = self.next

    def gprt_ipc_alloc(self, dev, nbytes, p, h):
        if self.fail == "alloc":
            return -4
        self._new(p)
        h._obj.bytes[0] = 1 + len(self.live)
        return 0

    def gprt_ipc_open(self, dev, h, p):
        if self.fail == "open":
            return -7
        self._new(p)
        return 0

    def gprt_ipc_close(self, dev, p):
        self.live.discard(int(p))
        return 0

    gprt_ipc_free = gprt_ipc_close


def _peer_failure_worker(rank, world, port, failing_rank, failing_call, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from gpuraytracer_b200 import multigpu as mg
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fake = _FakeIpcLib(fail=failing_call if rank == failing_rank else None)
    frames = mg.PeerFrames(3, 1024, rank, world, 0, exchange=lambda blob: mg.broadcast_bytes(blob, 64 * 3, "cpu"), barrier=dist.barrier, _lib=fake)
    bad = torch.tensor([1 if frames.error else 0])
    dist.all_reduce(bad, op=dist.ReduceOp.MAX)  # what bench.py does: every rank learns whether any rank failed
    frames.close()                                # collective, whatever happened
    with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
        f.write(f"{int(bad.item())} {1 if frames.error else 0} {len(fake.live)}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("failing_rank,failing_call,expect_bad", [(None, None, 0), (0, "alloc", 1), (1, "open", 1)])
def test_peer_frames_fail_collectively(tmp_path, failing_rank, failing_call, expect_bad):
    """A rank that cannot allocate (owner) or map (any other rank) the peer frames must not leave the others hanging in
    the handle exchange: all ranks finish construction, agree on the outcome, and close() frees / unmaps everything."""
    import torch.multiprocessing as mp
    world = 3
    mp.spawn(_peer_failure_worker, args=(world, _free_port(), failing_rank, failing_call, str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        bad, own_error, live = (int(x) for x in open(tmp_path / f"rank{rank}.txt").read().split())
        assert bad == expect_bad
        assert live == 0
        if failing_call == "alloc":
            assert own_error == 1          # the owner failed, the others saw an empty handle blob
        elif failing_call == "open":
            assert own_error == (1 if rank == failing_rank else 0)


@pytest.mark.gpu
@pytest.mark.parametrize("world,srows,height", [(2, 8, 180), (4, 5, 180), (8, 3, 181), (3, 7, 90)])
def test_striped_dispatch_equals_full_frame(world, srows, height):
    """All ranks' stripes (each rendered by gprt_dispatch_stripes on this one GPU) compose to a frame
    bit-identical to the full dispatch, hit ids included."""
    import torch
    from gpuraytracer_b200 import Renderer, default_config
    cfg = default_config()
    cfg.max_recursion_depth = cfg.metaballs_count = cfg.metaball_max_steps = cfg.sdf_max_steps = 0
    with Renderer(scene_id=0, width=320, height=height, device=0, config=cfg) as r:
        r.on_init()
        r.on_update(0.3)
        full, full_ids = r.on_render(want_hit_ids=True)
        r.set_stream(torch.cuda.current_stream().cuda_stream)
        pad = multigpu.max_local_rows(height, srows, world)
        gathered, gathered_ids = [], []
        for rank in range(world):
            buf = torch.zeros((pad, 320, 4), dtype=torch.uint8, device="cuda")
            ids = torch.zeros((pad, 320), dtype=torch.int32, device="cuda")
            r.do_raytracing_striped(buf.data_ptr(), srows, rank, world, d_hit_ids_ptr=ids.data_ptr())
            r.sync()
            gathered.append(buf)
            gathered_ids.append(ids)
        out = multigpu.compose(torch.zeros((height, 320, 4), dtype=torch.uint8, device="cuda"), gathered, height, srows, world)
        out_ids = multigpu.compose(torch.zeros((height, 320), dtype=torch.int32, device="cuda"), gathered_ids, height, srows, world)
    np.testing.assert_array_equal(out.cpu().numpy(), full)
    np.testing.assert_array_equal(out_ids.cpu().numpy().view(np.uint32), full_ids)
