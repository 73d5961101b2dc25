"""Sharding of the path across the GPUs of a box: row stripes of one frame, or whole frames.

The path shards by pixels: the scene is replicated (every rank builds the same LBVH from the same
inputs), rank r renders the row stripes r, r + world, ... (gprt_dispatch_stripes) and the RGBA8 rows
are gathered to rank 0 -- the only exchange step of the path.  Stripes are a few rows high and
interleaved so that the expensive regions of the image (the fractal pyramid, the metaballs) are
spread over all ranks.

torch.distributed is plumbing here: NCCL over NVLink on the GPU box (gather = grouped send/recv),
gloo in the CPU tests.  Nothing is computed in this module.
"""
import numpy as np

from ._capi import lib


def stripe_rows_for(height: int, world: int, max_rows: int = 8) -> int:
    """Largest stripe height <= max_rows that gives every rank the same number of whole stripes
    (so that the gather needs no padding); 1 if there is none."""
    if world <= 1:
        return max(int(height), 1)
    for s in range(max_rows, 0, -1):
        if height % (s * world) == 0:
            return s
    return 1


def local_rows(height: int, stripe_rows: int, rank: int, world: int) -> int:
    """Rows rank `rank` renders (C ABI: gprt_stripe_local_rows)."""
    return int(lib().gprt_stripe_local_rows(int(height), int(stripe_rows), int(rank), int(world)))


def global_rows(height: int, stripe_rows: int, rank: int, world: int) -> np.ndarray:
    """Global row index of every local row of `rank`, in local order."""
    n = local_rows(height, stripe_rows, rank, world)
    l = np.arange(n, dtype=np.int64)
    return ((l // stripe_rows) * world + rank) * stripe_rows + l % stripe_rows


def max_local_rows(height: int, stripe_rows: int, world: int) -> int:
    return max(local_rows(height, stripe_rows, r, world) for r in range(world))


def gather_stripes(local, gathered, dst=0, group=None):
    """Gather every rank's packed stripe buffer (same shape on all ranks, padded to max_local_rows)
    to rank `dst`.  `gathered` is a list of world tensors on dst, None elsewhere."""
    import torch.distributed as dist
    dist.gather(local, gathered if dist.get_rank(group) == dst else None, dst=dst, group=group)


def frame_owner(frame: int, world: int) -> int:
    """Frame sharding (small frames: a frame's serial floor does not shrink with its pixel count, so
    whole frames are dealt round-robin instead of stripes of each): rank that renders global frame `frame`."""
    return int(frame) % max(int(world), 1)


def gather_frames(local, gathered, dst=0, group=None):
    """Gather the whole frame every rank rendered this step to rank `dst`: gathered[k] is global frame
    step * world + k.  Same collective as gather_stripes."""
    gather_stripes(local, gathered, dst=dst, group=group)


def compose(full, gathered, height: int, stripe_rows: int, world: int):
    """Un-interleave the gathered stripe buffers into the full frame `full` [H, W, 4] (torch tensors
    or numpy arrays; index copies only)."""
    if height % (stripe_rows * world) == 0:
        # regular case: one strided copy per rank
        groups = height // (stripe_rows * world)
        view = full.reshape(groups, world, stripe_rows, *full.shape[1:])
        for r in range(world):
            src = gathered[r][:groups * stripe_rows].reshape(groups, stripe_rows, *full.shape[1:])
            if hasattr(view, "copy_"):
                view[:, r].copy_(src)
            else:
                view[:, r] = src
        return full
    for r in range(world):
        rows = global_rows(height, stripe_rows, r, world)
        if len(rows) == 0:
            continue
        if hasattr(full, "index_copy_"):
            import torch
            full.index_copy_(0, torch.as_tensor(rows, device=full.device), gathered[r][:len(rows)])
        else:
            full[rows] = gathered[r][:len(rows)]
    return full


class SharedHostFrames:
    """`n_slots` host buffers of `slot_bytes` that EVERY rank of the box can write with an asynchronous device-to-host
    copy: POSIX shared memory mapped by all ranks and page-locked in each with cudaHostRegister (one rank: plain
    page-locked memory).  Each rank then delivers its rows / frames over its own PCIe link straight into the frame the
    host consumer (rank 0) reads, instead of all pixels funnelling through rank 0's GPU and its one link.  `barrier`
    is called between creating, mapping and unlinking the file (dist.barrier on N ranks)."""

    def __init__(self, tag: str, n_slots: int, slot_bytes: int, rank: int, world: int, barrier=None):
        import mmap
        import os

        import torch
        self.n_slots, self.slot_bytes, self.world = int(n_slots), int(slot_bytes), int(world)
        self._registered = False
        self._mm = None
        total = self.n_slots * self.slot_bytes
        if world <= 1:
            self._t = torch.empty(total, dtype=torch.uint8, pin_memory=True)
            self.array = self._t.numpy()
        else:
            path = f"/dev/shm/gprt_{tag}"
            if rank == 0:
                fd = os.open(path, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o600)
                os.ftruncate(fd, total)
                os.close(fd)
            barrier()
            fd = os.open(path, os.O_RDWR)
            self._mm = mmap.mmap(fd, total)
            os.close(fd)
            self.array = np.frombuffer(self._mm, dtype=np.uint8)
            err = torch.cuda.cudart().cudaHostRegister(self.array.ctypes.data, total, 0)
            if int(err) != 0:
                raise RuntimeError(f"cudaHostRegister failed: {err}")
            self._registered = True
            barrier()
            if rank == 0:
                os.unlink(path)

    def ptr(self, slot: int, offset: int = 0) -> int:
        return int(self.array.ctypes.data) + int(slot) * self.slot_bytes + int(offset)

    def view(self, slot: int, shape):
        return self.array[slot * self.slot_bytes:(slot + 1) * self.slot_bytes].reshape(shape)

    def close(self):
        if self._registered:
            import torch
            torch.cuda.cudart().cudaHostUnregister(self.array.ctypes.data)
            self._registered = False
        self.array = None
        self._t = None
        if self._mm is not None:
            try:
                self._mm.close()
            except BufferError:
                pass
            self._mm = None


class _DevicePtr:
    """A raw device pointer as a __cuda_array_interface__ object (what torch.as_tensor wraps without copying)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


class PeerFrames:
    """`n` DEVICE frames of `frame_bytes` on rank `owner`'s GPU that every rank of the box writes directly (one process
    per GPU; C ABI: gprt_ipc_alloc / gprt_ipc_open).  With `ptr(k)` as the frame of on_render_striped_async every rank's
    rows go over NVLink straight to their global rows of the owner's frame: the frame needs no gather and no un-interleave
    pass, and the owner's GPU does nothing for it.  `exchange(bytes | None) -> bytes` hands the owner's handles to all
    ranks (a broadcast); `barrier()` separates unmapping from freeing in close()."""

    def __init__(self, n: int, frame_bytes: int, rank: int, world: int, device: int, exchange=None, barrier=None, owner: int = 0, _lib=None):
        """Collective over the `world` ranks and safe to fail in: a rank that cannot allocate or map never leaves the others
        waiting in `exchange` -- it records the reason in `self.error` (None when all is well) and the ranks decide together
        what to do (bench.py falls back to the NCCL gather when any rank reports an error; close() stays collective)."""
        import ctypes as C

        from ._capi import GprtError, IpcHandle, check
        self.n, self.frame_bytes, self.rank, self.world, self.device, self.owner = int(n), int(frame_bytes), int(rank), int(world), int(device), int(owner)
        self._barrier = barrier or (lambda: None)
        self._ptrs = []
        self.error = None
        self._L = L = _lib or lib()
        blob = None
        if rank == owner:
            handles = []
            try:
                for _ in range(self.n):
                    p, h = C.c_void_p(), IpcHandle()
                    check(L.gprt_ipc_alloc(self.device, self.frame_bytes, C.byref(p), C.byref(h)), "ipc_alloc")
                    self._ptrs.append(int(p.value))
                    handles.append(bytes(h.bytes))
                blob = b"".join(handles)
            except GprtError as e:
                self.error = str(e)
                for p in self._ptrs:
                    L.gprt_ipc_free(self.device, p)
                self._ptrs = []
                blob = bytes(64 * self.n)  # all zeros: "no frames"
        if world > 1:
            blob = exchange(blob)
            if rank != owner:
                if not any(blob):
                    self.error = f"rank {owner} could not allocate the frames"
                else:
                    try:
                        for k in range(self.n):
                            h = IpcHandle()
                            C.memmove(h.bytes, blob[64 * k:64 * (k + 1)], 64)
                            p = C.c_void_p()
                            check(L.gprt_ipc_open(self.device, C.byref(h), C.byref(p)), "ipc_open")
                            self._ptrs.append(int(p.value))
                    except GprtError as e:
                        self.error = str(e)
                        for p in self._ptrs:
                            L.gprt_ipc_close(self.device, p)
                        self._ptrs = []

    def ptr(self, k: int) -> int:
        """Frame k as a device pointer valid in THIS process."""
        return self._ptrs[k]

    def tensor(self, k: int, shape):
        """Frame k as a torch uint8 tensor (no copy); meant for the owner, which reads the finished frame."""
        import torch
        return torch.as_tensor(_DevicePtr(self._ptrs[k], self.frame_bytes), device=torch.device("cuda", self.device)).view(*shape)

    def close(self):
        L = self._L
        if self.rank != self.owner:
            for p in self._ptrs:
                L.gprt_ipc_close(self.device, p)
            self._ptrs = []
        if self.world > 1:
            self._barrier()
        if self.rank == self.owner:
            for p in self._ptrs:
                L.gprt_ipc_free(self.device, p)
            self._ptrs = []


def broadcast_bytes(blob, nbytes: int, device, src: int = 0, group=None) -> bytes:
    """`exchange` of PeerFrames over torch.distributed: `blob` (nbytes long) on rank `src`, None elsewhere."""
    import torch
    import torch.distributed as dist
    if dist.get_backend(group) == "gloo":
        device = "cpu"
    t = torch.zeros(nbytes, dtype=torch.uint8, device=device)
    if dist.get_rank(group) == src:
        t.copy_(torch.frombuffer(bytearray(blob), dtype=torch.uint8))
    dist.broadcast(t, src=src, group=group)
    return bytes(t.cpu().numpy().tobytes())


def stream_barrier(flag, group=None):
    """All ranks' CURRENT streams meet: work enqueued after it on any rank runs after the work enqueued before it on every
    rank (a 4-byte all-reduce; the call does not block the host).  After it the rows every rank copied into a PeerFrames
    frame are in place."""
    import torch.distributed as dist
    dist.all_reduce(flag, group=group)
