"""Multi-GPU plumbing for the render path: one process per GPU, scene replicated, image split into interleaved
scanline bands (band b → rank b mod world), every rank's kernel stores its bands straight into GPU 0's frame through
a CUDA-IPC mapping (peer stores over NVLink) — there is no collective on the data path and NCCL is not needed for it
(north_star). torch.distributed is used only to hand the 64-byte IPC handles around and for barriers.

The reference's counterpart is the tile work queue of renderer.cpp:403-445 (tiles pulled by 8 threads)."""
import ctypes as C
import os

import numpy as np

PLANES = ("bgrx", "t", "prim")


def n_bands(height, band_rows):
    return (height + band_rows - 1) // band_rows


def rows_of_rank(height, band_rows, rank, world):
    """Scanlines owned by `rank` (same arithmetic as k_render's tile → (x, y) mapping)."""
    rows = np.arange(height)
    return rows[(rows // band_rows) % world == rank]


def rays_of_rank(width, height, spp, band_rows, rank, world):
    return int(rows_of_rank(height, band_rows, rank, world).size) * width * spp


def exchange_handles(dist, rank, payload):
    """Rank 0's payload (dict of plane → 64-byte IPC handle) is broadcast to every rank; works on gloo and nccl."""
    box = [payload if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


class FrameGather:
    """GPU 0 owns the frame planes; other ranks map them. `planes[p]` is a device pointer valid in this process."""

    def __init__(self, ctx, dist, rank, world, width, height, planes=PLANES, local=()):
        """`local`: planes that are NOT gathered — every rank keeps its own full-size plane and fills only its bands
        (the live view needs the BGRX bitmap on GPU 0, north_star; t and primitive ids can stay where they were made,
        which cuts GPU 0's NVLink ingress from 12 to 4 bytes per pixel)."""
        self.ctx, self.rank, self.world = ctx, rank, world
        self.nbytes = width * height * 4
        self.planes, self._owned, self._opened = {}, [], []
        self.local = tuple(p for p in local if p in planes) if world > 1 else ()
        for p in self.local:
            ptr = ctx.dev_alloc(self.nbytes)
            self._owned.append(ptr)
            self.planes[p] = ptr
        planes = tuple(p for p in planes if p not in self.local)
        handles = None
        if rank == 0:
            handles = {}
            for p in planes:
                ptr = ctx.dev_alloc(self.nbytes)
                self._owned.append(ptr)
                self.planes[p] = ptr
                if world > 1:
                    handles[p] = ctx.ipc_export(ptr)
        if world > 1:
            handles = exchange_handles(dist, rank, handles)
            if rank != 0:
                for p in planes:
                    ptr = ctx.ipc_open(handles[p])
                    self._opened.append(ptr)
                    self.planes[p] = ptr

    def close(self):
        for p in self._opened:
            self.ctx.ipc_close(p)
        for p in self._owned:
            self.ctx.dev_free(p)
        self._opened, self._owned = [], []


def chunk_pixels_of_rank(p0, p1, chunk_px, rank, world):
    """Pixel indices of [p0, p1) that rank `rank` of `world` traces in a C5 job (atray_trace_incoherent_device): the range is cut
    into chunks of `chunk_px` pixels and chunk k belongs to rank k % world (render.cu: k_c5) — contiguous chunks, no exchange."""
    import numpy as np
    rel = np.arange(p1 - p0, dtype=np.int64)
    return p0 + rel[(rel // chunk_px) % world == rank]


class ViewShare:
    """Shared per-frame view pass (atray_view_share_*): every rank computes the pixel ranges of 1/world of the triangle
    slots and stores them into all ranks' tables over NVLink. One process per GPU: the blocks travel as CUDA-IPC handles
    (all_gather_object over the control-plane backend); `connect_local` wires several contexts of ONE process instead."""

    def __init__(self, ctx, dist, rank, world, width, height):
        self.ctx, self.world, self._opened = ctx, world, []
        if world <= 1:
            return
        block, _ = ctx.view_share_create(width, height)
        handles = [None] * world
        dist.all_gather_object(handles, ctx.ipc_export(block))
        peers = []
        for r in range(world):
            if r == rank:
                peers.append(block)
            else:
                ptr = ctx.ipc_open(handles[r])
                self._opened.append(ptr)
                peers.append(ptr)
        ctx.view_share_connect(rank, world, peers)

    @staticmethod
    def connect_local(ctxs, width, height):
        """Contexts of one process, one per GPU (rank r = ctxs[r]): peer access both ways, blocks handed over directly."""
        world = len(ctxs)
        for a in ctxs:
            for b in ctxs:
                if a is not b:
                    a.enable_peer(b)
        blocks = [c.view_share_create(width, height)[0] for c in ctxs]
        for r, c in enumerate(ctxs):
            c.view_share_connect(r, world, blocks)

    def close(self):
        if self.world > 1:
            self.ctx.view_share_off()
        for p in self._opened:
            self.ctx.ipc_close(p)
        self._opened = []


class SharedBitmap:
    """The host bitmap (Texture.bmb.buffer_memory) of a multi-process job: a file in /dev/shm mapped by every rank and
    page-locked in each process (atray_host_register). With a partition, atray_render_begin writes exactly the rank's
    scanlines, so all ranks render into this one bitmap — the analogue of the reference's pool threads writing their
    tiles into one texture (renderer.cpp:294-369) — and the N device→host copies run in parallel, each over its own GPU's
    PCIe link. Rank 0 creates the file (call with create=True first, then a barrier, then the other ranks open it)."""

    def __init__(self, name, shape, dtype=np.uint32, create=False, register=True):
        from . import api
        self.path = os.path.join("/dev/shm", name)
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.owner = create
        if create:
            with open(self.path, "wb") as f:
                f.truncate(self.nbytes)
        self.array = np.memmap(self.path, dtype=dtype, mode="r+", shape=tuple(shape))
        self._api, self._registered = api, False
        if register:
            api._ck(api._L.atray_host_register(C.c_void_p(self.array.ctypes.data), self.nbytes))
            self._registered = True

    def close(self):
        if self.array is not None:
            if self._registered:
                self._api._L.atray_host_unregister(C.c_void_p(self.array.ctypes.data))
            self.array = None
            if self.owner and os.path.exists(self.path):
                os.remove(self.path)


class ShmBarrier:
    """Barrier between the rank processes of one node through a file in /dev/shm: rank r publishes a step counter in its
    own slot and spins until every slot has reached it (one writer per slot, no atomics needed). Costs microseconds where
    a NCCL barrier costs a kernel launch and a device synchronisation — it brackets sub-millisecond frames."""

    def __init__(self, name, rank, world, create=False):
        self.path = os.path.join("/dev/shm", name)
        self.rank, self.world, self.owner, self.step = rank, world, create, 0
        if create:
            with open(self.path, "wb") as f:
                f.truncate(8 * 16 * world)
        self.slots = np.memmap(self.path, dtype=np.int64, mode="r+", shape=(world, 16))   # one cache line pair per rank

    def wait(self):
        self.step += 1
        self.slots[self.rank, 0] = self.step
        while int(self.slots[:, 0].min()) < self.step:
            pass

    def close(self):
        self.slots = None
        if self.owner and os.path.exists(self.path):
            os.remove(self.path)


def spread_devices(world, visible=None):
    """GPUs for the `world` local ranks of one node: evenly spread over the `visible` devices when there are at least twice
    as many (rank r → r * (visible // world)), else rank r → r. On the 8-GPU B200 boxes GPUs 0-3 and 4-7 reach host memory
    through two different paths, which is what bounds the multi-GPU host frame (profiles/r01d_summary.md, r01e)."""
    if visible is None:
        from . import api
        visible = api.device_count()
    if world > 1 and visible >= 2 * world:
        return [r * (visible // world) for r in range(world)]
    return list(range(world))
