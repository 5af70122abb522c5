"""CPU: host-side multi-GPU logic — band partition arithmetic and the IPC-handle exchange — with world_size 2 on the
gloo backend (no GPU). The data path itself (kernels storing into GPU 0's frame) is covered on the GPU by
test_partition_union_equals_whole and by bench.py --gpus N."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("H,band,world", [(2160, 8, 1), (2160, 8, 2), (2160, 8, 8), (301, 12, 3), (7, 4, 4), (1080, 16, 5)])
def test_band_partition_covers_every_row_once(H, band, world, built):
    from atray_b200 import multi
    seen = np.zeros(H, int)
    for r in range(world):
        rows = multi.rows_of_rank(H, band, r, world)
        seen[rows] += 1
        assert ((rows // band) % world == r).all()
        assert multi.rays_of_rank(100, H, 3, band, r, world) == rows.size * 300
    assert (seen == 1).all()
    assert multi.n_bands(H, band) == -(-H // band)
    sizes = [multi.rows_of_rank(H, band, r, world).size for r in range(world)]
    if H >= band * world:
        assert max(sizes) - min(sizes) <= band  # interleaving balances the ranks to within one band


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from atray_b200 import multi
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)

    class FakeCtx:  # stands in for the device context: records what FrameGather asks of it
        def __init__(self):
            self.log = []

        def dev_alloc(self, n):
            self.log.append(("alloc", n))
            return 0x1000 * (len(self.log) + 1)

        def dev_free(self, p):
            self.log.append(("free", p))

        def ipc_export(self, p):
            return (b"H%015d" % p) * 4

        def ipc_open(self, h):
            self.log.append(("open", h))
            return int(h[1:16]) + 7

        def ipc_close(self, p):
            self.log.append(("close", p))

        def view_share_create(self, w, h):
            self.log.append(("share_create", w, h))
            return 0xB10C0 + 0x100 * rank, 4096

        def view_share_connect(self, r, n, peers):
            self.log.append(("share_connect", r, n, tuple(peers)))

        def view_share_off(self):
            self.log.append(("share_off",))

    ctx = FakeCtx()
    # the shared view pass: every rank exports its block, maps the others' and connects with its own block at its own index
    sh = multi.ViewShare(ctx, dist, rank, world, 640, 360)
    conn = [e for e in ctx.log if e[0] == "share_connect"][0]
    own = 0xB10C0 + 0x100 * rank
    assert conn[1] == rank and conn[2] == world and conn[3][rank] == own
    assert all(conn[3][r] == 0xB10C0 + 0x100 * r + 7 for r in range(world) if r != rank)   # ipc_open of rank r's exported block
    sh.close()
    assert ctx.log[-2][0] == "share_off" and ctx.log[-1] == ("close", 0xB10C0 + 0x100 * (1 - rank) + 7)
    del ctx.log[:]
    g = multi.FrameGather(ctx, dist, rank, world, 64, 32)
    planes = dict(g.planes)
    g.close()
    n0 = len(ctx.log)
    g2 = multi.FrameGather(ctx, dist, rank, world, 64, 32, local=("t", "prim"))   # only the bitmap is gathered
    assert set(g2.planes) == {"bgrx", "t", "prim"} and g2.local == ("t", "prim")
    kinds = [e[0] for e in ctx.log[n0:]]
    assert kinds == (["alloc"] * 3 if rank == 0 else ["alloc", "alloc", "open"]), kinds
    g2.close()
    del ctx.log[n0:]
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, planes, ctx.log))


def test_frame_gather_handle_exchange_gloo_world2(built):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(2):
        r, planes, log = q.get(timeout=120)
        res[r] = (planes, log)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    p0, log0 = res[0]
    p1, log1 = res[1]
    assert set(p0) == set(p1) == {"bgrx", "t", "prim"}
    assert [e[0] for e in log0] == ["alloc"] * 3 + ["free"] * 3 and all(e[1] == 64 * 32 * 4 for e in log0[:3])
    for k in p0:  # rank 1 mapped exactly the allocations rank 0 exported
        assert p1[k] == p0[k] + 7
    assert [e[0] for e in log1] == ["open"] * 3 + ["close"] * 3


def _bitmap_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from atray_b200 import multi
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, W, band = 50, 16, 8
    name = "atray_test_%d" % port
    bm = multi.SharedBitmap(name, (H, W), np.uint32, create=True, register=False) if rank == 0 else None
    dist.barrier()
    if rank != 0:
        bm = multi.SharedBitmap(name, (H, W), np.uint32, create=False, register=False)
    bar = multi.ShmBarrier(name + "_bar", rank, world, create=True) if rank == 0 else None
    dist.barrier()
    if rank != 0:
        bar = multi.ShmBarrier(name + "_bar", rank, world, create=False)
    for _ in range(200):                                # the shared-memory barrier that brackets the timed e2e frames
        bar.wait()
    assert int(bar.slots[:, 0].min()) == 200
    dist.barrier()
    bar.close()
    rows = multi.rows_of_rank(H, band, rank, world)     # what atray_render_begin writes for this partition
    bm.array[rows] = (rank + 1) * 1000 + rows[:, None]
    bm.array.flush()
    dist.barrier()
    ok = True
    if rank == 0:
        for r in range(world):
            rr = multi.rows_of_rank(H, band, r, world)
            ok = ok and bool((np.asarray(bm.array)[rr] == (r + 1) * 1000 + rr[:, None]).all())
    dist.barrier()
    path = bm.path
    bm.close()
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok, os.path.exists(path)))


def test_shared_host_bitmap_gloo_world2(built):
    """Two rank processes map one host bitmap and write their own scanline bands into it (the multi-GPU e2e output path;
    page-locking is skipped here, there is no CUDA driver on the CPU box)."""
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_bitmap_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res) and not any(left for _, _, left in res)


def test_rank_placement_spreads_over_the_node():
    from atray_b200 import multi
    assert multi.spread_devices(1, 8) == [0]
    assert multi.spread_devices(2, 8) == [0, 4]
    assert multi.spread_devices(4, 8) == [0, 2, 4, 6]
    assert multi.spread_devices(8, 8) == list(range(8))
    assert multi.spread_devices(2, 2) == [0, 1] and multi.spread_devices(3, 4) == [0, 1, 2]
    import bench
    for w, v in ((2, 8), (4, 8), (8, 8), (2, 2), (2, 3)):
        assert [bench.pick_device(r, w, v) for r in range(w)] == multi.spread_devices(w, v)


@pytest.mark.parametrize("p0,p1,chunk,world", [(0, 100000, 4096, 8), (17, 9001, 100, 3), (0, 10, 4096, 2), (5, 5, 64, 4)])
def test_c5_chunks_cover_every_pixel_once(p0, p1, chunk, world, built):
    """The C5 split of SURVEY 8(e): contiguous chunks of the pixel range dealt round-robin, every pixel traced by exactly one rank."""
    from atray_b200 import multi
    parts = [multi.chunk_pixels_of_rank(p0, p1, chunk, r, world) for r in range(world)]
    allp = np.sort(np.concatenate(parts)) if parts else np.zeros(0, np.int64)
    assert (allp == np.arange(p0, p1)).all()
    for r, p in enumerate(parts):
        if p.size:
            k = (p - p0) // chunk
            assert (k % world == r).all()
            assert all(len(run) <= chunk for run in np.split(p, np.flatnonzero(np.diff(p) != 1) + 1))
