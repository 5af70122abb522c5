#!/usr/bin/env python3
"""Renders the C4 frame a few times on one GPU (device planes, no copy-out) and prints the frame's kernel time; under
`ncu --metrics gpu__time_duration.sum` it gives the per-kernel split of one frame.
With --host it times the end-to-end frame instead (pinned host bitmap, slab-pipelined copy-out, wall clock), which is
what ATRAY_SLABS / ATRAY_SLAB_WEIGHTS tune.
usage: python tools/frame_probe.py [--reps N] [--raster] [--world N] [--host]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import atray_b200 as A  # noqa: E402
from tools import gen_standin_obj  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--raster", action="store_true")
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--count", action="store_true")
ap.add_argument("--host", action="store_true")
ap.add_argument("--no-flush", action="store_true")
ap.add_argument("--band-rows", type=int, default=8)
ap.add_argument("--shards", type=int, default=0, help="run the view pass as N shard launches on this GPU (atray_debug_view_shards)")
args = ap.parse_args()
W, H = 3840, 2160
path = gen_standin_obj.ensure(os.path.join(os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes"), "standin_272x256.obj"))
m = A.load_model_data(path, 8)
A.place_default(m, 0.0)
sc = A.Scene(m)
A.prep_scene(sc)
ctx = sc.ctx
if args.shards:
    ctx.debug_view_shards(args.shards)
cam = A.set_camera(W, H)
if args.host:
    import numpy as np
    tex = A.PinnedArray((H, W), np.uint32)
    info = A.RenderInfo(cam, sc, tex.array, opts=ctx.opts(rank=0, world=1, band_rows=args.band_rows, raster=args.raster))
    best = 1e9
    for rep in range(3):
        for i in range(args.reps + 2):
            if i == 2:
                ctx.sync()
                t0 = time.perf_counter()
            A.start_render_from_camera(info)
            while A.wait_for_render_from_camera_to_finish(info, None, 33):
                pass
        dt = (time.perf_counter() - t0) / args.reps
        best = min(best, dt)
    print("host frame (slabs %s weights %s): %.1f us per frame (best of 3 x %d)" % (
        os.environ.get("ATRAY_SLABS", "default"), os.environ.get("ATRAY_SLAB_WEIGHTS", "default plan" if "ATRAY_SLABS" not in os.environ else "uniform"), 1e6 * best, args.reps), flush=True)
    sys.exit(0)
planes = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
flush = ctx.dev_alloc(256 << 20)
opts = ctx.opts(rank=args.rank, world=args.world, band_rows=args.band_rows, raster=args.raster, count_work=args.count)
ms, rms, vs = [], [], []
for i in range(args.reps + 2):
    if not args.no_flush:
        ctx.dev_memset(flush, 0, 256 << 20)
    ctx.render_device(cam, opts, *planes)
    ctx.sync()
    if i >= 2:
        ms.append(ctx.last_kernel_ms())
        if not args.raster:
            rms.append(ctx.last_render_kernel_ms())
            vs.append(ctx.last_view_split_ms())
print("world %d rank %d raster=%s: %.1f us per frame (min %.1f)" % (args.world, args.rank, args.raster, 1e3 * sum(ms) / len(ms), 1e3 * min(ms)), flush=True)
if rms:
    print("  render kernel alone %.1f us, view pass %.1f us (k_view_bounds %.1f, k_view_nodes %.1f)" % (
        1e3 * sum(rms) / len(rms), 1e3 * (sum(ms) - sum(rms)) / len(ms), 1e3 * sum(a for a, _ in vs) / len(vs), 1e3 * sum(b for _, b in vs) / len(vs)), flush=True)
if args.count:
    print(ctx.counters())
