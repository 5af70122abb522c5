#!/usr/bin/env python3
"""Per-tile timing of the culled render kernel on the C4 frame (ATRAY_TILE_LOG=1): duration per cost class, and what the
kernel's last microseconds were spent on. usage: ATRAY_TILE_LOG=1 python tools/tile_log_probe.py [--world N]"""
import argparse
import os
import sys

import numpy as np

os.environ.setdefault("ATRAY_TILE_LOG", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import atray_b200 as A  # noqa: E402
from tools import gen_standin_obj  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--brief", action="store_true")
args = ap.parse_args()
W, H = 3840, 2160
path = gen_standin_obj.ensure(os.path.join(os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes"), "standin_272x256.obj"))
m = A.load_model_data(path, 8)
A.place_default(m, 0.0)
sc = A.Scene(m)
A.prep_scene(sc)
ctx = sc.ctx
cam = A.set_camera(W, H)
planes = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
flush = ctx.dev_alloc(256 << 20)
opts = ctx.opts(rank=args.rank, world=args.world, band_rows=8)
for i in range(4):
    ctx.dev_memset(flush, 0, 256 << 20)
    ctx.render_device(cam, opts, *planes)
    ctx.sync()
n_bands = (H + 7) // 8
my_bands = (n_bands - args.rank + args.world - 1) // args.world
n_tiles = my_bands * 2 * ((W + 7) // 8)
log = ctx.tile_log(n_tiles)
busy = log[log[:, 1] > 0]
start = busy[:, 0].astype(np.int64)
t0 = start.min()
start = (start - t0) & 0xffffffff
dur = busy[:, 1].astype(np.int64)
end = start + dur
print("render kernel %.1f us (frame %.1f us); %d busy tiles, traced from %.1f to %.1f us after the first tile started" % (
    1e3 * ctx.last_render_kernel_ms(), 1e3 * ctx.last_kernel_ms(), len(busy), 0.0, end.max() / 1e3))
print("class  tiles   cost(min..max)   mean us   p50    p90    max    first start  last start  last end")
for c in range(15, -1, -1):
    k = busy[:, 3] == c
    if not k.any():
        continue
    d = dur[k] / 1e3
    print("%5d %6d %8d..%-7d %8.1f %6.1f %6.1f %6.1f %12.1f %11.1f %9.1f" % (
        c, k.sum(), busy[k, 2].min(), busy[k, 2].max(), d.mean(), np.percentile(d, 50), np.percentile(d, 90), d.max(),
        start[k].min() / 1e3, start[k].max() / 1e3, end[k].max() / 1e3))
if args.brief:
    print("world %d rank %d: busy %d, per class (7..0) %s, longest tile %.1f us" % (
        args.world, args.rank, len(busy), [int((busy[:, 3] == c).sum()) for c in range(7, -1, -1)], dur.max() / 1e3))
    sys.exit(0)
last_start = start.max()
print("list exhausted at %.1f us; tiles still running then: %d; of those, per class:" % (last_start / 1e3, int((end > last_start).sum())),
      {int(c): int(((end > last_start) & (busy[:, 3] == c)).sum()) for c in np.unique(busy[:, 3])})
late = np.argsort(end)[-12:]
print("last 12 tiles to finish: (class, cost, start us, duration us)")
for i in late:
    print("   ", int(busy[i, 3]), int(busy[i, 2]), round(start[i] / 1e3, 1), round(dur[i] / 1e3, 1))
print("correlation(log2 cost, duration) = %.3f" % np.corrcoef(np.log2(busy[:, 2].astype(np.float64) + 1), dur)[0, 1])
if os.path.isdir("gpurun_out"):
    np.save("gpurun_out/tile_log_world%d.npy" % args.world, log)
