#!/usr/bin/env python3
"""Times rank 0's share of the C4 frame for world = 1, 2, 4, 8 on ONE GPU (the other ranks' bands are simply not
rendered): shows how the per-rank kernel time splits into the replicated view pass and the partitioned render.
usage: python tools/part_probe.py [--once]   (--once: one frame per world size, for an ncu launch list)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import atray_b200 as A  # noqa: E402
from tools import gen_standin_obj  # noqa: E402

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
reps = 1 if "--once" in sys.argv else 10
for world in (1, 2, 4, 8):
    opts = ctx.opts(rank=0, world=world, band_rows=8)
    ms = []
    for i in range(reps + 2):
        ctx.dev_memset(flush, 0, 256 << 20)
        ctx.render_device(cam, opts, *planes)
        ctx.sync()
        if i >= 2 or reps == 1:
            ms.append(ctx.last_kernel_ms())
    print("world %d rank 0: %.1f us per frame (min %.1f)" % (world, 1e3 * sum(ms) / len(ms), 1e3 * min(ms)), flush=True)
