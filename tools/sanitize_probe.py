#!/usr/bin/env python3
"""Small frames through every kernel family (culled tile kernel, opt-in rasteriser, unculled, no-tree branch, jittered,
multi-bounce with primitives, ray list, band partition with host copy-out) for `compute-sanitizer --tool memcheck|racecheck`.
usage: compute-sanitizer --tool memcheck python tools/sanitize_probe.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import atray_b200 as A  # noqa: E402
from helpers import PATH_PLANES, PATH_SPHERES, random_rays  # noqa: E402
from tools import gen_standin_obj  # noqa: E402

path = gen_standin_obj.ensure("/tmp/atray_meshes/knot_48x32.obj", 48, 32)
m = A.load_model_data(path, 2)
A.place_default(m, 0.0)
sc = A.Scene(m, max_no_faces_per_node=64)
A.prep_scene(sc)
ctx = sc.ctx
W, H = 333, 187
cam = A.set_camera(W, H)
ref = ctx.render(cam, ctx.opts(cull=False))
for name, o in (("culled", ctx.opts()), ("raster", ctx.opts(raster=True)), ("counted", ctx.opts(count_work=True))):
    out = ctx.render(cam, o)
    assert (out["prim"] == ref["prim"]).all() and (out["bgrx"] == ref["bgrx"]).all(), name
big = A.set_camera(1280, 832)           # >= 1 Mpx: slab-pipelined copy-out
a, b = ctx.render(big), ctx.render(big, ctx.opts(cull=False))
assert (a["prim"] == b["prim"]).all()
host = {k: np.zeros((832, 1280), np.uint32) for k in ("bgrx", "t", "prim")}
for rank in range(3):
    ctx.render_begin(big, ctx.opts(rank=rank, world=3, band_rows=12), host["bgrx"], host["t"].view(np.float32), host["prim"])
    while ctx.render_poll(1000):
        pass
assert (host["prim"] == a["prim"]).all() and (host["bgrx"] == a["bgrx"]).all()
jit = A.set_camera(W, H, anti_aliasing=1, samples_per_pixel=4)
j1, j2 = ctx.render(jit), ctx.render(jit, ctx.opts(cull=False))
assert (j1["bgrx"] == j2["bgrx"]).all()
prim, t = ctx.trace_rays(random_rays(sc.model.dump()["aabb"], 20000, 3))
ctx.use_kd_tree(False)
nt = ctx.render(A.set_camera(96, 54))
ctx.use_kd_tree(True)
ctx.set_primitives(PATH_SPHERES, PATH_PLANES)
p = ctx.render(A.set_camera(160, 90, anti_aliasing=1, samples_per_pixel=2, bounce_limit=4))
print("sanitize probe ok: %d hits, %d ray hits, path casts %d" % (int((ref["prim"] != A.MISS).sum()), int((prim != A.MISS).sum()), ctx.last_ray_casts()))
