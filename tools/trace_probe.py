#!/usr/bin/env python3
"""C5-style incoherent rays through the ray-list kernel (k_trace) on one GPU: throughput with CUDA events; under
`ncu --set full -k regex:k_trace` it is the divergence-stress capture north_star asks for.
usage: python tools/trace_probe.py [--log2n 22] [--reps 3]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import atray_b200 as A  # noqa: E402
from helpers import random_rays  # noqa: E402
from tools import gen_standin_obj  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--log2n", type=int, default=22)
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
path = gen_standin_obj.ensure(os.path.join(os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes"), "standin_272x256.obj"))
m = A.load_model_data(path, 8)
A.place_default(m, 0.0)
sc = A.Scene(m)
A.prep_scene(sc)
ctx = sc.ctx
n = 1 << args.log2n
rays = random_rays(sc.model.dump()["aabb"], n, 1234)
d_r, d_p, d_t = ctx.dev_alloc(n * 24), ctx.dev_alloc(n * 4), ctx.dev_alloc(n * 4)
ctx.from_host(d_r, rays)
ms = []
for i in range(args.reps + 1):
    ctx.trace_rays_device(d_r, n, d_p, d_t)
    ctx.sync()
    if i >= 1:
        ms.append(ctx.last_kernel_ms())
print("k_trace: %d incoherent rays, %.2f ms per launch, %.1f Mrays/s" % (n, np.mean(ms), n / np.mean(ms) / 1e3), flush=True)
