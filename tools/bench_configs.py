#!/usr/bin/env python3
"""Device-resident throughput of every BASELINE.json config on ONE GPU (the headline bench.py line is C4 only):
C1 Simple 640x360 · C2 Monkey 1920x1080 · C3 Deer 1920x1080 16 spp jittered · C4 stand-in 3840x2160 · C5 stand-in,
incoherent random-direction rays through the ray-list kernel (a 2^24-ray sample of the 64 spp x 4K set per launch).
CUDA events on the library's stream, L2 flushed between frames, prints one JSON object.
usage: python tools/bench_configs.py [--reps N]"""
import argparse
import json
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
ap.add_argument("--reps", type=int, default=10)
args = ap.parse_args()
ASSETS = os.path.join(ROOT, "oracle", "_ref", "assets")
standin = gen_standin_obj.ensure(os.path.join(os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes"), "standin_272x256.obj"))


def scene(path, rescale, threads=4):
    m = A.load_model_data(path, threads)
    A.place_default(m, rescale)
    sc = A.Scene(m)
    A.prep_scene(sc)
    return sc


def time_frames(ctx, cam, opts, W, H):
    planes = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
    flush = ctx.dev_alloc(256 << 20)
    ms = []
    for i in range(args.reps + 3):
        ctx.dev_memset(flush, 0, 256 << 20)
        ctx.render_device(cam, opts, *planes)
        ctx.sync()
        if i >= 3:
            ms.append(ctx.last_kernel_ms())
    for p in planes + [flush]:
        ctx.dev_free(p)
    return float(np.mean(ms)), float(np.min(ms))


out = {}
cases = [("C1 Simple 640x360 1spp", os.path.join(ASSETS, "Simple.obj"), 40.0, 640, 360, 1, 0, 1),
         ("C2 Monkey 1920x1080 1spp", os.path.join(ASSETS, "Monkey.obj"), 40.0, 1920, 1080, 1, 0, 4),
         ("C3 Deer 1920x1080 16spp jittered", os.path.join(ASSETS, "Deer.obj"), 40.0, 1920, 1080, 16, 1, 4),
         ("C4 stand-in 3840x2160 1spp", standin, 0.0, 3840, 2160, 1, 0, 8)]
for name, path, rescale, W, H, spp, aa, thr in cases:
    if not os.path.exists(path):
        out[name] = "mesh not staged"
        continue
    sc = scene(path, rescale, thr)
    cam = A.set_camera(W, H, anti_aliasing=aa, samples_per_pixel=spp)
    mean, best = time_frames(sc.ctx, cam, sc.ctx.opts(), W, H)
    out[name] = {"rays": W * H * spp, "ms_per_frame": mean, "ms_min": best, "mrays_per_s": W * H * spp / mean / 1e3}
    if name.startswith("C4"):
        ctx = sc.ctx
        n = 1 << 24
        rays = random_rays(sc.model.dump()["aabb"], n, 1234)
        d_r, d_p, d_t = ctx.dev_alloc(n * 24), ctx.dev_alloc(n * 4), ctx.dev_alloc(n * 4)
        ctx.from_host(d_r, rays)
        ms = []
        for i in range(max(3, args.reps // 2) + 2):
            ctx.trace_rays_device(d_r, n, d_p, d_t, count_work=(i == 0))
            ctx.sync()
            if i == 0:
                ctr = ctx.counters()
            if i >= 2:
                ms.append(ctx.last_kernel_ms())
        hit = float((ctx.to_host(d_p, np.zeros(n, np.uint32)) != A.MISS).mean())
        out["C5 stand-in incoherent rays (2^24-ray sample of the 64spp x 4K set)"] = {
            "rays": n, "ms_per_launch": float(np.mean(ms)), "mrays_per_s": n / float(np.mean(ms)) / 1e3, "hit_fraction": hit,
            "box_tests_per_ray": ctr["box_tests"] / n, "tri_tests_per_ray": ctr["tri_tests"] / n,
            "full_c5_seconds_on_8_gpus_estimate": 530841600 / 8 / (n / float(np.mean(ms)) * 1e3)}
print(json.dumps(out, indent=1))
