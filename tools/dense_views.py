#!/usr/bin/env python3
"""Dense-view performance points: the C4 mesh at 3840x2160 from cameras where the model fills the
frame, next to the headline camera (app.cpp:88, 6 % hit pixels). For each view: device-resident frame time (CUDA events
on the library's stream, L2 flushed between frames), the render kernel alone, work counters of one counted frame, the hit
fraction — and, with --cpu, the reference's own CPU renderer (oracle/_ref) timed on the same camera.
usage: python tools/dense_views.py [--reps N] [--cpu] [--views a,b]   → one JSON object on stdout"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import atray_b200 as A  # noqa: E402
from tools import gen_standin_obj  # noqa: E402

VIEWS = {
    "headline (app.cpp:88)": dict(),
    "close-up, 76 % hit pixels": dict(eye=(10.0, -12.0, -20.0), facing=(-0.3, -0.1, -1.0), h_fov=0.5),
    "eye inside the model box, 89 % hit pixels": dict(eye=(0.0, -15.0, -38.0), facing=(0.3, 0.1, -1.0), h_fov=1.3),
    "wide angle from between the coils, 93 % hit pixels": dict(eye=(2.0, -14.0, -36.5), facing=(-1.0, -0.3, -0.7), h_fov=2.5),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--views", default="")
    args = ap.parse_args()
    W, H = 3840, 2160
    path = gen_standin_obj.ensure(os.path.join(os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes"), "standin_272x256.obj"))
    m = A.load_model_data(path, 8)
    A.place_default(m, 0.0)
    sc = A.Scene(m)
    A.prep_scene(sc)
    ctx = sc.ctx
    planes = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
    flush = ctx.dev_alloc(256 << 20)
    cpu = None
    if args.cpu:
        from oracle import bind as B
        cpu = B.Scene(B.ref(), path, loader_threads=8, rescale=0.0) if B.ref_available() else None
    threads = len(os.sched_getaffinity(0))
    out = {}
    for name, kw in VIEWS.items():
        if args.views and not any(v and v in name for v in args.views.split(",")):
            continue
        cam = A.set_camera(W, H, **kw)
        ms, rms = [], []
        for i in range(args.reps + 3):
            ctx.dev_memset(flush, 0, 256 << 20)
            ctx.render_device(cam, ctx.opts(), *planes)
            ctx.sync()
            if i >= 3:
                ms.append(ctx.last_kernel_ms())
                rms.append(ctx.last_render_kernel_ms())
        ctx.render_device(cam, ctx.opts(count_work=True), *planes)
        ctx.sync()
        ctr = ctx.counters()
        prim = ctx.to_host(planes[2], np.zeros((H, W), np.uint32))
        r = {"camera": kw, "ms_per_frame": float(np.mean(ms)), "ms_min": float(np.min(ms)),
             "render_kernel_ms": float(np.mean(rms)), "mrays_per_s": W * H / float(np.mean(ms)) / 1e3,
             "hit_fraction": float((prim != A.MISS).mean()), "counters": ctr}
        if cpu is not None:
            okw = dict(eye=kw.get("eye", A.api.DEFAULT_EYE), facing=kw.get("facing", A.api.DEFAULT_FACING), hfov=kw.get("h_fov", 1.0))
            img, sec, _ = cpu.render(W, H, threads=threads, **okw)
            bg = ctx.to_host(planes[0], np.zeros((H, W), np.uint32))
            r["cpu_reference"] = {"seconds_per_frame": sec, "mrays_per_s": W * H / sec / 1e6, "threads": threads,
                                  "same_pixels_as_gpu": bool((img == bg).all())}
        out[name] = r
        print(name, "%.3f ms" % r["ms_per_frame"], file=sys.stderr, flush=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
