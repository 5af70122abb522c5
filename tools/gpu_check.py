#!/usr/bin/env python3
"""First-contact GPU check (development aid): parity of the CUDA path against the CPU oracle on the stand-in at 4K,
kernel timings in both FP modes, work counters and the K0 microbenchmarks. Writes gpurun_out/gpu_check.json."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import atray_b200 as A  # noqa: E402
from oracle import bind as B  # noqa: E402
from tools import gen_standin_obj  # noqa: E402


def main():
    res = {}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    W, H = 3840, 2160
    path = gen_standin_obj.ensure("/tmp/atray_standin_272x256.obj")
    t0 = time.time()
    model = A.load_model_data(path, os.cpu_count() or 1)
    A.place_default(model, rescale=0.0)
    scene = A.Scene(model)
    A.prep_scene(scene)
    res["load_build_upload_s"] = time.time() - t0
    res["counts"] = model.counts()
    ctx = scene.ctx
    cam = A.set_camera(W, H)
    out = ctx.render(cam)
    res["strict_kernel_ms_first"] = ctx.last_kernel_ms()
    o = B.Scene(B.port(), path, rescale=0.0)
    t0 = time.time()
    face, t, img = o.trace_primary(W, H, threads=os.cpu_count() or 1)
    res["oracle_s"] = time.time() - t0
    res["oracle_threads"] = os.cpu_count()
    res["strict_prim_mismatch"] = int((out["prim"] != face).sum())
    res["strict_t_mismatch"] = int((out["t"].view(np.uint32) != t.view(np.uint32)).sum())
    res["strict_px_mismatch"] = int((out["bgrx"] != img).sum())
    res["hit_fraction"] = float((face != B.MISS).mean())
    fast = ctx.render(cam, ctx.opts(fp_mode=A.FP_FAST))
    res["fast_prim_mismatch"] = int((fast["prim"] != face).sum())
    both = (fast["prim"] == face) & (face != B.MISS)
    rel = np.abs(fast["t"][both].astype(np.float64) - t[both]) / t[both]
    res["fast_t_max_rel"] = float(rel.max()) if rel.size else 0.0
    res["fast_px_mismatch"] = int((fast["bgrx"] != img).sum())
    # device-resident timing
    d = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
    for mode, name in ((A.FP_STRICT, "strict"), (A.FP_FAST, "fast")):
        ms = []
        for i in range(8):
            ctx.render_device(cam, ctx.opts(fp_mode=mode), d[0], d[1], d[2])
            ctx.sync()
            ms.append(ctx.last_kernel_ms())
        res[name + "_kernel_ms"] = ms
        res[name + "_mrays_s"] = W * H / (min(ms[2:]) * 1e-3) / 1e6
    ctx.render_device(cam, ctx.opts(count_work=True), d[0], d[1], d[2])
    ctx.sync()
    res["counted_kernel_ms"] = ctx.last_kernel_ms()
    res["counters"] = ctx.counters()
    res["oracle_counters"] = o.primary_counters(W, H, threads=os.cpu_count() or 1)
    res["microbench"] = ctx.microbench()
    # ray-list entry on the same rays as a cross-check of K2
    rng = np.random.default_rng(7)
    n = 1 << 20
    box = model.dump()["aabb"]
    c, e = (box[:3] + box[3:]) / 2, (box[3:] - box[:3])
    org = c + 1.5 * np.linalg.norm(e) * (lambda v: v / np.linalg.norm(v, axis=1, keepdims=True))(rng.standard_normal((n, 3)))
    tgt = c + 0.5 * e * rng.uniform(-1, 1, (n, 3))
    dirs = tgt - org
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    rays = np.concatenate([org, dirs], axis=1).astype(np.float32)
    p2, t2 = ctx.trace_rays(rays)
    res["k2_kernel_ms_1M_incoherent"] = ctx.last_kernel_ms()
    f2, tt2 = o.trace_rays(rays, threads=os.cpu_count() or 1)
    res["k2_prim_mismatch"] = int((p2 != f2).sum())
    res["k2_t_mismatch"] = int((t2.view(np.uint32) != tt2.view(np.uint32)).sum())
    res["k2_hit_fraction"] = float((f2 != B.MISS).mean())
    print(json.dumps(res, indent=1))
    with open(os.path.join(ROOT, "gpurun_out", "gpu_check.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
