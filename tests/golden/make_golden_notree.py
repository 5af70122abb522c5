#!/usr/bin/env python3
"""Generates tests/golden/notree_*.npz with the REFERENCE compiled WITHOUT `#define USE_KD_TREE` (renderer.h:8 →
get_intersection_data's brute-force branch, renderer.cpp:58-82; oracle/_ref/libatray_ref_notree.so, built by
oracle/ref/build_ref.sh). Per mesh, default camera at 160x90: per-pixel {face, t bits, BGRX}, a 4-spp jittered image, a
2-spp x 4-bounce path image with the test spheres and plane, and 4 000 random rays.
Re-run only in the container that has /root/reference:  python tests/golden/make_golden_notree.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import bind as B  # noqa: E402
from tools import gen_standin_obj  # noqa: E402
from helpers import PATH_PLANES, PATH_SPHERES, PH, PW, SEED, random_rays  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    R = B.ref_notree()
    d = "/tmp/atray_meshes"
    os.makedirs(d, exist_ok=True)
    cases = {"knot_small": (gen_standin_obj.ensure(os.path.join(d, "knot_48x32.obj"), 48, 32), 0.0),
             "Cube": (os.path.join(B.ASSETS, "Cube.obj"), 40.0), "Deer": (os.path.join(B.ASSETS, "Deer.obj"), 40.0),
             "Monkey": (os.path.join(B.ASSETS, "Monkey.obj"), 40.0)}
    for name, (path, rescale) in cases.items():
        s = B.Scene(R, path, loader_threads=1, rescale=rescale)
        assert s.n_nodes == 0, "no-tree build must not build a tree"
        face, t, img = s.trace_primary(PW, PH, threads=8)
        jit, hits = s.render_jitter(PW, PH, 4, SEED, threads=8)
        rays = random_rays(s.model()["aabb"], 4000, 7)
        rf, rt = s.trace_rays(rays, threads=8)
        for c, r, m in PATH_SPHERES:
            s.add_sphere(c, r, m)
        for n, dd, m in PATH_PLANES:
            s.add_plane(n, dd, m)
        pimg, pt, pprim, pkind, casts = s.render_path(PW, PH, 2, 4, SEED + 1, aa=1, threads=8)
        np.savez_compressed(os.path.join(OUT, "notree_%s.npz" % name), obj_sha256=gen_standin_obj.sha256(path), rescale=rescale,
                            face=face, t_bits=t.view(np.uint32), bgrx=img, jitter_bgrx=jit, ray_seed=7, ray_face=rf,
                            ray_t_bits=rt.view(np.uint32), path_bgrx=pimg, path_t_bits=pt.view(np.uint32), path_prim=pprim,
                            path_kind=pkind.astype(np.uint8), path_casts=casts)
        print(name, "hit frac", float((face != B.MISS).mean()), "path casts", casts)
        s.close()


if __name__ == "__main__":
    main()
