#!/usr/bin/env python3
"""Generates tests/golden/*.npz by running the REFERENCE's own code (oracle/_ref/libatray_ref.so, built from
/root/reference/Source by oracle/ref/build_ref.sh) in this container. Re-run only here (needs /root/reference):
    python tests/golden/make_golden.py
Each fixture holds, for one mesh and the default camera (app.cpp:88) at 256x144: the tree summary and a digest of the
reference-built node array, per-pixel {face id, t bits, BGRX} from get_intersection_data, the image rendered by
start_render_from_camera, a 16-spp jittered image with the harness-defined per-pixel stream, and 20 000 random rays
traced through get_intersection_data. Plus known-answer vectors for PCG32 and set_camera."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import bind as B  # noqa: E402
from tools import gen_standin_obj  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
W, H = 256, 144
SEED = 0x853C49E6748FEA9B


def tree_digest(tree):
    h = hashlib.sha256()
    for k in ("boxes", "children_start", "prim_count", "prim_first", "tri", "face"):
        h.update(np.ascontiguousarray(tree[k]).tobytes())
    return h.hexdigest()


def random_rays(aabb, n, seed):
    """The C5 ray family of SURVEY §8(d): origins on a sphere around the model, targets inside its box."""
    rng = np.random.default_rng(seed)
    c, e = (aabb[:3] + aabb[3:]) / 2, (aabb[3:] - aabb[:3])
    v = rng.standard_normal((n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    org = c + 1.5 * np.linalg.norm(e) * v
    tgt = c + 0.5 * e * rng.uniform(-1, 1, (n, 3))
    d = tgt - org
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([org, d], axis=1).astype(np.float32)


def main():
    R = B.ref()
    d = "/tmp/atray_meshes"
    os.makedirs(d, exist_ok=True)
    cases = {
        "knot_small": (gen_standin_obj.ensure(os.path.join(d, "knot_48x32.obj"), 48, 32), 0.0),
        "Simple": (os.path.join(B.ASSETS, "Simple.obj"), 40.0),
        "Cube": (os.path.join(B.ASSETS, "Cube.obj"), 40.0),
        "Deer": (os.path.join(B.ASSETS, "Deer.obj"), 40.0),
        "Monkey": (os.path.join(B.ASSETS, "Monkey.obj"), 40.0),
    }
    for name, (path, rescale) in cases.items():
        for method, mname in ((1, "sah"), (0, "center")):
            if method == 0 and name not in ("Monkey", "knot_small"):
                continue
            s = B.Scene(R, path, loader_threads=1, rescale=rescale, method=method, cap=300 if method else 64)
            tree = s.tree()
            face, t, img, uv = s.trace_primary(W, H, want_uv=True)
            rendered, _, casts = s.render(W, H, threads=2)
            jit, hits = s.render_jitter(W // 2, H // 2, 16, SEED)
            rays = random_rays(s.model()["aabb"], 20000, 99)
            rf, rt = s.trace_rays(rays)
            np.savez_compressed(
                os.path.join(OUT, "%s_%s.npz" % (name, mname)),
                obj_sha256=gen_standin_obj.sha256(path), rescale=rescale, method=method, cap=300 if method else 64,
                counts=np.array([s.n_vertices, s.n_normals, s.n_tex, s.n_faces, s.n_nodes, s.n_refs, s.n_inner, s.max_leaf]),
                tree_sha256=tree_digest(tree), children_start=tree["children_start"], prim_count=tree["prim_count"],
                face=face, t_bits=t.view(np.uint32), bgrx=img, rendered=rendered, total_ray_casts=casts,
                jitter_bgrx=jit, jitter_hits=hits.astype(np.uint8), ray_seed=99, ray_face=rf, ray_t_bits=rt.view(np.uint32))
            print(name, mname, "nodes", s.n_nodes, "refs", s.n_refs, "hit frac", float((face != B.MISS).mean()))
            s.close()
    # multi-bounce path: the reference's own cast_ray / get_intersection_data with spheres and planes
    sys.path.insert(0, os.path.dirname(OUT))
    from helpers import PATH_CONFIGS, PATH_PLANES, PATH_SPHERES, PH, PW
    for name in ("knot_small", "Deer", "Monkey"):
        path, rescale = cases[name]
        out = {"obj_sha256": gen_standin_obj.sha256(path), "rescale": rescale}
        for prims in (0, 1):
            s = B.Scene(R, path, loader_threads=1, rescale=rescale)
            if prims:
                for c, r, m in PATH_SPHERES:
                    s.add_sphere(c, r, m)
                for n, dd, m in PATH_PLANES:
                    s.add_plane(n, dd, m)
            for k, (spp, bounce, aa) in enumerate(PATH_CONFIGS):
                img, t, prim, kind, casts = s.render_path(PW, PH, spp, bounce, SEED + k, aa=aa, threads=4)
                tag = "p%d_c%d_" % (prims, k)
                out.update({tag + "bgrx": img, tag + "t_bits": t.view(np.uint32), tag + "prim": prim,
                            tag + "kind": kind.astype(np.uint8), tag + "casts": casts})
            s.close()
        np.savez_compressed(os.path.join(OUT, "path_%s.npz" % name), **out)
        print("path", name)
    np.savez(os.path.join(OUT, "kat.npz"), pcg32_42_54=B.pcg32(R, 42, 54, 16),
             camera_1280x720=B.camera(R, 1280, 720).view(np.uint32), camera_3840x2160=B.camera(R, 3840, 2160).view(np.uint32),
             camera_custom=B.camera(R, 640, 360, eye=(1.5, -2.0, 3.0), facing=(0.3, 0.2, -1.0), hfov=0.8).view(np.uint32))


if __name__ == "__main__":
    main()
