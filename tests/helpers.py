"""Shared test helpers (tests only)."""
import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GW, GH = 256, 144
SEED = 0x853C49E6748FEA9B


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def same_bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and bool((a.view(np.uint32) == b.view(np.uint32)).all())


def tree_digest(tree):
    h = hashlib.sha256()
    for k in ("boxes", "children_start", "prim_count", "prim_first", "tri", "face"):
        h.update(np.ascontiguousarray(tree[k]).tobytes())
    return h.hexdigest()


def random_rays(aabb, n, seed):
    """C5 ray family: origins on a sphere around the model box, targets inside it. Same code as
    tests/golden/make_golden.py."""
    rng = np.random.default_rng(seed)
    c, e = (aabb[:3] + aabb[3:]) / 2, (aabb[3:] - aabb[:3])
    v = rng.standard_normal((n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    org = c + 1.5 * np.linalg.norm(e) * v
    tgt = c + 0.5 * e * rng.uniform(-1, 1, (n, 3))
    d = tgt - org
    with np.errstate(invalid="ignore"):
        d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([org, d], axis=1).astype(np.float32)


def golden_cases():
    out = []
    for f in sorted(os.listdir(GOLDEN)):
        if f.endswith(".npz") and f != "kat.npz" and not f.startswith(("path_", "notree_")):
            out.append(f[:-4])
    return out


def load_golden(case):
    return np.load(os.path.join(GOLDEN, case + ".npz"))


def file_sha256(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def mesh_for_case(case, meshes):
    """Returns (path, rescale) for a golden case or None when the mesh is not on this machine."""
    name = case.rsplit("_", 1)[0]
    return meshes.get(name)


def write_obj(path, vertices, faces):
    with open(path, "w") as f:
        for v in vertices:
            f.write("v %.6f %.6f %.6f\n" % tuple(v))
        for a, b, c in faces:
            f.write("f %d %d %d\n" % (a + 1, b + 1, c + 1))
    return path


def cube_mesh(size=1.0):
    s = size
    v = [(-s, -s, -s), (s, -s, -s), (s, s, -s), (-s, s, -s), (-s, -s, s), (s, -s, s), (s, s, s), (-s, s, s)]
    q = [(0, 3, 2, 1), (4, 5, 6, 7), (0, 1, 5, 4), (2, 3, 7, 6), (1, 2, 6, 5), (0, 4, 7, 3)]  # outward CCW
    f = []
    for a, b, c, d in q:
        f += [(a, b, c), (a, c, d)]
    return np.array(v, np.float32), np.array(f, np.int32)


# Scene primitives for the multi-bounce path tests (shapes as app.cpp:111-129 builds them): (center, radius, material)
# and (normal, distance, material) with material = (emission rgb, reflection rgb, scatter).
PATH_SPHERES = [((-6.0, -12.0, -30.0), 3.0, ((0.0, 0.0, 0.0), (0.9, 0.9, 0.9), 0.9)),
                ((8.0, -9.0, -34.0), 4.0, ((0.6, 0.5, 0.1), (0.3, 0.8, 0.3), 0.1))]
PATH_PLANES = [((0.0, 1.0, 0.05), -20.0, ((0.05, 0.05, 0.05), (0.6, 0.6, 0.7), 0.5))]
# (spp, bounce_limit, anti_aliasing)
PATH_CONFIGS = [(1, 1, 0), (2, 5, 0), (3, 4, 1), (1, 8, 1)]
PW, PH = 160, 90


def bmp_test_bitmap(w=7, h=5):
    """Deterministic BGRX bitmap (alpha byte 0, as Set_Pixel writes it) for the BMP writer tests."""
    y, x = np.mgrid[0:h, 0:w].astype(np.uint32)
    return ((x * 37 + y * 11) & 255) | (((x * 5 + y * 91 + 3) & 255) << 8) | (((x * 17 + y * 29 + 200) & 255) << 16)
