#!/usr/bin/env python3
"""Deterministic stand-in for the reference's missing Assets/Dragon.obj.

Frozen spec "standin-v1": (p,q)=(2,3) torus-knot tube, grid U x V (default 272 x 256 -> 139 264 faces /
417 792 unwelded vertices; README.md:19 of the reference says "139k faces, 416k verts"),
centre curve C(t) = ((cos 3t + 2) cos 2t, -sin 3t, (cos 3t + 2) sin 2t), tube radius
0.45 * (1 + 0.25 sin 5u cos 7v + 0.05 N(0,1)) with numpy.random.default_rng(1234).standard_normal((U, V)),
frame N = normalize(T x (0,1,0)), B = N x T, scaled so the largest extent is 47, two triangles per quad,
unwelded (3 "v" + 3 "vn" lines per face, "f a//a b//b c//c"), positions and normals written with %.6f.
Outward-facing winding (the reference's triangle test culls back faces, model.h:86).
"""
import argparse
import hashlib
import os

import numpy as np

SPEC = "standin-v1"


def standin_arrays(U=272, V=256, seed=1234):
    rng = np.random.default_rng(seed)
    noise = rng.standard_normal((U, V))
    u = np.arange(U, dtype=np.float64) * (2.0 * np.pi / U)
    v = np.arange(V, dtype=np.float64) * (2.0 * np.pi / V)
    c3, s3, c2, s2 = np.cos(3 * u), np.sin(3 * u), np.cos(2 * u), np.sin(2 * u)
    C = np.stack([(c3 + 2) * c2, -s3, (c3 + 2) * s2], axis=1)
    dC = np.stack([-3 * s3 * c2 - 2 * (c3 + 2) * s2, -3 * c3, -3 * s3 * s2 + 2 * (c3 + 2) * c2], axis=1)
    T = dC / np.linalg.norm(dC, axis=1, keepdims=True)
    N = np.cross(T, np.array([0.0, 1.0, 0.0]))
    N /= np.linalg.norm(N, axis=1, keepdims=True)
    B = np.cross(N, T)
    r = 0.45 * (1.0 + 0.25 * np.sin(5 * u)[:, None] * np.cos(7 * v)[None, :] + 0.05 * noise)
    radial = np.cos(v)[None, :, None] * N[:, None, :] + np.sin(v)[None, :, None] * B[:, None, :]
    P = C[:, None, :] + r[:, :, None] * radial
    ext = (P.reshape(-1, 3).max(0) - P.reshape(-1, 3).min(0)).max()
    P *= 47.0 / ext
    nrm = radial / np.linalg.norm(radial, axis=2, keepdims=True)
    i = np.arange(U)[:, None]
    j = np.arange(V)[None, :]
    i1, j1 = (i + 1) % U, (j + 1) % V
    idx = lambda a, b: (a * V + b + 0 * (i + j)).reshape(-1)
    q00, q10, q11, q01 = idx(i, j), idx(i1, j), idx(i1, j1), idx(i, j1)
    tri_a = np.stack([q00, q10, q11], axis=1)
    tri_b = np.stack([q00, q11, q01], axis=1)
    tris = np.empty((tri_a.shape[0] * 2, 3), dtype=np.int64)
    tris[0::2], tris[1::2] = tri_a, tri_b
    Pf, Nf = P.reshape(-1, 3), nrm.reshape(-1, 3)
    # make the winding outward: (b-a) x (c-a) must point along the radial normal
    a, b, c = Pf[tris[:, 0]], Pf[tris[:, 1]], Pf[tris[:, 2]]
    if (np.einsum("ij,ij->i", np.cross(b - a, c - a), Nf[tris[:, 0]]) < 0).mean() > 0.5:
        tris = tris[:, ::-1].copy()
    return Pf, Nf, tris


def write_obj(path, U=272, V=256, seed=1234):
    P, N, tris = standin_arrays(U, V, seed)
    pos = P[tris.reshape(-1)]
    nor = N[tris.reshape(-1)]
    nf = tris.shape[0]
    tmp = path + ".tmp%d" % os.getpid()
    with open(tmp, "w") as f:
        f.write("# %s U=%d V=%d seed=%d faces=%d\n" % (SPEC, U, V, seed, nf))
        CH = 30000
        for s in range(0, nf, CH):
            e = min(nf, s + CH)
            pv = pos[3 * s:3 * e]
            nv = nor[3 * s:3 * e]
            lines = []
            for k in range(e - s):
                p0, p1, p2 = pv[3 * k], pv[3 * k + 1], pv[3 * k + 2]
                n0, n1, n2 = nv[3 * k], nv[3 * k + 1], nv[3 * k + 2]
                base = 3 * (s + k) + 1
                lines.append(
                    "v %.6f %.6f %.6f\nv %.6f %.6f %.6f\nv %.6f %.6f %.6f\n"
                    "vn %.6f %.6f %.6f\nvn %.6f %.6f %.6f\nvn %.6f %.6f %.6f\n"
                    "f %d//%d %d//%d %d//%d\n"
                    % (p0[0], p0[1], p0[2], p1[0], p1[1], p1[2], p2[0], p2[1], p2[2],
                       n0[0], n0[1], n0[2], n1[0], n1[1], n1[2], n2[0], n2[1], n2[2],
                       base, base, base + 1, base + 1, base + 2, base + 2))
            f.write("".join(lines))
    os.replace(tmp, path)
    return nf


def ensure(path, U=272, V=256, seed=1234):
    """Generate once; later calls reuse the file."""
    if not os.path.exists(path):
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        write_obj(path, U, V, seed)
    return path


def sha256(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--U", type=int, default=272)
    ap.add_argument("--V", type=int, default=256)
    ap.add_argument("--seed", type=int, default=1234)
    a = ap.parse_args()
    n = write_obj(a.out, a.U, a.V, a.seed)
    print(a.out, n, "faces", sha256(a.out))
