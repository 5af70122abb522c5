"""CPU: the multi-bounce path of the C oracle (get_intersection_data in full: tree, spheres, planes, smooth/flat normals;
cast_ray at any bounce_limit — renderer.cpp:34-160, 213-262) against (a) the committed golden vectors made by the
reference's own code (tests/golden/path_*.npz, generator make_golden.py) and (b), where oracle/_ref exists, the
reference live. Random draws come from the harness-defined per-pixel stream (the reference seeds from rdtsc)."""
import os

import numpy as np
import pytest

from helpers import GOLDEN, PATH_CONFIGS, PATH_PLANES, PATH_SPHERES, PH, PW, SEED, bits, file_sha256, same_bits


def add_prims(s):
    for c, r, m in PATH_SPHERES:
        s.add_sphere(c, r, m)
    for n, d, m in PATH_PLANES:
        s.add_plane(n, d, m)


@pytest.mark.parametrize("name", ["knot_small", "Deer", "Monkey"])
def test_oracle_path_matches_reference_golden(name, port, meshes):
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    g = np.load(os.path.join(GOLDEN, "path_%s.npz" % name))
    path, rescale = meshes[name]
    assert file_sha256(path) == str(g["obj_sha256"])
    for prims in (0, 1):
        s = B.Scene(port, path, rescale=rescale)
        if prims:
            add_prims(s)
        for k, (spp, bounce, aa) in enumerate(PATH_CONFIGS):
            img, t, prim, kind, casts = s.render_path(PW, PH, spp, bounce, SEED + k, aa=aa, threads=4)
            tag = "p%d_c%d_" % (prims, k)
            assert (img == g[tag + "bgrx"]).all(), tag
            assert (bits(t) == g[tag + "t_bits"]).all() and (prim == g[tag + "prim"]).all() and (kind == g[tag + "kind"]).all()
            assert casts == int(g[tag + "casts"])
        s.close()


@pytest.mark.parametrize("name", ["Cube", "Deer", "Monkey", "knot_mid"])
def test_port_path_equals_reference_live(name, port, ref, meshes):
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    for prims in (0, 1):
        a, b = B.Scene(ref, path, rescale=rescale), B.Scene(port, path, rescale=rescale)
        if prims:
            add_prims(a), add_prims(b)
        for spp, bounce, aa, eye in ((2, 5, 0, B.DEFAULT_EYE), (2, 6, 1, (3.0, -8.0, -20.0))):
            ra = a.render_path(200, 112, spp, bounce, 77, aa=aa, threads=4, eye=eye)
            rb = b.render_path(200, 112, spp, bounce, 77, aa=aa, threads=4, eye=eye)
            assert (ra[0] == rb[0]).all() and same_bits(ra[1], rb[1]) and (ra[2] == rb[2]).all() and (ra[3] == rb[3]).all()
            assert ra[4] == rb[4]
        a.close(), b.close()


def test_bounce_one_without_primitives_is_the_flat_path(port, meshes):
    """bounce_limit = 1, no primitives: cast_ray returns the hit material's emission — the primary-ray path."""
    from oracle import bind as B
    path, rescale = meshes["knot_small"]
    s = B.Scene(port, path, rescale=rescale)
    face, t, img = s.trace_primary(PW, PH)
    img2, t2, prim2, kind2, casts = s.render_path(PW, PH, 1, 1, SEED)
    assert (img == img2).all() and same_bits(t, t2) and (face == prim2).all() and casts == int((face != B.MISS).sum())
    s.close()
