"""CPU: the no-tree branch of get_intersection_data (renderer.cpp:58-82, the reference compiled without USE_KD_TREE) in
the C oracle against golden vectors made by that reference build (tests/golden/notree_*.npz, generator
make_golden_notree.py) and, where oracle/_ref exists, against it live. Domain property: the brute-force distance is the
true minimum over all faces, so it never exceeds what the tree walk returns for the same ray."""
import os

import numpy as np
import pytest

from helpers import GOLDEN, PATH_PLANES, PATH_SPHERES, PH, PW, SEED, bits, file_sha256, random_rays, same_bits

NAMES = ["knot_small", "Cube", "Deer", "Monkey"]


@pytest.mark.parametrize("name", NAMES)
def test_oracle_notree_matches_reference_golden(name, port, meshes):
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    g = np.load(os.path.join(GOLDEN, "notree_%s.npz" % name))
    path, rescale = meshes[name]
    assert file_sha256(path) == str(g["obj_sha256"])
    s = B.Scene(port, path, rescale=rescale)
    tree_face, tree_t, _ = s.trace_primary(PW, PH, threads=4)
    s.use_kd_tree(False)
    face, t, img = s.trace_primary(PW, PH, threads=4)
    assert (face == g["face"]).all() and (bits(t) == g["t_bits"]).all() and (img == g["bgrx"]).all()
    both = (face != B.MISS) & (tree_face != B.MISS)
    assert (t[both] <= tree_t[both]).all() and ((tree_face != B.MISS) <= (face != B.MISS)).all()
    jit, _ = s.render_jitter(PW, PH, 4, SEED, threads=4)
    assert (jit == g["jitter_bgrx"]).all()
    rf, rt = s.trace_rays(random_rays(s.model()["aabb"], 4000, int(g["ray_seed"])), threads=4)
    assert (rf == g["ray_face"]).all() and (bits(rt) == g["ray_t_bits"]).all()
    for c, r, m in PATH_SPHERES:
        s.add_sphere(c, r, m)
    for n, d, m in PATH_PLANES:
        s.add_plane(n, d, m)
    pimg, pt, pprim, pkind, casts = s.render_path(PW, PH, 2, 4, SEED + 1, aa=1, threads=4)
    assert (pimg == g["path_bgrx"]).all() and (bits(pt) == g["path_t_bits"]).all() and (pprim == g["path_prim"]).all()
    assert (pkind == g["path_kind"]).all() and casts == int(g["path_casts"])
    s.close()


def test_port_notree_equals_reference_live(port, ref, meshes):
    from oracle import bind as B
    if not B.ref_notree_available():
        pytest.skip("oracle/_ref/libatray_ref_notree.so not built")
    path, rescale = meshes["knot_small"]
    a, b = B.Scene(B.ref_notree(), path, rescale=rescale), B.Scene(port, path, rescale=rescale)
    b.use_kd_tree(False)
    eye = (3.0, -8.0, -20.0)
    ra, rb = a.trace_primary(120, 67, threads=4, eye=eye), b.trace_primary(120, 67, threads=4, eye=eye)
    assert (ra[0] == rb[0]).all() and same_bits(ra[1], rb[1]) and (ra[2] == rb[2]).all()
    a.close(), b.close()
