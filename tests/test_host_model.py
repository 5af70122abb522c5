"""CPU: the product's host side (loader, transforms, tree builder, flattener, camera — atray_b200/csrc/host_model.cpp
through the C ABI) against the oracle and the reference-made golden vectors. No GPU compute is called here."""
import os

import numpy as np
import pytest

from helpers import bits, golden_cases, load_golden, mesh_for_case, same_bits, tree_digest, write_obj, cube_mesh


def build(path, rescale, method=1, cap=300, threads=3):
    import atray_b200 as A
    m = A.load_model_data(path, threads)
    A.place_default(m, rescale)
    A.build_KD_tree(m, method, cap)
    return m


@pytest.mark.parametrize("case", golden_cases())
def test_product_tree_matches_reference_golden(case, built, meshes):
    g = load_golden(case)
    mesh = mesh_for_case(case, meshes)
    if mesh is None:
        pytest.skip("mesh not on this machine")
    m = build(mesh[0], mesh[1], int(g["method"]), int(g["cap"]))
    c = m.counts()
    assert [c[k] for k in ("vertices", "normals", "tex_coords", "faces", "nodes", "refs", "inner", "max_leaf")] == \
        [int(x) for x in g["counts"]]
    assert tree_digest(m.tree()) == str(g["tree_sha256"])


@pytest.mark.parametrize("name", ["knot_small", "knot_mid", "Simple", "Cube", "Deer", "Monkey"])
@pytest.mark.parametrize("method,cap", [(1, 300), (0, 64), (1, 32)])
def test_product_host_equals_oracle(name, method, cap, port, meshes):
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    o = B.Scene(port, path, rescale=rescale, method=method, cap=cap)
    for threads in (1, 5):
        m = build(path, rescale, method, cap, threads)
        md, od = m.dump(), o.model()
        for k in od:
            assert same_bits(md[k], od[k]), k
        mt, ot = m.tree(), o.tree()
        for k in ot:
            assert same_bits(mt[k], ot[k]), k
    o.close()


def test_flat_layout_is_one_to_one(port, meshes):
    """north_star: the flat node array can be compared one-to-one with the reference tree."""
    path, rescale = meshes["knot_mid"]
    m = build(path, rescale, 1, 64)
    tree = m.tree()
    nodes, tris = m.flatten()
    n = nodes.shape[0]
    assert n == tree["boxes"].shape[0]
    assert same_bits(nodes[:, 0, :3], tree["boxes"][:, :3]) and same_bits(nodes[:, 1, :3], tree["boxes"][:, 3:])
    w0 = nodes[:, 0, 3].view(np.int32)
    w1 = nodes[:, 1, 3].view(np.int32)
    inner = tree["children_start"] != 0
    assert (w0[inner] == tree["children_start"][inner]).all() and (w0[inner] > 0).all()
    assert (w0[~inner] < 0).all() and (w1[~inner] == tree["prim_count"][~inner].astype(np.int32)).all()
    first = ~w0[~inner]
    assert (first % 8 == 0).all()
    # packets of 8: plane 0 = {a, face}, plane 1 = {ab, position in the reference's leaf order}, plane 2 = {ac, 0}
    A_, AB, AC = tris[:, 0].reshape(-1, 4), tris[:, 1].reshape(-1, 4), tris[:, 2].reshape(-1, 4)
    for leaf_i, f0 in zip(np.nonzero(~inner)[0], first):
        cnt, p0 = int(tree["prim_count"][leaf_i]), int(tree["prim_first"][leaf_i])
        tri = tree["tri"][p0:p0 + cnt]
        pos = AB[f0:f0 + cnt, 3].view(np.uint32)
        assert sorted(pos.tolist()) == list(range(cnt))          # a permutation of the leaf (Morton order inside the leaf)
        tri = tri[pos]
        assert same_bits(A_[f0:f0 + cnt, :3], tri[:, 0:3])
        assert same_bits(AB[f0:f0 + cnt, :3], tri[:, 3:6] - tri[:, 0:3])
        assert same_bits(AC[f0:f0 + cnt, :3], tri[:, 6:9] - tri[:, 0:3])
        assert (A_[f0:f0 + cnt, 3].view(np.uint32) == tree["face"][p0:p0 + cnt][pos]).all()
        pad = (cnt + 7) // 8 * 8
        assert (A_[f0 + cnt:f0 + pad, 3].view(np.uint32) == 0xFFFFFFFF).all()
        assert (AB[f0 + cnt:f0 + pad, :3] == 0).all() and (AC[f0 + cnt:f0 + pad] == 0).all()


def test_camera_matches_reference_kat(built):
    import atray_b200 as A
    from atray_b200 import api
    import helpers
    kat = np.load(os.path.join(helpers.GOLDEN, "kat.npz"))
    assert (bits(api.camera_as_array(A.set_camera(1280, 720))) == kat["camera_1280x720"]).all()
    assert (bits(api.camera_as_array(A.set_camera(3840, 2160))) == kat["camera_3840x2160"]).all()
    cam = A.set_camera(640, 360, eye=(1.5, -2.0, 3.0), facing=(0.3, 0.2, -1.0), h_fov=0.8)
    assert (bits(api.camera_as_array(cam)) == kat["camera_custom"]).all()
    # update_camera_pos (camera.h:23-34): moving a default camera to the custom pose gives the custom camera's basis, settings kept
    moved = A.update_camera_pos(A.set_camera(640, 360, h_fov=0.8), (1.5, -2.0, 3.0), (0.3, 0.2, -1.0))
    assert (bits(api.camera_as_array(moved)) == kat["camera_custom"]).all() and (moved.width, moved.height) == (640, 360)


def test_loader_edge_cases(built, port, tmp_path):
    """Ragged input the reference's loader accepts: quads (first 3 corners kept), negative indices, v/vt, v//vn,
    v/vt/vn, missing trailing newline, exponents, blank and comment lines, leading-space lines (ignored)."""
    import atray_b200 as A
    from oracle import bind as B
    text = ("# comment\n\nv 0 0 0\nv 1.5e0 0 0\nv 0 +2.25 0\nv 0 0 -3.125E-1\n  v 9 9 9\nvn 0 0 1\nvt 0.5 0.5\n"
            "f 1 2 3 4\nf -1 -2 -3\nf 1/1 2/1 3/1\nf 1//1 2//1 4//1\nf 1/1/1 3/1/1 4/1/1\nusemtl x\nf 2 3 4")
    p = tmp_path / "ragged.obj"
    p.write_text(text)
    o = B.Scene(port, str(p), rescale=0.0, cap=300)
    for threads in (1, 2, 7):
        m = A.load_model_data(str(p), threads)
        A.place_default(m, 0.0)
        A.build_KD_tree(m, A.DIV_SAH, 300)
        md, od = m.dump(), o.model()
        for k in od:
            assert same_bits(md[k], od[k]), (k, threads)
        assert m.counts()["faces"] == 6 and m.counts()["vertices"] == 4
        mt, ot = m.tree(), o.tree()
        for k in ot:
            assert same_bits(mt[k], ot[k]), k


def test_empty_and_root_leaf_models(built, port, tmp_path):
    import atray_b200 as A
    from oracle import bind as B
    e = tmp_path / "empty.obj"
    e.write_text(" n ")  # the reference's Assets/Simple.obj placeholder
    m = A.load_model_data(str(e), 4)
    A.place_default(m, 40.0)
    A.build_KD_tree(m)
    assert m.counts()["nodes"] == 1 and m.counts()["refs"] == 0
    v, f = cube_mesh()
    c = write_obj(str(tmp_path / "cube.obj"), v, f)
    m = build(c, 40.0)
    o = B.Scene(port, c, rescale=40.0)
    assert m.counts()["nodes"] == 1 and m.counts()["refs"] == 12
    for k, val in o.tree().items():
        assert same_bits(m.tree()[k], val)
    m2 = A.model_from_arrays(v, f)
    A.place_default(m2, 40.0)
    A.build_KD_tree(m2)
    assert same_bits(m2.tree()["tri"], m.tree()["tri"])


def test_error_behaviour(built, tmp_path):
    import atray_b200 as A
    from atray_b200 import api
    with pytest.raises(A.AtrayError):
        A.load_model_data(str(tmp_path / "nope.obj"))
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nf 1 2 3\n")  # face references vertices that do not exist (reference: out-of-bounds read)
    with pytest.raises(A.AtrayError):
        A.load_model_data(str(bad))
    v, f = cube_mesh()
    m = A.model_from_arrays(v, f)
    with pytest.raises(A.AtrayError):
        A.resize_scale(m, 2.0)  # needs get_AABB first
    with pytest.raises(A.AtrayError):
        m.flatten()  # needs build_KD_tree first
    with pytest.raises(A.AtrayError):
        A.build_KD_tree(m, 7, 300)
    with pytest.raises(A.AtrayError):
        A.set_camera(0, 10)
    with pytest.raises(A.AtrayError):
        A.model_from_arrays(v, np.array([[0, 1, 99]], np.int32))


def test_degenerate_split_guard(built):
    """The reference builder never terminates when every primitive lands in the same child again;
    the product returns ATRAY_ELIMIT instead."""
    import atray_b200 as A
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    f = np.tile(np.array([[0, 1, 2]], np.int32), (40, 1))  # 40 identical triangles, cap 4, CENTER split
    m = A.model_from_arrays(v, f)
    A.place_default(m, 0.0)
    with pytest.raises(A.AtrayError):
        A.build_KD_tree(m, A.DIV_CENTER, 4)


@pytest.mark.parametrize("method,cap", [(1, 300), (0, 32), (1, 64)])
def test_parallel_tree_build_equals_the_sequential_one(method, cap, standin, monkeypatch):
    """SURVEY 8 f4 (first half): the subtrees below the root built by several threads and spliced in the reference's node order —
    every box, child index and leaf primitive list identical to the one-thread build (which the golden tests pin to the reference)."""
    import atray_b200 as A
    from helpers import tree_digest
    digests = []
    for threads in ("1", "8", "3"):
        monkeypatch.setenv("ATRAY_BUILD_THREADS", threads)
        m = A.load_model_data(standin, 4)
        A.place_default(m, 0.0)
        A.build_KD_tree(m, method, cap)
        t = m.tree()
        digests.append((tree_digest(t), len(t["children_start"])))
    assert digests[0] == digests[1] == digests[2] and digests[0][1] > 9
