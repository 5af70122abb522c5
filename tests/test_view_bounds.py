"""CPU: soundness of the per-frame film-space culling (atray_b200/csrc/view_bounds.h) against the oracle.

For every pixel of a frame the oracle lists every (leaf, triangle) its float triangle test ACCEPTS in any listed leaf
(a superset of what the reference's leaf scan can accept). The product's host evaluation of the film rectangles
(atray_debug_film_rects, the same code the CUDA pre-pass runs) must contain that pixel's film point in the triangle's
rectangle, and the pixel's 8x4 tile must be stamped busy. The test also checks that the bounds actually cull."""
import os

import numpy as np
import pytest

NT = min(16, os.cpu_count() or 1)

CAMERAS = [
    dict(),                                                                   # app.cpp:88
    dict(eye=(3.0, -4.0, 9.0), facing=(-0.2, -0.4, -1.0), h_fov=0.7),
    dict(eye=(0.0, -15.0, -38.0), facing=(0.3, 0.1, -1.0), h_fov=1.3),        # eye at the model centre
    dict(eye=(40.0, 5.0, -38.0), facing=(-1.0, -0.4, 0.02), h_fov=0.5),       # from the side
    dict(eye=(0.1, 2.0, 0.0), facing=(0.1, 0.5, 1.0), h_fov=1.0),             # model behind the camera
    dict(eye=(1000.0, 500.0, 2000.0), facing=(-1000.0, -515.0, -2038.0), h_fov=0.02),  # far away, narrow
]


def film_coords(cam, W, H):
    """renderer.cpp:317,329 in float32, as the kernel and the oracle evaluate them."""
    f = np.float32
    y = np.arange(H, dtype=np.float32)
    x = np.arange(W, dtype=np.float32)
    fy = f(-1.0) + f(2.0) * (y / f(H))
    fx = ((f(-1.0) + f(2.0) * (x / f(W))) * f(cam.h_fov)) * f(cam.aspect_ratio)
    return fx.astype(np.float32), fy.astype(np.float32)


def check(A, B, port, path, rescale, method, cap, W, H, camkw):
    m = A.load_model_data(path, 2)
    A.place_default(m, rescale)
    A.build_KD_tree(m, method, cap)
    o = B.Scene(port, path, rescale=rescale, method=method, cap=cap)
    cam = A.set_camera(W, H, **camkw)
    rects, leaf, pos, busy, hulls, pxr = m.film_rects(cam, hulls=True, px=True)
    okw = dict(eye=camkw.get("eye", A.api.DEFAULT_EYE), facing=camkw.get("facing", A.api.DEFAULT_FACING),
               hfov=camkw.get("h_fov", 1.0))
    px, node, ppos = o.accepted_hits(W, H, threads=NT, **okw)
    o.close()
    # slot of (leaf node, position in the reference's leaf order)
    real = pos != 0xFFFFFFFF
    key = leaf[real].astype(np.uint64) << np.uint64(32) | pos[real].astype(np.uint64)
    slots = np.nonzero(real)[0]
    order = np.argsort(key)
    key, slots = key[order], slots[order]
    q = node.astype(np.uint64) << np.uint64(32) | ppos.astype(np.uint64)
    at = np.searchsorted(key, q)
    assert (key[at] == q).all(), "flat form lost a leaf record"
    s = slots[at]
    fx, fy = film_coords(cam, W, H)
    x, y = px % W, px // W
    r = rects[s]
    inside = (fx[x] >= r[:, 0]) & (fx[x] <= r[:, 2]) & (fy[y] >= r[:, 1]) & (fy[y] <= r[:, 3])
    assert inside.all(), "%d accepted hits outside their film rectangle" % int((~inside).sum())
    assert busy[y // 4, x // 8].all(), "accepted hit in a tile stamped as sky"
    q = pxr[s]      # the pixel range the kernels cull with (film_rect_px) holds the pixel of every accepted hit
    inpx = (x >= q[:, 0]) & (x <= q[:, 1]) & (y >= q[:, 2]) & (y <= q[:, 3])
    assert inpx.all(), "%d accepted hits outside their pixel range" % int((~inpx).sum())
    # the three half-planes of the slot, evaluated in float32 the way the kernel does (both association orders)
    hl = hulls[s]
    for k in range(3):
        a, b, c = hl[:, k, 0], hl[:, k, 1], hl[:, k, 2]
        e = (a * fx[x]).astype(np.float32) + (b * fy[y]).astype(np.float32)
        assert (e.astype(np.float32) <= c).all(), "%d accepted hits outside half-plane %d" % (int((e > c).sum()), k)
        e64 = a.astype(np.float64) * fx[x] + b.astype(np.float64) * fy[y]      # what an FMA contraction rounds from
        assert (e64.astype(np.float32) <= c).all()
    return px.size, rects, busy, hulls, s


@pytest.mark.parametrize("camkw", CAMERAS)
@pytest.mark.parametrize("name,method,cap", [("knot_mid", 1, 300), ("knot_small", 0, 64), ("Monkey", 1, 300), ("Deer", 1, 32),
                                             ("Cube", 1, 300)])
def test_accepted_hits_lie_in_their_film_rectangles(name, method, cap, camkw, built, port, meshes):
    import atray_b200 as A
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    check(A, B, port, path, rescale, method, cap, 480, 270, camkw)


def test_bounds_are_tight_on_the_default_view(built, port, meshes):
    """The rectangles must cull: on the default view most tiles are sky and the average rectangle is small."""
    import atray_b200 as A
    from oracle import bind as B
    path, rescale = meshes["knot_mid"]
    n, rects, busy, hulls, hit_slots = check(A, B, port, path, rescale, 1, 300, 960, 540, {})
    # the half-planes cull: inside its own rectangle a triangle's hull keeps well under the whole pixel count
    cam = A.set_camera(960, 540)
    fx, fy = film_coords(cam, 960, 540)
    kept = total = 0
    for sl in np.unique(hit_slots)[:400]:
        r, h = rects[sl], hulls[sl]
        xs, ys = fx[(fx >= r[0]) & (fx <= r[2])], fy[(fy >= r[1]) & (fy <= r[3])]
        gx, gy = np.meshgrid(xs, ys)
        ok = np.ones(gx.shape, bool)
        for k in range(3):
            ok &= (h[k, 0] * gx + h[k, 1] * gy) <= h[k, 2]
        kept += int(ok.sum()); total += ok.size
    assert kept < 0.85 * total, (kept, total)   # (sample of 400 slots in slot order: depends on the slot layout)
    assert n > 10000
    assert busy.mean() < 0.5
    fin = np.isfinite(rects).all(axis=1)
    assert fin.sum() > 0.3 * rects.shape[0]            # front-facing triangles have finite rectangles
    empty = rects[:, 0] > rects[:, 2]
    assert empty.sum() > 0.3 * rects.shape[0]          # back-facing ones are removed for this eye
    w = (rects[fin, 2] - rects[fin, 0]) / (2 * 1.7778 / 960)
    assert np.median(w) < 40                           # pixels


def test_degenerate_camera_disables_culling(built, meshes):
    import atray_b200 as A
    path, rescale = meshes["knot_small"]
    m = A.load_model_data(path, 1)
    A.place_default(m, rescale)
    A.build_KD_tree(m)
    cam = A.set_camera(64, 36, facing=(0.0, 1.0, 0.0))     # facing ∥ up: camera_x = normalize(0) is not finite
    with pytest.raises(A.AtrayError):
        m.film_rects(cam)


@pytest.mark.parametrize("name", ["knot_sah300", "knot_center32", "monkey", "slivers"])
def test_random_cameras_keep_every_accepted_hit(name, built, port, meshes, tmp_path):
    """The fuzz cameras of tests/test_gpu_fuzz.py (a prefix of the same seeded sequence) on the CPU: every hit the oracle
    accepts lies in the triangle's film rectangle and in a tile stamped busy."""
    import atray_b200 as A
    from oracle import bind as B
    from fuzz_cams import fuzz_cameras, sliver_mesh
    from helpers import write_obj
    if name == "slivers":
        v, f = sliver_mesh()
        path, rescale, method, cap = write_obj(str(tmp_path / "slivers.obj"), v, f), 40.0, 1, 64
    else:
        mesh, method, cap = {"knot_sah300": ("knot_mid", 1, 300), "knot_center32": ("knot_mid", 0, 32),
                             "monkey": ("Monkey", 1, 300)}[name]
        if mesh not in meshes:
            pytest.skip("mesh not staged")
        path, rescale = meshes[mesh]
    m = A.load_model_data(path, 2)
    A.place_default(m, rescale)
    A.build_KD_tree(m, method, cap)
    tree = m.tree()
    cams = fuzz_cameras(120, 1000 + len(name), m.dump()["aabb"], tree["boxes"][tree["children_start"] == 0])
    checked = hits = 0
    for kw in cams:
        kw = dict(kw)
        W, H = kw.pop("width"), kw.pop("height")
        kw.pop("anti_aliasing"), kw.pop("samples_per_pixel")
        try:
            hits += check(A, B, port, path, rescale, method, cap, W, H, kw)[0]
        except A.AtrayError:
            continue        # camera the culling refuses (degenerate basis): the product renders it unculled
        checked += 1
    assert checked > 100 and hits > 1000
