"""CPU, only where oracle/_ref exists (this container): the C restatement against the reference's own code, live."""
import numpy as np
import pytest

from helpers import SEED, random_rays, same_bits


@pytest.mark.parametrize("name", ["Simple", "Cube", "Deer", "Monkey", "knot_small", "knot_mid"])
@pytest.mark.parametrize("method,cap", [(1, 300), (0, 64)])
def test_port_equals_reference(name, method, cap, port, ref, meshes):
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    a = B.Scene(ref, path, loader_threads=1 if name in ("Simple", "Cube") else 3, rescale=rescale, method=method, cap=cap)
    b = B.Scene(port, path, rescale=rescale, method=method, cap=cap)
    ma, mb = a.model(), b.model()
    for k in ma:
        assert same_bits(ma[k], mb[k]), k
    ta, tb = a.tree(), b.tree()
    for k in ta:
        assert same_bits(ta[k], tb[k]), k
    W, H = 480, 270
    fa, tA, ia, ua = a.trace_primary(W, H, threads=4, want_uv=True)
    fb, tB, ib, ub = b.trace_primary(W, H, threads=4, want_uv=True)
    assert (fa == fb).all() and same_bits(tA, tB) and (ia == ib).all() and same_bits(ua, ub)
    img, _, _ = a.render(W, H, threads=4)  # the reference's own tile thread pool
    assert (img == ib).all()
    ja, ha = a.render_jitter(160, 90, 8, SEED, threads=4)
    jb, hb = b.render_jitter(160, 90, 8, SEED, threads=4)
    assert (ja == jb).all() and (ha == hb).all()
    rays = random_rays(ma["aabb"], 50000, 5)
    ra, rta = a.trace_rays(rays, threads=4)
    rb, rtb = b.trace_rays(rays, threads=4)
    assert (ra == rb).all() and same_bits(rta, rtb)
    # TriangleIntersectionData's u, v of arbitrary rays (get_ray_kd_tree_intersection, kd_tree.cpp:302-313, 447-452)
    ua, uta, uva = a.trace_rays_uv(rays[:20000], threads=4)
    ub, utb, uvb = b.trace_rays_uv(rays[:20000], threads=4)
    assert (ua == ub).all() and same_bits(uta, utb) and same_bits(uva, uvb) and ((ua != 0xFFFFFFFF).any() or not (ra != 0xFFFFFFFF).any())
    a.close(); b.close()


def test_port_equals_reference_standin_1080p(port, ref, standin):
    from oracle import bind as B
    a = B.Scene(ref, standin, loader_threads=4, rescale=0.0)
    b = B.Scene(port, standin, rescale=0.0)
    assert (a.n_nodes, a.n_refs, a.n_inner, a.max_leaf) == (b.n_nodes, b.n_refs, b.n_inner, b.max_leaf)
    fa, tA, ia = a.trace_primary(1920, 1080, threads=8)
    fb, tB, ib = b.trace_primary(1920, 1080, threads=8)
    assert (fa == fb).all() and same_bits(tA, tB) and (ia == ib).all()
    a.close(); b.close()
