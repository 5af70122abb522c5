"""CPU: the oracle's C5 stream (oracle/atray_oracle.c: or_trace_c5 — incoherent rays per pixel index from the per-ray PCG32
stream) is self-consistent: the rays it reports, traced through the ray-list entry point, give the same ids and
distances; pixel means follow from the per-ray hits with the reference's sequential f32 sum, clamp and truncation; the
restatement agrees with the reference build on those rays."""
import numpy as np

from helpers import SEED, same_bits


def pixel_means(hit, spp, sky=(0.3, 0.4, 0.5), mdl=(0.4, 0.2, 0.2)):
    """renderer.cpp:353-365 in float32: sequential sum of the samples' emissions, / spp, clamp, (uint8)(c*255), BGRX."""
    f = np.float32
    hit = hit.reshape(-1, spp)
    out = np.zeros(hit.shape[0], np.uint32)
    ch = []
    for k in range(3):
        acc = np.zeros(hit.shape[0], np.float32)
        for s in range(spp):
            acc = (acc + np.where(hit[:, s], f(mdl[k]), f(sky[k]))).astype(np.float32)
        c = (acc / f(spp)).astype(np.float32)
        c = np.maximum(f(0), np.minimum(c, f(1)))
        ch.append(((c * f(255)).astype(np.float32)).astype(np.int32).astype(np.uint32) & 255)
    return ch[2] | (ch[1] << 8) | (ch[0] << 16)


def test_c5_stream_is_consistent(port, meshes):
    from oracle import bind as B
    path, rescale = meshes["knot_mid"]
    o = B.Scene(port, path, rescale=rescale)
    p0, p1, spp = 12345, 12345 + 700, 64
    face, t, bgrx, rays = o.trace_c5(p0, p1, spp, SEED, threads=4, want_rays=True)
    assert np.allclose(np.linalg.norm(rays[:, 3:], axis=1), 1.0, atol=1e-6)
    f2, t2 = o.trace_rays(rays, threads=4)
    assert (f2 == face).all() and same_bits(t2, t)
    assert (bgrx == pixel_means(face != B.MISS, spp)).all()
    assert 0.2 < float((face != B.MISS).mean()) < 0.98
    # a pixel's rays do not depend on the range it is traced in; another seed gives other rays
    fa, ta, ba = o.trace_c5(p0 + 100, p0 + 200, spp, SEED, threads=2)
    assert (fa == face[100 * spp:200 * spp]).all() and (ba == bgrx[100:200]).all()
    fb, _, _ = o.trace_c5(p0, p1, spp, SEED + 1, threads=4)
    assert not (fb == face).all()
    if B.ref_available():   # the reference's own get_ray_kd_tree_intersection on the same rays
        r = B.Scene(B.ref(), path, rescale=rescale)
        f3, t3 = r.trace_rays(rays, threads=4)
        assert (f3 == face).all() and same_bits(t3, t)
        r.close()
    o.close()
