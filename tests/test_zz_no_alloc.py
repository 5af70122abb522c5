"""GPU: atray_trace_rays_static — the ray-list trace on the library's static device scratch (a call that allocates nothing,
include/atray_b200.h) — against the existing entry point atray_trace_rays_device, through the C ABI, bit for bit (the tolerance
of the STRICT parity tests), on two calls in a row that share the scratch. Collected last on purpose:
a failure here must not hide the parity tests."""
import numpy as np
import pytest

from helpers import random_rays, same_bits

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def A(built):
    import atray_b200 as A
    assert A.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    return A


def scene(A, path, rescale, method, cap):
    m = A.load_model_data(path, 4)
    A.place_default(m, rescale)
    sc = A.Scene(m, division_method=method, max_no_faces_per_node=cap)
    A.prep_scene(sc)
    return sc


def both_entries(ctx, rays, bufs, fp_mode):
    """prim / t of the same device-resident ray list from the existing entry and from the static-scratch entry."""
    n = rays.shape[0]
    d_rays, d_prim, d_t = bufs
    ctx.from_host(d_rays, rays)
    out = []
    for call in (ctx.trace_rays_device, ctx.trace_rays_static):
        ctx.dev_memset(d_prim, 0xEE, n * 4)
        ctx.dev_memset(d_t, 0xEE, n * 4)
        call(d_rays, n, d_prim, d_t, fp_mode=fp_mode)
        ctx.sync()
        out.append((ctx.to_host(d_prim, np.zeros(n, np.uint32)), ctx.to_host(d_t, np.zeros(n, np.float32))))
    return out


# SAH / 300: short leaf lists, the whole launch grid fits the static scratch; CENTER / 6: 30 k tiny leaves, rays spill into the
# per-warp overflow areas and only a few thread blocks fit (the entry shrinks its grid)
@pytest.mark.parametrize("method,cap,n1,n2", [(1, 300, 150000, 40001), (0, 6, 60000, 20011)])
def test_static_scratch_entry_matches_the_allocating_entry_twice_in_a_row(A, meshes, method, cap, n1, n2):
    path, rescale = meshes["knot_mid"]
    sc = scene(A, path, rescale, method, cap)
    ctx = sc.ctx
    box = sc.model.dump()["aabb"]
    cap_n = max(n1, n2)
    bufs = (ctx.dev_alloc(cap_n * 24), ctx.dev_alloc(cap_n * 4), ctx.dev_alloc(cap_n * 4))   # the caller's inputs and outputs
    try:
        hits = 0
        for k, n in enumerate((n1, n2)):                       # the second call finds the first call's scratch contents
            rays = random_rays(box, n, 11 + k)
            (p_ref, t_ref), (p_st, t_st) = both_entries(ctx, rays, bufs, A.FP_STRICT)
            assert (p_st == p_ref).all() and same_bits(t_st, t_ref), "call %d differs from atray_trace_rays_device" % (k + 1)
            hits += int((p_ref != A.MISS).sum())
        assert hits > 0, "the rays never hit the model"
        rays = random_rays(box, n2, 5)
        (p_ref, t_ref), (p_st, t_st) = both_entries(ctx, rays, bufs, A.FP_FAST)   # same kernel, same arithmetic: still bit-equal
        assert (p_st == p_ref).all() and same_bits(t_st, t_ref)
        with pytest.raises(A.api.AtrayError, match="too long"):
            ctx.trace_rays_static(bufs[0], (1 << 36) + 1, bufs[1], bufs[2])
        ctx.trace_rays_static(bufs[0], 0, bufs[1], bufs[2])    # empty list: nothing launched
        ctx.sync()
    finally:
        for b in bufs:
            ctx.dev_free(b)
