"""GPU: the CUDA path, called through the C ABI, against the CPU oracle and the reference-made golden vectors.
Bar (prompt ③ / north_star): STRICT mode bit-exact on primitive id, t and pixels; FAST (FMA) mode within
≤1e-5 id mismatches, t within 1e-4 relative, pixels within 1 LSB."""
import os

import numpy as np
import pytest

from helpers import (GH, GOLDEN, GW, PATH_CONFIGS, PATH_PLANES, PATH_SPHERES, PH, PW, SEED, bits, cube_mesh, golden_cases,
                     load_golden, mesh_for_case, random_rays, same_bits, write_obj)

pytestmark = pytest.mark.gpu

NTHREADS = min(16, os.cpu_count() or 1)


@pytest.fixture(scope="module")
def A(built):
    import atray_b200 as A
    assert A.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    return A


def gpu_scene(A, path, rescale, method=1, cap=300):
    m = A.load_model_data(path, 4)
    A.place_default(m, rescale)
    sc = A.Scene(m, division_method=method, max_no_faces_per_node=cap)
    A.prep_scene(sc)
    return sc


@pytest.mark.parametrize("case", golden_cases())
def test_cuda_matches_reference_golden(case, A, meshes):
    g = load_golden(case)
    mesh = mesh_for_case(case, meshes)
    if mesh is None:
        pytest.skip("mesh not on this machine")
    sc = gpu_scene(A, mesh[0], mesh[1], int(g["method"]), int(g["cap"]))
    out = sc.ctx.render(A.set_camera(GW, GH))
    assert (out["prim"] == g["face"]).all()
    assert (bits(out["t"]) == g["t_bits"]).all()
    assert (out["bgrx"] == g["bgrx"]).all() and (out["bgrx"] == g["rendered"]).all()
    jit = sc.ctx.render(A.set_camera(GW // 2, GH // 2, anti_aliasing=1, samples_per_pixel=16), sc.ctx.opts(seed=SEED))
    assert (jit["bgrx"] == g["jitter_bgrx"]).all()
    rays = random_rays(sc.model.dump()["aabb"], 20000, int(g["ray_seed"]))
    prim, t = sc.ctx.trace_rays(rays)
    assert (prim == g["ray_face"]).all() and (bits(t) == g["ray_t_bits"]).all()


@pytest.mark.parametrize("name", ["knot_small", "knot_mid", "Simple", "Cube", "Deer", "Monkey"])
@pytest.mark.parametrize("method,cap", [(1, 300), (0, 64), (1, 32)])
def test_cuda_equals_oracle(name, method, cap, A, port, meshes):
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    sc = gpu_scene(A, path, rescale, method, cap)
    o = B.Scene(port, path, rescale=rescale, method=method, cap=cap)
    for (W, H) in ((1920, 1080), (333, 187)):  # second size: not a multiple of the 8x4 tile
        out = sc.ctx.render(A.set_camera(W, H))
        face, t, img = o.trace_primary(W, H, threads=NTHREADS)
        assert (out["prim"] == face).all() and same_bits(out["t"], t) and (out["bgrx"] == img).all()
    # other camera
    eye, facing = (3.0, -4.0, 9.0), (-0.2, -0.4, -1.0)
    out = sc.ctx.render(A.set_camera(640, 360, eye=eye, facing=facing, h_fov=0.7))
    face, t, img = o.trace_primary(640, 360, threads=NTHREADS, eye=eye, facing=facing, hfov=0.7)
    assert (out["prim"] == face).all() and same_bits(out["t"], t) and (out["bgrx"] == img).all()
    o.close()


def test_c2_monkey_1080p_and_c3_deer_jitter(A, port, meshes):
    """BASELINE.json configs[1] and configs[2]."""
    from oracle import bind as B
    if "Monkey" not in meshes:
        pytest.skip("reference assets not staged")
    sc = gpu_scene(A, *meshes["Monkey"])
    o = B.Scene(port, *meshes["Monkey"][:1], rescale=40.0)
    out = sc.ctx.render(A.set_camera(1920, 1080))
    face, t, img = o.trace_primary(1920, 1080, threads=NTHREADS)
    assert (out["prim"] == face).all() and same_bits(out["t"], t) and (out["bgrx"] == img).all()
    assert abs(float((face != B.MISS).mean()) - 0.0556) < 1e-3  # SURVEY §6 probe
    o.close()
    sc = gpu_scene(A, *meshes["Deer"])
    o = B.Scene(port, meshes["Deer"][0], rescale=40.0)
    jit = sc.ctx.render(A.set_camera(1920, 1080, anti_aliasing=1, samples_per_pixel=16), sc.ctx.opts(seed=SEED), want=("bgrx",))
    ref_img, hits = o.render_jitter(1920, 1080, 16, SEED, threads=NTHREADS)
    assert (jit["bgrx"] == ref_img).all()
    assert len(np.unique(ref_img)) > 2  # edge pixels really are blends
    o.close()


@pytest.mark.parametrize("W,H", [(96, 720), (40, 1080), (1080, 64)])
def test_jittered_portrait_and_strip_frames_equal_oracle(W, H, A, port, meshes):
    """Anti-aliased frames far from 16:9: the jitter is half_pixel_width = 0.5 h_fov / W in film units while a pixel
    is 2 h_fov aspect / W wide, i.e. 0.25 / aspect pixels in x — 4.5 pixels at 40x1080; the pixel ranges' slack is derived from
    the camera, so no busy tile may be filled as sky. Pixels against the oracle's jitter loop, culled against unculled."""
    from oracle import bind as B
    path, rescale = meshes["knot_mid"]
    sc = gpu_scene(A, path, rescale)
    o = B.Scene(port, path, rescale=rescale)
    cam = A.set_camera(W, H, anti_aliasing=1, samples_per_pixel=8)
    a = sc.ctx.render(cam, sc.ctx.opts(seed=SEED))
    b = sc.ctx.render(cam, sc.ctx.opts(seed=SEED, cull=False))
    ref_img, hits = o.render_jitter(W, H, 8, SEED, threads=NTHREADS)
    assert (a["bgrx"] == ref_img).all() and (b["bgrx"] == ref_img).all()
    assert (a["prim"] == b["prim"]).all() and same_bits(a["t"], b["t"])
    assert (hits > 0).any() and len(np.unique(ref_img)) > 2
    o.close()


def test_c4_standin_4k_full_size(A, port, standin):
    """BASELINE.json configs[3] at full size: bit-exact in STRICT mode, tolerance in FAST mode, work counters equal."""
    from oracle import bind as B
    sc = gpu_scene(A, standin, 0.0)
    o = B.Scene(port, standin, rescale=0.0)
    W, H = 3840, 2160
    cam = A.set_camera(W, H)
    face, t, img = o.trace_primary(W, H, threads=NTHREADS)
    # unculled kernel: same work as the reference, counter for counter
    out = sc.ctx.render(cam, sc.ctx.opts(count_work=True, cull=False))
    ctr = sc.ctx.counters()
    assert (out["prim"] == face).all() and same_bits(out["t"], t) and (out["bgrx"] == img).all()
    oc = o.primary_counters(W, H, threads=NTHREADS)
    assert ctr["rays"] == W * H and ctr["box_tests"] == oc["box_tests"] and ctr["tri_tests"] == oc["tri_tests"]
    assert ctr["leaves_listed"] == oc["leaves_listed"] and ctr["leaves_scanned"] == oc["leaves_scanned"]
    # default (film-space culling): same bits, a fraction of the work
    out = sc.ctx.render(cam, sc.ctx.opts(count_work=True))
    cc = sc.ctx.counters()
    assert (out["prim"] == face).all() and same_bits(out["t"], t) and (out["bgrx"] == img).all()
    assert cc["rays"] == W * H and cc["tiles_sky"] + cc["tiles_busy"] == (W // 8) * (H // 4)
    assert cc["tiles_sky"] > 0.8 * (W // 8) * (H // 4) and cc["tri_tests"] < 0.02 * ctr["tri_tests"]
    assert cc["pass_v"] >= int((face != B.MISS).sum())
    ras = sc.ctx.render(cam, sc.ctx.opts(raster=True))   # film-space rasteriser (opt-in, raster.cuh)
    assert (ras["prim"] == face).all() and same_bits(ras["t"], t) and (ras["bgrx"] == img).all()
    fast = sc.ctx.render(cam, sc.ctx.opts(fp_mode=A.FP_FAST))
    mism = int((fast["prim"] != face).sum())
    assert mism <= 1e-5 * W * H, mism
    both = (fast["prim"] == face) & (face != B.MISS)
    rel = np.abs(fast["t"][both].astype(np.float64) - t[both]) / t[both]
    assert rel.max() <= 1e-4
    px = np.abs(((fast["bgrx"][..., None] >> np.array([0, 8, 16])) & 255).astype(int) -
                ((img[..., None] >> np.array([0, 8, 16])) & 255).astype(int))
    assert (px.max(axis=-1) > 1).sum() <= 1e-5 * W * H  # only flipped hit/miss pixels differ by more than 1 LSB
    # idempotence: same frame twice → same bits; partition union == whole frame
    again = sc.ctx.render(cam)
    assert (again["prim"] == out["prim"]).all() and same_bits(again["t"], out["t"])
    o.close()


CAMS = [dict(), dict(eye=(3.0, -4.0, 9.0), facing=(-0.2, -0.4, -1.0), h_fov=0.7),
        dict(eye=(0.0, -15.0, -38.0), facing=(0.3, 0.1, -1.0), h_fov=1.3),        # eye inside the model's box
        dict(eye=(40.0, 5.0, -38.0), facing=(-1.0, -0.4, 0.02), h_fov=0.5),
        dict(eye=(0.1, 2.0, 0.0), facing=(0.1, 0.5, 1.0)),                         # model behind the camera
        dict(eye=(1000.0, 500.0, 2000.0), facing=(-1000.0, -515.0, -2038.0), h_fov=0.02),
        dict(eye=(2.0, -14.0, -36.5), facing=(-1.0, -0.3, -0.7), h_fov=2.5)]       # wide angle from between the coils


@pytest.mark.parametrize("name,method,cap", [("knot_mid", 1, 300), ("knot_mid", 0, 32), ("Monkey", 1, 300), ("Deer", 1, 64), ("Cube", 1, 300)])
def test_culled_equals_unculled_and_oracle(name, method, cap, A, port, meshes):
    """Film-space culling (default) never changes a bit: against the unculled kernel and the oracle, for cameras outside,
    inside and facing away from the model, plain and jittered."""
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    sc = gpu_scene(A, path, rescale, method, cap)
    o = B.Scene(port, path, rescale=rescale, method=method, cap=cap)
    for i, kw in enumerate(CAMS):
        W, H = (801, 453) if i % 2 else (640, 360)
        cam = A.set_camera(W, H, **kw)
        a = sc.ctx.render(cam)
        b = sc.ctx.render(cam, sc.ctx.opts(cull=False))
        assert (a["prim"] == b["prim"]).all() and same_bits(a["t"], b["t"]) and (a["bgrx"] == b["bgrx"]).all()
        face, t, img = o.trace_primary(W, H, threads=NTHREADS, eye=kw.get("eye", A.api.DEFAULT_EYE),
                                       facing=kw.get("facing", A.api.DEFAULT_FACING), hfov=kw.get("h_fov", 1.0))
        assert (a["prim"] == face).all() and same_bits(a["t"], t) and (a["bgrx"] == img).all()
        camj = A.set_camera(W // 2, H // 2, anti_aliasing=1, samples_per_pixel=4, **kw)
        aj = sc.ctx.render(camj, sc.ctx.opts(seed=SEED + i))
        bj = sc.ctx.render(camj, sc.ctx.opts(seed=SEED + i, cull=False))
        assert (aj["bgrx"] == bj["bgrx"]).all() and (aj["prim"] == bj["prim"]).all() and same_bits(aj["t"], bj["t"])
    # camera the culling cannot use (facing ∥ up → non-finite basis): falls back to the unculled kernel, same as the oracle
    cam = A.set_camera(64, 36, facing=(0.0, 1.0, 0.0))
    a = sc.ctx.render(cam)
    face, t, img = o.trace_primary(64, 36, threads=2, facing=(0.0, 1.0, 0.0))
    assert (a["prim"] == face).all() and (a["bgrx"] == img).all()
    o.close()


@pytest.mark.parametrize("name,method,cap", [("knot_mid", 1, 300), ("knot_mid", 0, 32), ("Monkey", 1, 300), ("Deer", 1, 64)])
def test_rasterised_frame_equals_tile_kernel(name, method, cap, A, meshes, monkeypatch):
    """raster=True rasterises a frame without anti-aliasing triangle by triangle (raster.cuh) instead of traversing it
    tile by tile. Same bits for every camera, with a band partition, and when the hit list overflows (those pixels go
    through the exact per-ray walker inside k_resolve)."""
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    sc = gpu_scene(A, path, rescale, method, cap)
    frames = []
    for i, kw in enumerate(CAMS):
        W, H = (801, 453) if i % 2 else (640, 360)
        cam = A.set_camera(W, H, samples_per_pixel=1 + (i % 3), **kw)
        a = sc.ctx.render(cam, sc.ctx.opts(count_work=True, raster=True))
        ca = sc.ctx.counters()
        b = sc.ctx.render(cam)
        assert (a["prim"] == b["prim"]).all() and same_bits(a["t"], b["t"]) and (a["bgrx"] == b["bgrx"]).all()
        assert ca["rays"] == W * H * (1 + (i % 3))
        again = sc.ctx.render(cam, sc.ctx.opts(raster=True))   # the pixel keys were restored: the next frame starts clean
        assert (again["prim"] == a["prim"]).all() and same_bits(again["t"], a["t"])
        frames.append((cam, a))
    # band partition executed rank by rank on this GPU
    cam, a = frames[1]
    W, H = cam.width, cam.height
    ctx = sc.ctx
    d = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
    for world, band in ((2, 8), (3, 12)):
        for rank in range(world):
            ctx.render_device(cam, ctx.opts(rank=rank, world=world, band_rows=band, raster=True), d[0], d[1], d[2])
        ctx.sync()
        assert (ctx.to_host(d[0], np.zeros((H, W), np.uint32)) == a["bgrx"]).all()
        assert same_bits(ctx.to_host(d[1], np.zeros((H, W), np.float32)), a["t"])
        assert (ctx.to_host(d[2], np.zeros((H, W), np.uint32)) == a["prim"]).all()
    for p in d:
        ctx.dev_free(p)
    # tiny hit list: nearly every hit pixel is flagged and re-traced exactly
    monkeypatch.setenv("ATRAY_HIT_CAP", "64")
    sc2 = gpu_scene(A, path, rescale, method, cap)
    for cam, a in frames[:3]:
        c = sc2.ctx.render(cam, sc2.ctx.opts(count_work=True, raster=True))
        assert (c["prim"] == a["prim"]).all() and same_bits(c["t"], a["t"]) and (c["bgrx"] == a["bgrx"]).all()
    assert sc2.ctx.counters()["overflow_rays"] > 0


@pytest.mark.parametrize("cap", [12, 6])
def test_cut_rule_on_split_edges(cap, A, port, tmp_path):
    """The "at most 5 accepted children" cut (kd_tree.cpp:374) only triggers for rays through a split edge. Eight cubes
    mirrored around the origin with CENTER splits put the split planes at exactly 0 / ±2; an eye at (-3,-3,10) looking
    down -z makes every diagonal pixel's ray cross x = 0 and y = 0 at the same parameter (bit-equal slab values), inside
    the tree. The culled kernel decides per node whether the cut can trigger (pairwise different split parameters ⇒
    ≤ 4 children pass); these rays take its exact sibling evaluation. Culled = unculled = oracle, bit for bit."""
    from oracle import bind as B
    v0, f0 = cube_mesh(1.0)
    vs, fs = [], []
    for k, c in enumerate([(x, y, z) for x in (-2, 2) for y in (-2, 2) for z in (-2, 2)]):
        vs.append(v0 + np.array(c, np.float32))
        fs.append(f0 + 8 * k)
    path = write_obj(str(tmp_path / "cubes8.obj"), np.concatenate(vs), np.concatenate(fs))
    m = A.load_model_data(path, 1)
    A.place_default(m, 0.0, translate=(0.0, 0.0, 0.0))
    sc = A.Scene(m, division_method=A.DIV_CENTER, max_no_faces_per_node=cap)
    A.prep_scene(sc)
    tree = m.tree()
    assert (tree["boxes"][tree["children_start"][0], 3:] == 0.0).all()   # root split point = hi corner of child 0 = origin
    o = B.Scene(port, path, rescale=0.0, translate=(0.0, 0.0, 0.0), method=0, cap=cap)
    for eye in ((-3.0, -3.0, 10.0), (-3.0, -3.0, 2.5), (0.0, 0.0, 10.0), (1.0, 1.0, 0.5)):
        cam = A.set_camera(256, 256, eye=eye, facing=(0.0, 0.0, -1.0))
        a = sc.ctx.render(cam)
        b = sc.ctx.render(cam, sc.ctx.opts(cull=False))
        c = sc.ctx.render(cam, sc.ctx.opts(raster=True))
        face, t, img = o.trace_primary(256, 256, threads=NTHREADS, eye=eye, facing=(0.0, 0.0, -1.0))
        assert (a["prim"] == face).all() and same_bits(a["t"], t) and (a["bgrx"] == img).all()
        assert (b["prim"] == face).all() and same_bits(b["t"], t)
        assert (c["prim"] == face).all() and same_bits(c["t"], t)
        assert (face != B.MISS).any()
    o.close()


def test_c5_incoherent_rays(A, port, standin):
    """BASELINE.json configs[4] (sub-sample): random-direction rays through the ray-list entry point."""
    from oracle import bind as B
    sc = gpu_scene(A, standin, 0.0)
    o = B.Scene(port, standin, rescale=0.0)
    rays = random_rays(sc.model.dump()["aabb"], 1 << 21, 1234)
    prim, t = sc.ctx.trace_rays(rays, count_work=True)
    assert sc.ctx.counters()["rays"] == rays.shape[0]
    f, tt = o.trace_rays(rays, threads=NTHREADS)
    assert (prim == f).all() and same_bits(t, tt)
    assert 0.3 < float((f != B.MISS).mean()) < 0.95
    prim_f, t_f = sc.ctx.trace_rays(rays, fp_mode=A.FP_FAST)
    # FAST (FMA) is opt-in; on incoherent rays its measured id mismatch rate is ~3.5e-5, STRICT is exact
    assert int((prim_f != f).sum()) <= 1e-4 * rays.shape[0]
    o.close()


def test_c5_device_generated_rays_and_pixel_means(A, port, standin):
    """BASELINE.json configs[4] as defined in SURVEY §8(d): 64 incoherent rays per pixel index, generated ON THE DEVICE from
    the per-ray PCG32 stream, traced, reduced to the pixel's mean colour. 131072 pixels x 64 spp = 8.39 M rays against the
    oracle's restatement of the same stream: per-ray primitive id and t bit-exact, per-pixel BGRX equal; the round-robin
    chunk split over ranks (executed rank by rank here) reproduces the whole job; a second seed gives other rays."""
    from oracle import bind as B
    sc = gpu_scene(A, standin, 0.0)
    o = B.Scene(port, standin, rescale=0.0)
    p0, p1, spp = 4_000_000, 4_000_000 + 131072, 64
    got = sc.ctx.trace_incoherent(p0, p1, spp, SEED, count_work=True)
    assert sc.ctx.counters()["rays"] == (p1 - p0) * spp
    face, t, bgrx = o.trace_c5(p0, p1, spp, SEED, threads=NTHREADS)
    assert (got["prim"] == face).all() and same_bits(got["t"], t)
    assert (got["bgrx"] == bgrx).all()
    assert 0.3 < float((face != B.MISS).mean()) < 0.95 and len(np.unique(bgrx)) > 8     # the means really are blends
    # rank split: chunks of 1000 pixels dealt to 3 ranks, one rank after the other into the same device buffers
    ctx = sc.ctx
    n_px, n = p1 - p0, (p1 - p0) * spp
    d_b, d_p, d_t = ctx.dev_alloc(n_px * 4), ctx.dev_alloc(n * 4), ctx.dev_alloc(n * 4)
    for d, nb in ((d_b, n_px * 4), (d_p, n * 4), (d_t, n * 4)):
        ctx.dev_memset(d, 0xEE, nb)
    ctx.trace_incoherent_device(p0, p1, spp, d_b, d_p, d_t, seed=SEED, rank=1, world=3, chunk_px=1000)
    ctx.sync()
    part = ctx.to_host(d_b, np.zeros(n_px, np.uint32))
    mine = (np.arange(n_px) // 1000) % 3 == 1
    assert (part[mine] == bgrx[mine]).all() and (part[~mine] == 0xEEEEEEEE).all()     # a rank touches only its chunks
    for r in (0, 2):
        ctx.trace_incoherent_device(p0, p1, spp, d_b, d_p, d_t, seed=SEED, rank=r, world=3, chunk_px=1000)
    ctx.sync()
    assert (ctx.to_host(d_b, np.zeros(n_px, np.uint32)) == bgrx).all()
    assert (ctx.to_host(d_p, np.zeros(n, np.uint32)) == face).all()
    assert same_bits(ctx.to_host(d_t, np.zeros(n, np.float32)), t)
    for d in (d_b, d_p, d_t):
        ctx.dev_free(d)
    # odd sample count (not a multiple of the 32 samples a warp takes at a time), other seed, small range
    g2 = sc.ctx.trace_incoherent(7, 7 + 999, 37, SEED + 1)
    f2, t2, b2 = o.trace_c5(7, 7 + 999, 37, SEED + 1, threads=NTHREADS)
    assert (g2["prim"] == f2).all() and same_bits(g2["t"], t2) and (g2["bgrx"] == b2).all()
    assert not (g2["bgrx"][:900] == got["bgrx"][:900]).all()
    o.close()


def test_partition_union_equals_whole(A, meshes):
    """Multi-GPU split (interleaved scanline bands) executed sequentially on one GPU reproduces the 1-GPU frame."""
    path, rescale = meshes["knot_mid"]
    sc = gpu_scene(A, path, rescale)
    W, H = 500, 301
    cam = A.set_camera(W, H)
    whole = sc.ctx.render(cam)
    ctx = sc.ctx
    d = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
    for world, band in ((2, 8), (4, 16), (8, 4), (3, 12)):
        for p in d:
            ctx.from_host(p, np.full(W * H, 0xDEADBEEF, np.uint32))
        for rank in range(world):
            ctx.render_device(cam, ctx.opts(rank=rank, world=world, band_rows=band), d[0], d[1], d[2])
        ctx.sync()
        bg = ctx.to_host(d[0], np.zeros((H, W), np.uint32))
        t = ctx.to_host(d[1], np.zeros((H, W), np.float32))
        pr = ctx.to_host(d[2], np.zeros((H, W), np.uint32))
        assert (bg == whole["bgrx"]).all() and same_bits(t, whole["t"]) and (pr == whole["prim"]).all()
    # one rank alone writes only its own bands
    for p in d:
        ctx.from_host(p, np.full(W * H, 0xDEADBEEF, np.uint32))
    ctx.render_device(cam, ctx.opts(rank=1, world=2, band_rows=8), d[0], d[1], d[2])
    ctx.sync()
    bg = ctx.to_host(d[0], np.zeros((H, W), np.uint32))
    rows = np.arange(H)
    mine = (rows // 8) % 2 == 1
    assert (bg[mine] == whole["bgrx"][mine]).all() and (bg[~mine] == 0xDEADBEEF).all()
    for p in d:
        ctx.dev_free(p)


def test_partitioned_host_copy_out_writes_only_own_scanlines(A, meshes):
    """atray_render_begin with a partition writes exactly the rank's scanlines of the host frame (pinned and pageable
    hosts): the ranks of a multi-GPU job render into ONE host bitmap, as the reference's pool threads do."""
    path, rescale = meshes["knot_mid"]
    sc = gpu_scene(A, path, rescale)
    ctx = sc.ctx
    for (W, H) in ((500, 301), (1920, 1084)):   # the second size takes the slab-pipelined copy-out (>= 1 Mpx), clipped last band
        cam = A.set_camera(W, H)
        whole = ctx.render(cam)
        for world, band in ((2, 8), (3, 12), (8, 8)):
            pin = {k: A.PinnedArray((H, W), np.uint32) for k in ("bgrx", "t", "prim")}
            page = {k: np.full((H, W), 0xDEADBEEF, np.uint32) for k in ("bgrx", "t", "prim")}
            for a in pin.values():
                a.array[:] = 0xDEADBEEF
            for rank in range(world):
                for host in (pin, page):
                    arr = {k: (v.array if host is pin else v) for k, v in host.items()}
                    ctx.render_begin(cam, ctx.opts(rank=rank, world=world, band_rows=band), arr["bgrx"], arr["t"].view(np.float32), arr["prim"])
                    while ctx.render_poll(1000):
                        pass
                if rank == 0:   # after the first rank only its bands are written
                    rows = np.arange(H)
                    mine = (rows // band) % world == 0
                    assert (pin["bgrx"].array[mine] == whole["bgrx"][mine]).all() and (pin["bgrx"].array[~mine] == 0xDEADBEEF).all()
                    assert (page["prim"][mine] == whole["prim"][mine]).all() and (page["prim"][~mine] == 0xDEADBEEF).all()
            for host in (pin, page):
                arr = {k: (v.array if host is pin else v) for k, v in host.items()}
                assert (arr["bgrx"] == whole["bgrx"]).all() and (arr["prim"] == whole["prim"]).all()
                assert (arr["t"] == whole["t"].view(np.uint32)).all()
            for a in pin.values():
                a.free()


def test_spill_path_long_leaf_lists(A, port, meshes):
    """Rays whose leaf list outgrows the per-lane array take the per-warp spill path; results stay bit-exact."""
    from oracle import bind as B
    path, rescale = meshes["knot_mid"]
    sc = gpu_scene(A, path, rescale, 0, 6)  # CENTER, cap 6 → 30 k tiny leaves, long lists (caps < 6 trip the build guard)
    o = B.Scene(port, path, rescale=rescale, method=0, cap=6)
    rays = random_rays(sc.model.dump()["aabb"], 200000, 3)
    prim, t = sc.ctx.trace_rays(rays, count_work=True)
    ctr = sc.ctx.counters()
    f, tt = o.trace_rays(rays, threads=NTHREADS)
    assert (prim == f).all() and same_bits(t, tt)
    assert ctr["overflow_rays"] > 0, "test mesh no longer exercises the spill path"
    out = sc.ctx.render(A.set_camera(640, 360))
    face, t2, img = o.trace_primary(640, 360, threads=NTHREADS)
    assert (out["prim"] == face).all() and same_bits(out["t"], t2)
    o.close()


def test_edge_rays_and_empty_scene(A, port, tmp_path):
    """Axis-parallel rays (1/0 = inf, 0*inf = NaN in the slab test), rays starting inside boxes, zero directions,
    an empty model (the reference's Simple.obj placeholder) and a root-leaf model."""
    from oracle import bind as B
    v, f = cube_mesh()
    c = write_obj(str(tmp_path / "cube.obj"), v, f)
    sc = gpu_scene(A, c, 40.0)
    o = B.Scene(port, c, rescale=40.0)
    box = sc.model.dump()["aabb"]
    ctr = (box[:3] + box[3:]) / 2
    rays = []
    for ax in range(3):
        for sgn in (1.0, -1.0):
            d = np.zeros(3); d[ax] = sgn
            for off in (0.0, 1.0, 19.99, 20.0, 20.0001, 25.0):
                org = ctr - 60 * d
                org[(ax + 1) % 3] += off
                rays.append(np.concatenate([org, d]))
            rays.append(np.concatenate([ctr, d]))           # from inside
            rays.append(np.concatenate([box[:3], d]))       # from a corner
    rays.append(np.concatenate([ctr, np.zeros(3)]))         # zero direction
    rays.append(np.concatenate([ctr - [0, 0, 60], [0, 1e-30, 1]]))
    rays = np.array(rays, np.float32)
    prim, t = sc.ctx.trace_rays(rays)
    ff, tt = o.trace_rays(rays)
    assert (prim == ff).all() and same_bits(t, tt)
    assert (ff != B.MISS).any() and (ff == B.MISS).any()
    o.close()
    e = tmp_path / "empty.obj"
    e.write_text(" n ")
    sc = gpu_scene(A, str(e), 40.0)
    out = sc.ctx.render(A.set_camera(640, 360))
    assert (out["prim"] == A.MISS).all() and (out["bgrx"] == 0x004C667F).all() and (out["t"] == A.MAX_FLOAT).all()
    prim, t = sc.ctx.trace_rays(rays)
    assert (prim == A.MISS).all()
    prim, t = sc.ctx.trace_rays(np.zeros((0, 6), np.float32))
    assert prim.size == 0


def test_reference_call_sequence(A, port, meshes):
    """The drop-in surface: load → get_AABB/transforms → set_camera → prep_scene → start_… → poll wait_…(33 ms)
    (app.cpp:62-187), pinned host bitmap, TRUE-while-running / FALSE-when-done semantics."""
    from oracle import bind as B
    path, rescale = meshes["knot_mid"]
    model = A.load_model_data(path, 2)
    A.get_AABB(model)
    A.translate_to(model, (0.0, -15.0, -38.0))
    cam = A.set_camera(1280, 720)
    scene = A.Scene(model)
    nodes = A.prep_scene(scene)
    assert nodes == model.counts()["nodes"]
    tex = A.PinnedArray((720, 1280), np.uint32)
    prim = np.zeros((720, 1280), np.uint32)  # pageable on purpose: exercises the staging path
    info = A.RenderInfo(cam, scene, tex.array, prim=prim)
    A.start_render_from_camera(info)
    polls = 0
    while A.wait_for_render_from_camera_to_finish(info, None, 33):
        polls += 1
    assert info.jobs_done == info.jobs_size == 720 // 8
    o = B.Scene(port, path, rescale=0.0)
    face, t, img = o.trace_primary(1280, 720, threads=NTHREADS)
    assert (tex.array == img).all() and (prim == face).all()
    assert info.total_ray_casts == int((face != B.MISS).sum())
    assert A.wait_for_render_from_camera_to_finish(info, None, 0) == 0  # idle → finished
    with pytest.raises(A.AtrayError):
        scene.ctx.render(A.set_camera(64, 64, bounce_limit=0))
    tex.free()
    o.close()


def test_heaviest_first_tile_schedule(A, meshes, monkeypatch):
    """The busy tiles are listed per cost class and started most expensive class first (ATRAY_TILE_LOG records it); the
    frame does not depend on the schedule: any class shift, including a single class, gives the same bits."""
    path, rescale = meshes["knot_mid"]
    W, H = 1920, 1080
    cam = A.set_camera(W, H)
    monkeypatch.setenv("ATRAY_TILE_LOG", "1")
    frames = []
    for shift in ("-3", "0", "31"):     # -3 = the default (one packet in 8 is counted, classes three octaves lower)
        monkeypatch.setenv("ATRAY_COST_SHIFT", shift)
        sc = gpu_scene(A, path, rescale)
        ctx = sc.ctx
        planes = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
        ctx.render_device(cam, ctx.opts(count_work=True), *planes)   # device frame: one slab, one list per class
        ctx.sync()
        out = {}
        for k, dt, d in zip(("bgrx", "t", "prim"), (np.uint32, np.float32, np.uint32), planes):
            out[k] = ctx.to_host(d, np.zeros((H, W), dt))
            ctx.dev_free(d)
        frames.append(out)
        ctr = ctx.counters()
        log = ctx.tile_log(((H + 7) // 8) * 2 * ((W + 7) // 8))
        busy = log[log[:, 1] > 0]
        assert len(busy) == ctr["tiles_busy"] > 0
        if shift == "31":
            assert (busy[:, 3] == 0).all()
        else:
            cls = busy[:, 3].astype(np.int64)
            assert len(np.unique(cls)) > 1, "scene no longer spreads over several cost classes"
            expect = np.clip(np.floor(np.log2(np.maximum(busy[:, 2], 1))).astype(np.int64) - int(shift), 0, 7)
            assert (cls == expect).all()
            # one cursor walks the classes downwards: a class's first tile starts after the first tile of every heavier
            # class (2 us of slack: warps that draw neighbouring list positions read the timer in either order)
            start = (busy[:, 0].astype(np.int64) - int(busy[:, 0].min())) & 0xffffffff
            first = [start[cls == c].min() for c in sorted(np.unique(cls), reverse=True)]
            assert all(a <= b + 2000 for a, b in zip(first, first[1:])), first
    for f in frames[1:]:
        assert (f["bgrx"] == frames[0]["bgrx"]).all() and (f["prim"] == frames[0]["prim"]).all() and same_bits(f["t"], frames[0]["t"])


def test_view_pass_in_shards_gives_the_same_frame(A, meshes):
    """The per-frame view pass split into shards of the triangle slots (what N GPUs that share the pass run, one shard per
    GPU with stores into every peer's tables) — executed here as shard launches on ONE GPU with this GPU standing in for
    every peer (atray_debug_view_shards): whole frame, a rank of 8, the jittered path and the slab-pipelined host frame
    give the same bits and the same tile counts as the unsharded pass."""
    path, rescale = meshes["knot_mid"]
    W, H = 1920, 1084
    results = []
    # ATRAY_SHARE_MODE (read per frame): default = ranges only + k_view_stamp on the owning rank; 0 = stamps stored by the computing rank
    for shards, mode in ((0, None), (2, None), (3, None), (8, None), (3, "0"), (8, "0")):
        os.environ.pop("ATRAY_SHARE_MODE", None)
        if mode is not None:
            os.environ["ATRAY_SHARE_MODE"] = mode
        sc = gpu_scene(A, path, rescale)
        ctx = sc.ctx
        ctx.debug_view_shards(shards)
        out = {}
        for tag, cam, opts in (("whole", A.set_camera(W, H), ctx.opts(count_work=True)),
                               ("rank5of8", A.set_camera(W, H), ctx.opts(rank=5, world=8, band_rows=8, count_work=True)),
                               ("aa", A.set_camera(640, 362, anti_aliasing=True, samples_per_pixel=4), ctx.opts(count_work=True))):
            h, w = cam.height, cam.width
            planes = [ctx.dev_alloc(w * h * 4) for _ in range(3)]
            for d in planes:
                ctx.dev_memset(d, 0, w * h * 4)
            for _ in range(2):      # twice: the second frame meets the tables the first one left behind
                ctx.render_device(cam, opts, *planes)
            ctx.sync()
            for k, dt, d in zip(("bgrx", "t", "prim"), (np.uint32, np.float32, np.uint32), planes):
                out[tag + "." + k] = ctx.to_host(d, np.zeros((h, w), dt)).view(np.uint32)
                ctx.dev_free(d)
            ctr = ctx.counters()
            out[tag + ".ctr"] = np.array([ctr["tiles_busy"], ctr["tiles_sky"], ctr["rays"]], np.uint64)
            assert ctr["tiles_busy"] > 0
        host = ctx.render(A.set_camera(W, H))
        for k in ("bgrx", "t", "prim"):
            out["host." + k] = np.ascontiguousarray(host[k]).view(np.uint32)
        results.append(out)
    os.environ.pop("ATRAY_SHARE_MODE", None)
    for other in results[1:]:
        assert other.keys() == results[0].keys()
        for k in results[0]:
            assert (other[k] == results[0][k]).all(), k
    assert (results[0]["host.bgrx"] == results[0]["whole.bgrx"]).all()


def test_shared_view_pass_on_two_gpus(A, meshes):
    """Two GPUs of one process share the view pass (atray_view_share_*): each computes half of the slot ranges and stores
    them into both GPUs' tables over NVLink, a flag per rank orders the exchange, and each renders its bands straight into
    GPU 0's frame (peer stores). The gathered frame equals the 1-GPU frame bit for bit, frame after frame (the tables are
    double-buffered), for two cameras."""
    if A.device_count() < 2:
        pytest.skip("needs two GPUs (run under gpurun --gpus 2)")
    from atray_b200 import multi
    from atray_b200.renderer import Context
    path, rescale = meshes["knot_mid"]
    W, H = 1920, 1084
    m = A.load_model_data(path, 4)
    A.place_default(m, rescale)
    A.build_KD_tree(m)
    ctxs = [Context(0), Context(1)]
    for c in ctxs:
        c.upload(m)
    multi.ViewShare.connect_local(ctxs, W, H)
    planes = [ctxs[0].dev_alloc(W * H * 4) for _ in range(3)]      # GPU 0 owns the frame; GPU 1 stores into it
    for cam in (A.set_camera(W, H), A.set_camera(W, H, eye=(3.0, -4.0, 9.0), facing=(-0.2, -0.4, -1.0), h_fov=0.7)):
        ref = ctxs[0].render(cam)                                   # whole frame, pass computed by one GPU
        for frame in range(5):
            for d in planes:
                ctxs[0].dev_memset(d, 0xEE, W * H * 4)
            ctxs[0].sync()
            for r, c in enumerate(ctxs):                            # both ranks are queued before either is waited for
                c.render_device(cam, c.opts(rank=r, world=2, band_rows=8), *planes)
            for c in ctxs:
                c.sync()
            got = [ctxs[0].to_host(d, np.zeros((H, W), np.uint32)) for d in planes]
            assert (got[0] == ref["bgrx"]).all(), frame
            assert (got[1] == ref["t"].view(np.uint32)).all() and (got[2] == ref["prim"]).all(), frame
    for c in ctxs:
        c.view_share_check()       # no wait for a peer's flag timed out
    for d in planes:
        ctxs[0].dev_free(d)
    for c in ctxs:
        c.close()


def test_pipelined_copy_out_equals_whole_frame(A, meshes, monkeypatch):
    """Frames of ≥ 1 Mpx through atray_render_begin are rendered slab by slab and copied out while later slabs are
    still being traced (copy stream waiting on the kernel's per-slab counters). The host frame must equal the
    device-resident frame of atray_render_device and the unpipelined copy-out (ATRAY_SLABS=1) bit for bit; progress
    is monotone and ends at the frame height; repeated frames reuse the counters correctly."""
    path, rescale = meshes["knot_mid"]
    W, H = 1920, 1084   # height not a multiple of the band height: the last band is partial
    cam = A.set_camera(W, H)
    ref_sc = gpu_scene(A, path, rescale)
    whole = ref_sc.ctx.render(cam)   # synchronous convenience call = begin + poll on pageable arrays (staged)
    for slabs in ("1", "3", "8", "16", "w:2,6,8,16", "w:1,1,1,9", "w:5,1", "w:2,6,8,8,6,2/1"):
        monkeypatch.delenv("ATRAY_SLABS", raising=False)
        monkeypatch.delenv("ATRAY_SLAB_WEIGHTS", raising=False)
        monkeypatch.delenv("ATRAY_COPY_STREAMS", raising=False)
        if slabs.startswith("w:"):   # graded slab heights; "/1" = one copy stream instead of two
            plan = slabs[2:].split("/")
            monkeypatch.setenv("ATRAY_SLAB_WEIGHTS", plan[0])
            if len(plan) > 1:
                monkeypatch.setenv("ATRAY_COPY_STREAMS", plan[1])
        else:
            monkeypatch.setenv("ATRAY_SLABS", slabs)
        sc = gpu_scene(A, path, rescale)
        tex, tt, tp = A.PinnedArray((H, W), np.uint32), A.PinnedArray((H, W), np.float32), A.PinnedArray((H, W), np.uint32)
        for rep in range(3):
            tex.array[:] = 0xdeadbeef
            tp.array[:] = 7
            info = A.RenderInfo(cam, sc, tex.array, t=tt.array, prim=tp.array)
            A.start_render_from_camera(info)
            seen = [0]
            while A.wait_for_render_from_camera_to_finish(info, None, 0):
                seen.append(sc.ctx.render_progress()[0])
            assert all(b >= a for a, b in zip(seen, seen[1:])), seen
            assert sc.ctx.render_progress() == (H, H)
            assert info.jobs_done == info.jobs_size == (H + 7) // 8
            assert (tex.array == whole["bgrx"]).all() and (tp.array == whole["prim"]).all()
            assert same_bits(tt.array, whole["t"])
        for a in (tex, tt, tp):
            a.free()


@pytest.mark.parametrize("name", ["knot_small", "Deer", "Monkey"])
def test_multi_bounce_path_matches_reference_golden(name, A, meshes):
    """cast_ray at bounce_limit > 1, smooth/flat normals, spheres and planes — k_path against the
    vectors the reference's own code produced (tests/golden/path_*.npz): pixels, t, primitive id, ray_casts bit-exact."""
    if name not in meshes:
        pytest.skip("mesh not staged")
    g = np.load(os.path.join(GOLDEN, "path_%s.npz" % name))
    path, rescale = meshes[name]
    for prims in (0, 1):
        sc = gpu_scene(A, path, rescale)
        if prims:
            sc.ctx.set_primitives(PATH_SPHERES, PATH_PLANES)
        for k, (spp, bounce, aa) in enumerate(PATH_CONFIGS):
            cam = A.set_camera(PW, PH, anti_aliasing=aa, samples_per_pixel=spp, bounce_limit=bounce)
            out = sc.ctx.render(cam, sc.ctx.opts(seed=SEED + k))
            tag = "p%d_c%d_" % (prims, k)
            assert (out["bgrx"] == g[tag + "bgrx"]).all(), tag
            assert (bits(out["t"]) == g[tag + "t_bits"]).all() and (out["prim"] == g[tag + "prim"]).all(), tag
            if bounce != 1 or prims:
                assert sc.ctx.last_ray_casts() == int(g[tag + "casts"])


@pytest.mark.parametrize("name", ["knot_mid", "Cube", "Monkey"])
def test_multi_bounce_path_equals_oracle(name, A, port, meshes):
    """Larger frames, other cameras (inside the model, grazing the plane), 5 spp x 5 bounces (the reference's default
    workload, app.cpp:84-85), through the reference-named entry points with Scene.spheres / Scene.planes."""
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    path, rescale = meshes[name]
    m = A.load_model_data(path, 2)
    A.place_default(m, rescale)
    scene = A.Scene(m)
    scene.spheres, scene.planes = list(PATH_SPHERES), list(PATH_PLANES)
    A.prep_scene(scene)
    o = B.Scene(port, path, rescale=rescale)
    for c, r, mm in PATH_SPHERES:
        o.add_sphere(c, r, mm)
    for n, d, mm in PATH_PLANES:
        o.add_plane(n, d, mm)
    for i, (eye, facing, aa) in enumerate([(B.DEFAULT_EYE, B.DEFAULT_FACING, 0), ((0.0, -14.0, -36.0), (0.2, 0.1, -1.0), 1),
                                           ((20.0, -19.5, -10.0), (-0.6, 0.0, -1.0), 1)]):
        W, H = 401, 227
        cam = A.set_camera(W, H, eye=eye, facing=facing, anti_aliasing=aa, samples_per_pixel=5, bounce_limit=5)
        tex, tt, tp = np.zeros((H, W), np.uint32), np.zeros((H, W), np.float32), np.zeros((H, W), np.uint32)
        info = A.RenderInfo(cam, scene, tex, t=tt, prim=tp, opts=scene.ctx.opts(seed=SEED + i))
        A.start_render_from_camera(info)
        while A.wait_for_render_from_camera_to_finish(info, None, 33):
            pass
        img, t, prim, kind, casts = o.render_path(W, H, 5, 5, SEED + i, aa=aa, threads=NTHREADS, eye=eye, facing=facing)
        assert (tex == img).all() and same_bits(tt, t) and (tp == prim).all()
        assert info.total_ray_casts == casts
        assert len(np.unique(kind)) >= 2
    # removing the primitives restores the culled primary-ray path
    scene.ctx.set_primitives()
    a = scene.ctx.render(A.set_camera(320, 180))
    face, t, img = B.Scene(port, path, rescale=rescale).trace_primary(320, 180, threads=NTHREADS)
    assert (a["prim"] == face).all() and (a["bgrx"] == img).all()
    o.close()


@pytest.mark.parametrize("name", ["knot_small", "Cube", "Deer", "Monkey"])
def test_no_tree_branch_matches_reference_golden_and_oracle(name, A, port, meshes):
    """atray_scene_use_kd_tree(ctx, 0) = the reference compiled without USE_KD_TREE (renderer.cpp:58-82): frames, jittered
    frames, ray lists and the multi-bounce path against golden vectors made by that reference build, then the tree back."""
    from oracle import bind as B
    if name not in meshes:
        pytest.skip("mesh not staged")
    g = np.load(os.path.join(GOLDEN, "notree_%s.npz" % name))
    path, rescale = meshes[name]
    sc = gpu_scene(A, path, rescale)
    with_tree = sc.ctx.render(A.set_camera(PW, PH))
    sc.ctx.use_kd_tree(False)
    out = sc.ctx.render(A.set_camera(PW, PH))
    assert (out["prim"] == g["face"]).all() and (bits(out["t"]) == g["t_bits"]).all() and (out["bgrx"] == g["bgrx"]).all()
    jit = sc.ctx.render(A.set_camera(PW, PH, anti_aliasing=1, samples_per_pixel=4), sc.ctx.opts(seed=SEED))
    assert (jit["bgrx"] == g["jitter_bgrx"]).all()
    prim, t = sc.ctx.trace_rays(random_rays(sc.model.dump()["aabb"], 4000, int(g["ray_seed"])))
    assert (prim == g["ray_face"]).all() and (bits(t) == g["ray_t_bits"]).all()
    sc.ctx.set_primitives(PATH_SPHERES, PATH_PLANES)
    p = sc.ctx.render(A.set_camera(PW, PH, anti_aliasing=1, samples_per_pixel=2, bounce_limit=4), sc.ctx.opts(seed=SEED + 1))
    assert (p["bgrx"] == g["path_bgrx"]).all() and (bits(p["t"]) == g["path_t_bits"]).all() and (p["prim"] == g["path_prim"]).all()
    assert sc.ctx.last_ray_casts() == int(g["path_casts"])
    sc.ctx.set_primitives([], [])
    # other camera and size against the oracle's no-tree branch
    o = B.Scene(port, path, rescale=rescale)
    o.use_kd_tree(False)
    eye, facing = (3.0, -4.0, 9.0), (-0.2, -0.4, -1.0)
    a = sc.ctx.render(A.set_camera(333, 187, eye=eye, facing=facing, h_fov=0.7))
    face, tt, img = o.trace_primary(333, 187, threads=NTHREADS, eye=eye, facing=facing, hfov=0.7)
    assert (a["prim"] == face).all() and same_bits(a["t"], tt) and (a["bgrx"] == img).all()
    o.close()
    sc.ctx.use_kd_tree(True)
    back = sc.ctx.render(A.set_camera(PW, PH))
    assert (back["prim"] == with_tree["prim"]).all() and same_bits(back["t"], with_tree["t"])


def test_errors_on_device_side(A, meshes):
    ctx = A.Context(0)
    with pytest.raises(A.AtrayError):
        ctx.render(A.set_camera(64, 64))  # no scene uploaded
    with pytest.raises(A.AtrayError):
        A.Context(9999)
    v, f = cube_mesh()
    m = A.model_from_arrays(v, f)
    with pytest.raises(A.AtrayError):
        ctx.upload(m)  # tree not built
    A.place_default(m, 40.0)
    A.build_KD_tree(m)
    ctx.upload(m)
    with pytest.raises(A.AtrayError):
        ctx.render(A.set_camera(64, 64), ctx.opts(band_rows=6))
    with pytest.raises(A.AtrayError):
        ctx.render(A.set_camera(64, 64), ctx.opts(rank=2, world=2))
    out = ctx.render(A.set_camera(64, 64))
    assert (out["prim"] != A.MISS).any()
    ctx.close()


def test_cpp_host_layer_replays_app_cpp(A, port, meshes, tmp_path):
    """Source/engine (C++, reference entry-point names) driven by Source/engine/app_headless.cpp — the call order of
    render_app (app.cpp:59-195) — produces the oracle's bitmap, t and primitive planes."""
    import subprocess
    from oracle import bind as B
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "Source", "engine", "app_headless")
    assert os.path.exists(exe), "run __graft_entry__.build()"
    path, rescale = meshes["knot_mid"]
    planes = str(tmp_path / "planes.bin")
    bmp = str(tmp_path / "resultings")   # Write_To_File appends "_<id>.bmp" (texture.cpp:98-106)
    W, H = 1280, 720
    out = subprocess.run([exe, path, str(W), str(H), "0", bmp, planes], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert out.returncode == 0, out.stdout
    raw = np.fromfile(planes, np.uint32).reshape(3, H, W)
    o = B.Scene(port, path, rescale=0.0)
    face, t, img = o.trace_primary(W, H, threads=NTHREADS)
    assert (raw[0] == img).all() and (raw[1] == t.view(np.uint32)).all() and (raw[2] == face).all()
    assert "total_ray_casts %d " % int((face != B.MISS).sum()) in out.stdout
    b = open(bmp + "_0.bmp", "rb").read()
    assert b[:2] == b"BM" and len(b) == 70 + W * H * 4 and (np.frombuffer(b[70:], np.uint32).reshape(H, W) == img).all()
    o.close()
