"""GPU: the rest of the drop-in boundary through the C ABI and the C++ host layer:
the cancel that reaches the kernels (info.twq.clear(), app.cpp:165-172), the single-ray query (get_ray_kd_tree_intersection,
kd_tree.h:56), several GPUs from one process (atray_multi, Scene.gpu_devices), the flag CLI of app_headless."""
import os
import subprocess
import time

import numpy as np
import pytest

from helpers import random_rays, same_bits

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "Source", "engine", "app_headless")
NTHREADS = min(16, os.cpu_count() or 1)


@pytest.fixture(scope="module")
def A(built):
    import atray_b200 as A
    assert A.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    return A


def gpu_scene(A, path, rescale, method=1, cap=300, devices=None):
    m = A.load_model_data(path, 4)
    A.place_default(m, rescale)
    sc = A.Scene(m, division_method=method, max_no_faces_per_node=cap)
    sc.gpu_devices = devices
    A.prep_scene(sc)
    return sc


def test_single_ray_query_gives_the_oracles_u_v(A, port, meshes):
    from oracle import bind as B
    path, rescale = meshes["knot_mid"]
    sc = gpu_scene(A, path, rescale)
    o = B.Scene(port, path, rescale=rescale)
    rays = random_rays(sc.model.dump()["aabb"], 30000, 21)
    prim, t, uv = sc.ctx.trace_rays_uv(rays)
    f, tt, ouv = o.trace_rays_uv(rays, threads=NTHREADS)
    assert (prim == f).all() and same_bits(t, tt) and same_bits(uv, ouv)      # rows of misses stay NaN on both sides
    hit = np.flatnonzero(f != B.MISS)
    assert hit.size > 100
    for i in list(hit[:3]) + [int(np.flatnonzero(f == B.MISS)[0])]:          # the reference's call shape: one ray per call
        td = {"face_index": 123456, "u": -1.0, "v": -1.0}
        dist = A.get_ray_kd_tree_intersection(rays[i], sc, td)
        assert np.float32(dist).view(np.uint32) == tt[i].view(np.uint32)
        if f[i] != B.MISS:
            assert td["face_index"] == f[i] and np.float32(td["u"]) == ouv[i, 0] and np.float32(td["v"]) == ouv[i, 1]
        else:
            assert td == {"face_index": 123456, "u": -1.0, "v": -1.0} and dist == A.MAX_FLOAT   # tri_data untouched on a miss
    o.close()


def test_cancel_stops_the_kernels_and_leaves_the_context_usable(A, standin):
    """A long job (C5 rays of 2 M pixels, ~0.5 s) is cut short by atray_render_cancel within milliseconds; entries not yet traced
    keep the buffer's old contents; the next frames are bit-identical to frames of a fresh context."""
    m = A.load_model_data(standin, 8)
    A.place_default(m, 0.0)
    sc = A.Scene(m)
    A.prep_scene(sc)
    ctx = sc.ctx
    cam = A.set_camera(640, 360)
    clean = ctx.render(cam)
    npx, spp = 1 << 21, 64
    d_bgrx = ctx.dev_alloc(npx * 4)
    ctx.dev_memset(d_bgrx, 0xEE, npx * 4)
    ctx.sync()
    t0 = time.perf_counter()
    ctx.trace_incoherent_device(0, npx, spp, d_bgrx=d_bgrx)
    ctx.sync()
    full_s = time.perf_counter() - t0
    done = ctx.to_host(d_bgrx, np.zeros(npx, np.uint32))
    assert (done != 0xEEEEEEEE).all()
    assert ctx.cancel() is False                      # idle context: nothing to cut short
    ctx.dev_memset(d_bgrx, 0xEE, npx * 4)
    ctx.sync()
    ctx.trace_incoherent_device(0, npx, spp, d_bgrx=d_bgrx)
    time.sleep(min(0.02, full_s / 10))
    t0 = time.perf_counter()
    assert ctx.cancel() is True
    cancel_s = time.perf_counter() - t0
    part = ctx.to_host(d_bgrx, np.zeros(npx, np.uint32))
    untouched = int((part == 0xEEEEEEEE).sum())
    assert untouched > npx // 4, "the kernel ran to the end (%d of %d pixels untouched, full job %.3f s)" % (untouched, npx, full_s)
    assert cancel_s < max(0.05, full_s / 4), "cancel took %.3f s of a %.3f s job" % (cancel_s, full_s)
    traced = part != 0xEEEEEEEE
    assert (part[traced] == done[traced]).all()       # what was traced before the cancel is final
    again = ctx.render(cam)                           # the cancel word is lowered again
    for k in clean:
        assert same_bits(again[k], clean[k]), k
    ctx.trace_incoherent_device(0, 4096, spp, d_bgrx=d_bgrx)
    ctx.sync()
    assert (ctx.to_host(d_bgrx, np.zeros(4096, np.uint32)) == done[:4096]).all()
    # begin / poll: a 5-bounce frame at 4K, cancelled while in flight through the host mirror of info.twq.clear()
    cam5 = A.set_camera(3840, 2160, anti_aliasing=1, samples_per_pixel=8, bounce_limit=5)
    tex = A.PinnedArray((2160, 3840), np.uint32)
    info = A.RenderInfo(cam5, sc, tex.array)
    A.start_render_from_camera(info)
    assert A.wait_for_render_from_camera_to_finish(info, None, 0) == 1
    t0 = time.perf_counter()
    assert A.clear_work_queue(info) is True
    assert A.wait_for_pool([], 0xFFFFFFFF) == 0
    assert time.perf_counter() - t0 < 0.1
    assert A.wait_for_render_from_camera_to_finish(info, None, 0) == 0
    A.start_render_from_camera(info)                  # starting over a frame in flight replaces it (no "already in flight" error)
    A.start_render_from_camera(info)
    while A.wait_for_render_from_camera_to_finish(info, None, 33):
        pass
    ref5 = gpu_scene(A, standin, 0.0).ctx.render(cam5, want=("bgrx",))["bgrx"]
    assert (tex.array == ref5).all()
    tex.free()
    ctx.dev_free(d_bgrx)


def test_two_contexts_of_one_process_fill_one_frame(A, meshes):
    """atray_multi with both contexts on GPU 0 (any box has one): partition, per-context scanline copies into one host frame,
    progress, ray-cast totals — the shared view pass stays off (it needs distinct GPUs)."""
    path, rescale = meshes["knot_mid"]
    one = gpu_scene(A, path, rescale)
    two = gpu_scene(A, path, rescale, devices=[0, 0])
    assert isinstance(two.ctx, A.MultiContext) and len(two.ctx.ctx) == 2
    for cam in (A.set_camera(1280, 720), A.set_camera(333, 101), A.set_camera(256, 128, anti_aliasing=1, samples_per_pixel=4)):
        a, b = one.ctx.render(cam), two.ctx.render(cam)
        for k in a:
            assert same_bits(a[k], b[k]), k
    cam = A.set_camera(1600, 900)
    tex = A.PinnedArray((900, 1600), np.uint32)
    prim = np.zeros((900, 1600), np.uint32)
    info = A.RenderInfo(cam, two, tex.array, prim=prim)
    A.start_render_from_camera(info)
    while A.wait_for_render_from_camera_to_finish(info, None, 33):
        pass
    ref = one.ctx.render(cam)
    assert (tex.array == ref["bgrx"]).all() and (prim == ref["prim"]).all()
    assert info.jobs_done == info.jobs_size and info.total_ray_casts == int((ref["prim"] != A.MISS).sum())
    # device-resident gather: both contexts store their bands into planes of context 0
    c0 = two.ctx.ctx[0]
    d = [c0.dev_alloc(1600 * 900 * 4) for _ in range(3)]
    two.ctx.render_device(cam, two.ctx.opts(), *d)
    two.ctx.sync()
    assert (c0.to_host(d[0], np.zeros((900, 1600), np.uint32)) == ref["bgrx"]).all()
    assert (c0.to_host(d[2], np.zeros((900, 1600), np.uint32)) == ref["prim"]).all()
    for p in d:
        c0.dev_free(p)
    tex.free()


def test_two_gpus_from_one_process(A, meshes):
    if A.device_count() < 2:
        pytest.skip("needs two GPUs")
    path, rescale = meshes["knot_mid"]
    one = gpu_scene(A, path, rescale)
    two = gpu_scene(A, path, rescale, devices=[0, 1])
    for cam in (A.set_camera(1920, 1080), A.set_camera(333, 101)):
        a = one.ctx.render(cam)
        b = two.ctx.render(cam)
        for k in a:
            assert same_bits(a[k], b[k]), k
    os.environ["ATRAY_MULTI_SHARE"] = "1"             # the opt-in shared view pass (ranges over NVLink, stamps on the owning GPU)
    two = gpu_scene(A, path, rescale, devices=[0, 1])
    del os.environ["ATRAY_MULTI_SHARE"]
    for cam in (A.set_camera(1920, 1080), A.set_camera(333, 101)):
        a = one.ctx.render(cam)
        for rep in range(3):                           # the shared view pass alternates its table buffers
            b = two.ctx.render(cam)
            for k in a:
                assert same_bits(a[k], b[k]), (k, rep)
    for c in two.ctx.ctx:
        assert A.api._L.atray_view_share_check(c.h) == 0
    planes = "/tmp/atray_two_gpus.bin"
    out = subprocess.run([EXE, "--obj", path, "--res", "1280x720", "--gpus", "2", "--planes", planes], stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    assert out.returncode == 0 and "gpus 2" in out.stdout, out.stdout
    raw = np.fromfile(planes, np.uint32).reshape(3, 720, 1280)
    ref = one.ctx.render(A.set_camera(1280, 720))
    assert (raw[0] == ref["bgrx"]).all() and (raw[1] == ref["t"].view(np.uint32)).all() and (raw[2] == ref["prim"]).all()


def test_app_headless_flags(A, port, meshes, tmp_path):
    """The flag CLI: tree method / leaf cap / resolution / spp / aa / mode, a probe ray through get_ray_kd_tree_intersection,
    and --cancel-after-ms (info.twq.clear() + wait_for_pool on a frame in flight)."""
    from oracle import bind as B
    assert os.path.exists(EXE), "run __graft_entry__.build()"
    path, rescale = meshes["knot_mid"]
    W, H = 640, 360
    o = B.Scene(port, path, rescale=0.0, method=0, cap=32)
    box = gpu_scene(A, path, 0.0).model.dump()["aabb"]
    ray = random_rays(box, 64, 3)
    f, tt, ouv = o.trace_rays_uv(ray)
    i = int(np.flatnonzero(f != B.MISS)[0])
    planes = str(tmp_path / "p.bin")
    out = subprocess.run([EXE, "--obj", path, "--res", "%dx%d" % (W, H), "--method", "center", "--leaf-cap", "32", "--planes", planes,
                          "--probe-ray", ",".join(repr(float(x)) for x in ray[i])], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert out.returncode == 0, out.stdout
    raw = np.fromfile(planes, np.uint32).reshape(3, H, W)
    face, t, img = o.trace_primary(W, H, threads=NTHREADS)
    assert (raw[0] == img).all() and (raw[1] == t.view(np.uint32)).all() and (raw[2] == face).all()
    want = "probe t_bits %08x face %u u_bits %08x v_bits %08x" % (tt[i].view(np.uint32), f[i], ouv[i, 0].view(np.uint32), ouv[i, 1].view(np.uint32))
    assert want in out.stdout, (want, out.stdout)
    o.close()
    # no-tree mode = the reference built without USE_KD_TREE (renderer.cpp:58-82)
    o = B.Scene(port, path, rescale=0.0)
    o.use_kd_tree(False)
    out = subprocess.run([EXE, "--obj", path, "--res", "160x90", "--mode", "notree", "--planes", planes], stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    assert out.returncode == 0, out.stdout
    raw = np.fromfile(planes, np.uint32).reshape(3, 90, 160)
    face, t, img = o.trace_primary(160, 90, threads=NTHREADS)
    assert (raw[0] == img).all() and (raw[2] == face).all()
    o.close()
    out = subprocess.run([EXE, "--obj", path, "--res", "3840x2160", "--spp", "8", "--aa", "1", "--bounces", "5", "--cancel-after-ms", "1"],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert out.returncode == 0 and "frame was running" in out.stdout and "still running: 0" in out.stdout, out.stdout
    assert subprocess.run([EXE, "--res", "10x10"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT).returncode == 2


def test_ipc_gather_from_a_second_process(A, meshes):
    """The device-resident gather of bench.py's N > 1 leg on one GPU: this process owns the three frame planes and renders rank 0
    of 2; a SECOND process opens the planes through CUDA IPC (atray_ipc_export / atray_ipc_open) and renders rank 1 into them.
    The union is the frame one context renders alone, and each process touched only its own bands."""
    import sys
    path, rescale = meshes["knot_mid"]
    W, H = 1280, 716
    sc = gpu_scene(A, path, rescale)
    ctx = sc.ctx
    cam = A.set_camera(W, H)
    whole = ctx.render(cam)
    planes = [ctx.dev_alloc(W * H * 4) for _ in range(3)]
    for p in planes:
        ctx.dev_memset(p, 0xEE, W * H * 4)
    ctx.sync()
    handles = [ctx.ipc_export(p).hex() for p in planes]
    child = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ipc_child.py"), path, repr(float(rescale)), str(W), str(H), "1", "2"] + handles,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert child.returncode == 0 and "child rank 1 of 2 done" in child.stdout, child.stdout[-2000:]
    half = [ctx.to_host(p, np.zeros((H, W), np.uint32)) for p in planes]
    mine = (np.arange(H) // 8) % 2 == 0                     # bands of 8 scanlines: even bands are rank 0's
    for got, k in zip(half, ("bgrx", "t", "prim")):
        ref = np.ascontiguousarray(whole[k]).view(np.uint32)
        assert (got[~mine] == ref[~mine]).all(), k          # the child's bands, stored from the other process
        assert (got[mine] == 0xEEEEEEEE).all(), k           # and nothing else
    ctx.render_device(cam, ctx.opts(rank=0, world=2, band_rows=8), *planes)
    ctx.sync()
    for p, k in zip(planes, ("bgrx", "t", "prim")):
        assert (ctx.to_host(p, np.zeros((H, W), np.uint32)) == np.ascontiguousarray(whole[k]).view(np.uint32)).all(), k
        ctx.dev_free(p)
