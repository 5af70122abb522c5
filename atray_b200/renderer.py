"""Device side of the Python mirror: prep_scene / start_render_from_camera / wait_for_render_from_camera_to_finish
(renderer.h:32-35) on top of the C ABI. Every call here ends in a CUDA kernel of libatray_b200.so."""
import ctypes as C

import numpy as np

from . import api
from .api import _L, _ck, _p, _vp


class PinnedArray:
    """numpy view over page-locked host memory (the host bitmap of Setup_Texture, texture.cpp:133-146)."""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        ptr = _vp()
        _ck(_L.atray_host_alloc_pinned(self.nbytes, C.byref(ptr)))
        self.ptr = ptr.value
        buf = (C.c_char * self.nbytes).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            _L.atray_host_free_pinned(self.ptr)
            self.ptr = None


class Context:
    """One GPU: uploaded scene + stream (atray_ctx)."""

    def __init__(self, device=0):
        h = _vp()
        _ck(_L.atray_ctx_create(device, C.byref(h)))
        self.h = h
        self.device = device
        self._keep = None

    def close(self):
        if getattr(self, "h", None) and not getattr(self, "_borrowed", False):
            _L.atray_ctx_destroy(self.h)
        self.h = None

    def __del__(self):
        self.close()

    # -- scene ------------------------------------------------------------------------------------------------------
    def upload(self, model, sky=api.SKY_MATERIAL, material=api.MODEL_MATERIAL):
        s, m = api.make_material(sky), api.make_material(material)
        _ck(_L.atray_scene_upload(self.h, model.h, C.byref(s), C.byref(m)))

    def set_primitives(self, spheres=(), planes=()):
        """scene.spheres / scene.planes (scene.h:13-24): spheres = [(center, radius, material)], planes = [(normal,
        distance, material)], material = (emission rgb, reflection rgb, scatter)."""
        sp = (api.Sphere * max(1, len(spheres)))()
        for i, (c, r, m) in enumerate(spheres):
            sp[i] = api.Sphere((C.c_float * 3)(*c), r, api.make_material(m))
        pl = (api.Plane * max(1, len(planes)))()
        for i, (n, d, m) in enumerate(planes):
            pl[i] = api.Plane((C.c_float * 3)(*n), d, api.make_material(m))
        _ck(_L.atray_scene_set_primitives(self.h, C.cast(sp, C.c_void_p), len(spheres), C.cast(pl, C.c_void_p), len(planes)))

    def use_kd_tree(self, use=True):
        """renderer.h:8 USE_KD_TREE as a run-time switch: False → every model query is the no-tree branch
        (renderer.cpp:58-82: model box, then every face in file order)."""
        _ck(_L.atray_scene_use_kd_tree(self.h, 1 if use else 0))

    def last_ray_casts(self):
        v = C.c_uint64(0)
        _ck(_L.atray_last_ray_casts(self.h, C.byref(v)))
        return v.value

    # -- frames -----------------------------------------------------------------------------------------------------
    @staticmethod
    def opts(fp_mode=api.FP_STRICT, count_work=False, seed=api.DEFAULT_SEED, rank=0, world=1, band_rows=8, cull=True, raster=False):
        """cull=False selects the unculled kernel (every listed leaf scanned completely, as the reference does);
        raster=True (frames without anti-aliasing) rasterises the frame triangle by triangle in film space instead of
        traversing tile by tile (raster.cuh) — same bits."""
        return api.RenderOpts(fp_mode, 1 if count_work else 0, seed, api.Partition(rank, world, band_rows),
                              (0 if cull else api.RENDER_NO_CULL) | (api.RENDER_RASTER if raster else 0))

    def render_begin(self, cam, opts, bgrx=None, t=None, prim=None):
        fr = api.Frame(_p(bgrx) if bgrx is not None else None, _p(t) if t is not None else None,
                       _p(prim) if prim is not None else None)
        self._keep = (bgrx, t, prim)
        _ck(_L.atray_render_begin(self.h, C.byref(cam), C.byref(opts), C.byref(fr)))

    def render_poll(self, timeout_ms=0xFFFFFFFF):
        return _ck(_L.atray_render_poll(self.h, timeout_ms))

    def render_progress(self):
        """(rows_done, rows_total): scanlines [0, rows_done) of the host frame are final (twq.jobs_done analogue)."""
        a, b = C.c_int32(0), C.c_int32(0)
        _ck(_L.atray_render_progress(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def render(self, cam, opts=None, want=("bgrx", "t", "prim")):
        """Synchronous convenience: one frame into fresh numpy arrays (row 0 = bottom scanline)."""
        opts = opts or self.opts()
        H, W = cam.height, cam.width
        out = {"bgrx": np.zeros((H, W), np.uint32) if "bgrx" in want else None,
               "t": np.zeros((H, W), np.float32) if "t" in want else None,
               "prim": np.zeros((H, W), np.uint32) if "prim" in want else None}
        self.render_begin(cam, opts, out["bgrx"], out["t"], out["prim"])
        while self.render_poll(1000):
            pass
        return out

    def render_device(self, cam, opts, d_bgrx=0, d_t=0, d_prim=0):
        fr = api.Frame(d_bgrx or None, d_t or None, d_prim or None)
        _ck(_L.atray_render_device(self.h, C.byref(cam), C.byref(opts), C.byref(fr)))

    def sync(self):
        _ck(_L.atray_ctx_sync(self.h))

    def last_kernel_ms(self):
        ms = C.c_float(0)
        _ck(_L.atray_last_kernel_ms(self.h, C.byref(ms)))
        return ms.value

    def last_render_kernel_ms(self):
        """Device time of the render kernel alone (k_render_culled) in the last culled frame."""
        ms = C.c_float(0)
        _ck(_L.atray_last_render_kernel_ms(self.h, C.byref(ms)))
        return ms.value

    def last_view_split_ms(self):
        """(k_view_bounds ms, k_view_nodes ms) of the last culled frame."""
        a, b = C.c_float(0), C.c_float(0)
        _ck(_L.atray_last_view_split_ms(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def tile_log(self, n_tiles):
        """(n_tiles, 4) uint32 {start ns, duration ns, sampled cost, class} of the last culled frame (ATRAY_TILE_LOG=1)."""
        out = np.zeros((n_tiles, 4), np.uint32)
        _ck(_L.atray_debug_tile_log(self.h, api._p(out), n_tiles))
        return out

    def counters(self):
        c = api.Counters()
        _ck(_L.atray_get_counters(self.h, C.byref(c)))
        return c.as_dict()

    # -- ray lists ----------------------------------------------------------------------------------------------------
    def trace_rays(self, rays, fp_mode=api.FP_STRICT, count_work=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        prim = np.zeros(n, np.uint32)
        t = np.zeros(n, np.float32)
        _ck(_L.atray_trace_rays(self.h, _p(rays), n, fp_mode, 1 if count_work else 0, _p(prim), _p(t)))
        return prim, t

    def trace_rays_device(self, d_rays, n, d_prim, d_t, fp_mode=api.FP_STRICT, count_work=False):
        _ck(_L.atray_trace_rays_device(self.h, d_rays, n, fp_mode, 1 if count_work else 0, d_prim, d_t))

    def trace_rays_uv(self, rays):
        """get_ray_kd_tree_intersection for a ray list (kd_tree.cpp:302-313): prim, t and TriangleIntersectionData's (u, v);
        rows of misses stay NaN."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        prim = np.zeros(n, np.uint32)
        t = np.zeros(n, np.float32)
        uv = np.full((n, 2), np.nan, np.float32)
        _ck(_L.atray_trace_rays_uv(self.h, _p(rays), n, _p(prim), _p(t), _p(uv)))
        return prim, t, uv

    def cancel(self):
        """atray_render_cancel: True when work was on the stream and has been cut short."""
        return _ck(_L.atray_render_cancel(self.h)) == 1

    def trace_rays_static(self, d_rays, n, d_prim, d_t, fp_mode=api.FP_STRICT):
        """trace_rays_device on the library's static device scratch: the call allocates nothing (include/atray_b200.h)."""
        _ck(_L.atray_trace_rays_static(self.h, d_rays, n, fp_mode, d_prim, d_t))

    def trace_incoherent(self, p0, p1, spp, seed=api.DEFAULT_SEED, fp_mode=api.FP_STRICT, count_work=False, want=("bgrx", "prim", "t")):
        """C5: `spp` device-generated incoherent rays per pixel index in [p0, p1) → dict of bgrx (per pixel), prim, t (per ray)."""
        n = (p1 - p0) * spp
        out = {"bgrx": np.zeros(p1 - p0, np.uint32) if "bgrx" in want else None, "prim": np.zeros(n, np.uint32) if "prim" in want else None,
               "t": np.zeros(n, np.float32) if "t" in want else None}
        _ck(_L.atray_trace_incoherent(self.h, p0, p1, spp, seed, fp_mode, 1 if count_work else 0, _p(out["bgrx"]), _p(out["prim"]), _p(out["t"])))
        return out

    def trace_incoherent_device(self, p0, p1, spp, d_bgrx=0, d_prim=0, d_t=0, seed=api.DEFAULT_SEED, rank=0, world=1, chunk_px=4096,
                                fp_mode=api.FP_STRICT, count_work=False):
        part = api.Partition(rank, world, 8)
        _ck(_L.atray_trace_incoherent_device(self.h, p0, p1, spp, seed, C.byref(part), chunk_px, fp_mode, 1 if count_work else 0,
                                             d_bgrx or None, d_prim or None, d_t or None))

    # -- memory -------------------------------------------------------------------------------------------------------
    def dev_alloc(self, nbytes):
        p = _vp()
        _ck(_L.atray_dev_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def dev_free(self, p):
        _ck(_L.atray_dev_free(self.h, p))

    def to_host(self, dev_ptr, arr):
        _ck(_L.atray_dev_copy_to_host(self.h, _p(arr), dev_ptr, arr.nbytes))
        return arr

    def from_host(self, dev_ptr, arr):
        arr = np.ascontiguousarray(arr)
        _ck(_L.atray_dev_copy_from_host(self.h, dev_ptr, _p(arr), arr.nbytes))

    def dev_memset(self, dev_ptr, value, nbytes):
        _ck(_L.atray_dev_memset(self.h, dev_ptr, value, nbytes))

    def ipc_export(self, dev_ptr):
        h = np.zeros(64, np.uint8)
        _ck(_L.atray_ipc_export(self.h, dev_ptr, _p(h)))
        return h.tobytes()

    def ipc_open(self, handle_bytes):
        h = np.frombuffer(handle_bytes, np.uint8).copy()
        p = _vp()
        _ck(_L.atray_ipc_open(self.h, _p(h), C.byref(p)))
        return p.value

    def ipc_close(self, p):
        _ck(_L.atray_ipc_close(self.h, p))

    # -- shared view pass (multi-GPU) ---------------------------------------------------------------------------------
    def view_share_create(self, max_width, max_height):
        """Allocates this context's block of the shared view pass; returns (device pointer, bytes)."""
        p, n = _vp(), C.c_size_t(0)
        _ck(_L.atray_view_share_create(self.h, max_width, max_height, C.byref(p), C.byref(n)))
        return p.value, n.value

    def view_share_connect(self, rank, world, peer_blocks):
        arr = (C.c_void_p * world)(*peer_blocks)
        _ck(_L.atray_view_share_connect(self.h, rank, world, C.cast(arr, C.c_void_p)))

    def view_share_rendezvous(self):
        _ck(_L.atray_view_share_rendezvous(self.h))

    def view_share_check(self):
        _ck(_L.atray_view_share_check(self.h))

    def view_share_off(self):
        _ck(_L.atray_view_share_off(self.h))

    def enable_peer(self, other):
        _ck(_L.atray_ctx_enable_peer(self.h, other.h))

    def debug_view_shards(self, shards):
        _ck(_L.atray_debug_view_shards(self.h, shards))

    def microbench(self):
        a, b = C.c_double(0), C.c_double(0)
        _ck(_L.atray_microbench(self.h, C.byref(a), C.byref(b)))
        return {"fp32_lane_ops_per_s": a.value, "l2_read_gbs": b.value}


# ---- reference-named entry points (renderer.h:32-35) -----------------------------------------------------------------
class MultiContext:
    """Several GPUs driven from this one process (atray_multi: atray_ctx_create_multi …): the reference's tile pool with GPUs as the
    workers. `ctx[i]` are borrowed Context views of the member contexts (counters, kernel times, device memory …)."""

    def __init__(self, devices=None, n=None):
        n = len(devices) if devices is not None else int(n)
        arr = (C.c_int32 * n)(*devices) if devices is not None else None
        h = _vp()
        _ck(_L.atray_ctx_create_multi(C.cast(arr, C.c_void_p) if arr is not None else None, n, C.byref(h)))
        self.h = h
        self.ctx = []
        for i in range(n):
            c = Context.__new__(Context)
            c.h, c.device, c._keep, c._borrowed = _vp(_L.atray_multi_ctx(self.h, i)), _L.atray_multi_device(self.h, i), None, True
            self.ctx.append(c)
        self.devices = [c.device for c in self.ctx]
        self._keep = None

    def close(self):
        if getattr(self, "h", None):
            for c in self.ctx:
                c.h = None
            _L.atray_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def upload(self, model, sky=api.SKY_MATERIAL, material=api.MODEL_MATERIAL):
        s, m = api.make_material(sky), api.make_material(material)
        _ck(_L.atray_multi_scene_upload(self.h, model.h, C.byref(s), C.byref(m)))

    def set_primitives(self, spheres=(), planes=()):
        for c in self.ctx:
            c.set_primitives(spheres, planes)

    def use_kd_tree(self, use=True):
        _ck(_L.atray_multi_scene_use_kd_tree(self.h, 1 if use else 0))

    opts = staticmethod(Context.opts)

    def render_begin(self, cam, opts, bgrx=None, t=None, prim=None):
        fr = api.Frame(_p(bgrx) if bgrx is not None else None, _p(t) if t is not None else None,
                       _p(prim) if prim is not None else None)
        self._keep = (bgrx, t, prim)
        _ck(_L.atray_multi_render_begin(self.h, C.byref(cam), C.byref(opts), C.byref(fr)))

    def render_poll(self, timeout_ms=0xFFFFFFFF):
        return _ck(_L.atray_multi_render_poll(self.h, timeout_ms))

    def render_progress(self):
        a, b = C.c_int32(0), C.c_int32(0)
        _ck(_L.atray_multi_render_progress(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def cancel(self):
        return _ck(_L.atray_multi_render_cancel(self.h)) == 1

    def render(self, cam, opts=None, want=("bgrx", "t", "prim")):
        opts = opts or self.opts()
        H, W = cam.height, cam.width
        out = {"bgrx": np.zeros((H, W), np.uint32) if "bgrx" in want else None,
               "t": np.zeros((H, W), np.float32) if "t" in want else None,
               "prim": np.zeros((H, W), np.uint32) if "prim" in want else None}
        self.render_begin(cam, opts, out["bgrx"], out["t"], out["prim"])
        while self.render_poll(1000):
            pass
        return out

    def render_device(self, cam, opts, d_bgrx=0, d_t=0, d_prim=0):
        """Planes on one device (normally ctx[0].dev_alloc): the other GPUs store their bands into them over NVLink."""
        fr = api.Frame(d_bgrx or None, d_t or None, d_prim or None)
        _ck(_L.atray_multi_render_device(self.h, C.byref(cam), C.byref(opts), C.byref(fr)))

    def sync(self):
        _ck(_L.atray_multi_sync(self.h))

    def last_ray_casts(self):
        return sum(c.last_ray_casts() for c in self.ctx)


class Scene:
    """Scene (scene.h:13-24) restricted to what the path uses: materials[0] = sky, one model with its material."""

    def __init__(self, model, sky=api.SKY_MATERIAL, material=api.MODEL_MATERIAL, division_method=api.DIV_SAH,
                 max_no_faces_per_node=300):
        self.model, self.sky, self.material = model, sky, material
        self.division_method, self.max_no_faces_per_node = division_method, max_no_faces_per_node
        self.spheres, self.planes = [], []   # scene.spheres / scene.planes: (center, radius, material) / (normal, distance, material)
        self.ctx = None
        self.gpu_devices = None   # e.g. [0, 4]: prep_scene creates a MultiContext over these GPUs (None: one GPU, `device`)


class RenderInfo:
    """RenderInfo (renderer.h:17-30): camera, scene, camera_tex (host BGRX bitmap) + the two extra planes."""

    def __init__(self, camera, scene, camera_tex, t=None, prim=None, opts=None):
        self.camera, self.scene, self.camera_tex, self.t, self.prim = camera, scene, camera_tex, t, prim
        self.opts = opts
        self.jobs_done, self.jobs_size = 0, 0
        self.total_ray_casts = 0


def prep_scene(scene, device=0):
    """prep_scene(Scene&, uint32&) — renderer.cpp:264-291: build the tree if the model has none, then (new) flatten and
    upload once. Returns the node count (kd_tree_max_nodes)."""
    n = scene.model.counts()
    if n["nodes"] == 0:
        api.build_KD_tree(scene.model, scene.division_method, scene.max_no_faces_per_node)
        n = scene.model.counts()
    if scene.ctx is None:
        scene.ctx = MultiContext(list(scene.gpu_devices)) if scene.gpu_devices and len(scene.gpu_devices) > 1 else \
            Context(scene.gpu_devices[0] if scene.gpu_devices else device)
    scene.ctx.upload(scene.model, scene.sky, scene.material)
    if scene.spheres or scene.planes:
        scene.ctx.set_primitives(scene.spheres, scene.planes)
    return n["nodes"]


def start_render_from_camera(info, tpool=None):
    """start_render_from_camera(RenderInfo&, ThreadPool&) — renderer.cpp:403-455. Returns immediately; `tpool` is
    accepted for signature parity and ignored (the tile thread pool is replaced by the kernel)."""
    opts = info.opts or Context.opts()
    if info.scene.ctx.render_poll(0):       # a frame is still in flight on this scene: the new one replaces it
        info.scene.ctx.cancel()
    # jobs = bands of 8 scanlines (the reference lists square tiles, renderer.cpp:420-445; callers only show k/n)
    info.jobs_size, info.jobs_done = (info.camera.height + 7) // 8, 0
    info.scene.ctx.render_begin(info.camera, opts, info.camera_tex, info.t, info.prim)


def clear_work_queue(info):
    """info.twq.clear() — work_queue.h:11-15, the cancel of app.cpp:165-172: stops the frame in flight (the kernels' persistent
    loops poll the context's cancel word) and returns when the GPU has drained. True when a frame was cut short."""
    info.jobs_size = 0
    return info.scene.ctx.cancel()


def activate_pool(tpool, proc, data):
    """thread_pool.h:14-20: one thread per pool slot on proc(data). `tpool` is a list the started threads are kept in."""
    import threading
    n = max(1, len(tpool))
    tpool[:] = [threading.Thread(target=proc, args=(data,)) for _ in range(n)]
    for th in tpool:
        th.start()


def wait_for_pool(tpool, time_out_in_ms=0xFFFFFFFF):
    """thread_pool.h:22-26: TRUE (1) when the wait timed out, FALSE (0) when every thread has finished. The renderer never
    activates the pool, so after clear_work_queue() this returns 0 at once (app.cpp:169)."""
    import time
    deadline = None if time_out_in_ms == 0xFFFFFFFF else time.monotonic() + time_out_in_ms / 1e3
    for th in tpool or []:
        if not hasattr(th, "join"):
            continue
        th.join(None if deadline is None else max(0.0, deadline - time.monotonic()))
        if th.is_alive():
            return 1
    return 0


def get_ray_kd_tree_intersection(op_ray, scene, tri_data=None):
    """get_ray_kd_tree_intersection (kd_tree.h:56, kd_tree.cpp:302-313) for one ray: op_ray = (origin3, direction3); returns the
    distance (MAX_FLOAT on a miss) and fills tri_data = {"face_index", "u", "v"} on a hit only."""
    ctx = scene.ctx.ctx[0] if isinstance(scene.ctx, MultiContext) else scene.ctx
    prim, t, uv = ctx.trace_rays_uv(np.asarray(op_ray, np.float32).reshape(1, 6))
    if tri_data is not None and prim[0] != api.MISS:
        tri_data.update(face_index=int(prim[0]), u=float(uv[0, 0]), v=float(uv[0, 1]))
    return float(t[0])


def wait_for_render_from_camera_to_finish(info, tpool=None, ms_to_wait_for=33):
    """renderer.cpp:457-471: TRUE (1) while still running, FALSE (0) when finished."""
    running = info.scene.ctx.render_poll(ms_to_wait_for)
    if running:
        info.jobs_done = info.scene.ctx.render_progress()[0] // 8   # bands whose scanlines have landed in camera_tex
    if not running:
        info.jobs_done = info.jobs_size
        # ray_casts counts non-sky bounce iterations (renderer.cpp:222-231,260): at bounce 1 = samples that hit geometry
        if info.camera.bounce_limit != 1 or info.scene.spheres or info.scene.planes:
            info.total_ray_casts = info.scene.ctx.last_ray_casts()
        elif info.prim is not None and info.camera.samples_per_pixel == 1:
            info.total_ray_casts = int((info.prim != api.MISS).sum())
    return running
