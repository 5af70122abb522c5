"""ctypes binding of libatray_b200.so (include/atray_b200.h) with a thin host-side mirror of the reference's
interface for the render path: names and argument meaning follow /root/reference/Source/engine
(load_model_data, get_AABB, resize_scale, translate_to, build_KD_tree, set_camera, prep_scene,
start_render_from_camera, wait_for_render_from_camera_to_finish — see app.cpp:62-187 for the call order).

Python is test/bench orchestration only; the product is the C-ABI library. There is no CPU fallback: if the library
is missing this module raises at import, and every device call raises AtrayError when CUDA is unavailable."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ATRAY_B200_LIB") or os.path.join(HERE, "libatray_b200.so")  # override: A/B builds of the same ABI

MISS = 0xFFFFFFFF
MAX_FLOAT = np.float32(3.402823466e+38)
DIV_CENTER, DIV_SAH = 0, 1
FP_STRICT, FP_FAST = 0, 1
RENDER_NO_CULL = 1
RENDER_RASTER = 2

# app.cpp:88,91,96
DEFAULT_EYE = (0.1, 2.0, 0.0)
DEFAULT_FACING = (-0.1, -0.5, -1.0)
DEFAULT_HFOV = 1.0
SKY_MATERIAL = ((0.3, 0.4, 0.5), (0.2, 0.3, 0.4), 0.3)
MODEL_MATERIAL = ((0.4, 0.2, 0.2), (0.92, 0.5, 0.0), 0.3)
DEFAULT_SEED = 0x853C49E6748FEA9B


class AtrayError(RuntimeError):
    pass


class Material(C.Structure):
    _fields_ = [("emission_color", C.c_float * 3), ("reflection_color", C.c_float * 3), ("scatter", C.c_float)]


class Sphere(C.Structure):   # AtraySphere (sphere.h:5-10 with the material by value)
    _fields_ = [("center", C.c_float * 3), ("radius", C.c_float), ("material", Material)]


class Plane(C.Structure):    # AtrayPlane (plane.h:5-10)
    _fields_ = [("normal", C.c_float * 3), ("distance", C.c_float), ("material", Material)]


class CameraStruct(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("anti_aliasing", C.c_int32),
                ("samples_per_pixel", C.c_uint32), ("bounce_limit", C.c_int32), ("aspect_ratio", C.c_float),
                ("camera_z", C.c_float * 3), ("camera_x", C.c_float * 3), ("camera_y", C.c_float * 3),
                ("eye", C.c_float * 3), ("frame_center", C.c_float * 3), ("h_fov", C.c_float),
                ("half_pixel_width", C.c_float), ("half_pixel_height", C.c_float)]


class Partition(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("band_rows", C.c_int32)]


class RenderOpts(C.Structure):
    _fields_ = [("fp_mode", C.c_int32), ("count_work", C.c_int32), ("jitter_seed", C.c_uint64), ("part", Partition),
                ("flags", C.c_uint32)]


class Frame(C.Structure):
    _fields_ = [("bgrx", C.c_void_p), ("t", C.c_void_p), ("prim", C.c_void_p)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "box_tests", "tri_tests", "pass_det", "pass_u", "pass_v",
                                          "leaves_listed", "leaves_scanned", "overflow_rays", "rect_tests", "tiles_sky",
                                          "tiles_busy")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


if not os.path.exists(LIB_PATH):
    raise ImportError("atray_b200: %s is missing — run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback)" % LIB_PATH)
_L = C.CDLL(LIB_PATH)
_L.atray_last_error.restype = C.c_char_p

_vp, _i, _u, _f, _u64, _sz = C.c_void_p, C.c_int32, C.c_uint32, C.c_float, C.c_uint64, C.c_size_t
_SIGS = {
    "atray_version": [],
    "atray_model_load_obj": [C.c_char_p, _i, C.POINTER(_vp)],
    "atray_model_from_arrays": [_vp, _u, _vp, _u, C.POINTER(_vp)],
    "atray_model_compute_aabb": [_vp, _vp],
    "atray_model_resize_scale": [_vp, _f],
    "atray_model_translate_to": [_vp, _vp],
    "atray_model_build_tree": [_vp, _i, _u],
    "atray_model_counts": [_vp, _vp],
    "atray_model_dump": [_vp] * 8,
    "atray_model_tree_dump": [_vp] * 7,
    "atray_model_flatten": [_vp, _vp, _vp, _vp, _vp],
    "atray_write_bmp": [_vp, C.c_uint32, C.c_uint32, C.c_char_p, C.c_char_p, C.c_size_t],
    "atray_debug_film_rects": [_vp, C.POINTER(CameraStruct), _vp, _vp, _vp, _vp, _vp, _vp],
    "atray_camera_set": [C.POINTER(CameraStruct), _i, _i, _i, _u, _i, _vp, _vp, _f],
    "atray_device_count": [C.POINTER(_i)],
    "atray_ctx_create": [_i, C.POINTER(_vp)],
    "atray_scene_upload": [_vp, _vp, C.POINTER(Material), C.POINTER(Material)],
    "atray_scene_set_primitives": [_vp, _vp, _u, _vp, _u],
    "atray_scene_use_kd_tree": [_vp, _i],
    "atray_last_ray_casts": [_vp, C.POINTER(_u64)],
    "atray_debug_tile_log": [_vp, _vp, C.c_size_t],
    "atray_render_begin": [_vp, C.POINTER(CameraStruct), C.POINTER(RenderOpts), C.POINTER(Frame)],
    "atray_render_poll": [_vp, _u],
    "atray_render_progress": [_vp, C.POINTER(_i), C.POINTER(_i)],
    "atray_render_cancel": [_vp],
    "atray_render_device": [_vp, C.POINTER(CameraStruct), C.POINTER(RenderOpts), C.POINTER(Frame)],
    "atray_ctx_sync": [_vp],
    "atray_last_kernel_ms": [_vp, C.POINTER(_f)],
    "atray_last_render_kernel_ms": [_vp, C.POINTER(_f)],
    "atray_last_view_split_ms": [_vp, C.POINTER(_f), C.POINTER(_f)],
    "atray_trace_rays": [_vp, _vp, _u64, _i, _i, _vp, _vp],
    "atray_trace_rays_device": [_vp, _vp, _u64, _i, _i, _vp, _vp],
    "atray_trace_rays_static": [_vp, _vp, _u64, _i, _vp, _vp],
    "atray_trace_rays_uv": [_vp, _vp, _u64, _vp, _vp, _vp],
    "atray_get_counters": [_vp, C.POINTER(Counters)],
    "atray_trace_incoherent_device": [_vp, _u64, _u64, _u, _u64, C.POINTER(Partition), _u, _i, _i, _vp, _vp, _vp],
    "atray_trace_incoherent": [_vp, _u64, _u64, _u, _u64, _i, _i, _vp, _vp, _vp],
    "atray_dev_alloc": [_vp, _sz, C.POINTER(_vp)],
    "atray_dev_free": [_vp, _vp],
    "atray_dev_copy_to_host": [_vp, _vp, _vp, _sz],
    "atray_dev_copy_from_host": [_vp, _vp, _vp, _sz],
    "atray_dev_memset": [_vp, _vp, C.c_int, _sz],
    "atray_host_alloc_pinned": [_sz, C.POINTER(_vp)],
    "atray_host_free_pinned": [_vp],
    "atray_host_register": [_vp, _sz],
    "atray_host_unregister": [_vp],
    "atray_ipc_export": [_vp, _vp, _vp],
    "atray_ipc_open": [_vp, _vp, C.POINTER(_vp)],
    "atray_ipc_close": [_vp, _vp],
    "atray_microbench": [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)],
    "atray_view_share_create": [_vp, _i, _i, C.POINTER(_vp), C.POINTER(_sz)],
    "atray_view_share_connect": [_vp, _i, _i, _vp],
    "atray_view_share_off": [_vp],
    "atray_view_share_rendezvous": [_vp],
    "atray_view_share_check": [_vp],
    "atray_ctx_enable_peer": [_vp, _vp],
    "atray_debug_view_shards": [_vp, _i],
    "atray_ctx_create_multi": [_vp, _i, C.POINTER(_vp)],
    "atray_multi_count": [_vp],
    "atray_multi_device": [_vp, _i],
    "atray_multi_scene_upload": [_vp, _vp, C.POINTER(Material), C.POINTER(Material)],
    "atray_multi_scene_set_primitives": [_vp, _vp, _u, _vp, _u],
    "atray_multi_scene_use_kd_tree": [_vp, _i],
    "atray_multi_render_begin": [_vp, C.POINTER(CameraStruct), C.POINTER(RenderOpts), C.POINTER(Frame)],
    "atray_multi_render_poll": [_vp, _u],
    "atray_multi_render_progress": [_vp, C.POINTER(_i), C.POINTER(_i)],
    "atray_multi_render_cancel": [_vp],
    "atray_multi_render_device": [_vp, C.POINTER(CameraStruct), C.POINTER(RenderOpts), C.POINTER(Frame)],
    "atray_multi_sync": [_vp],
}
for _n, _a in _SIGS.items():
    getattr(_L, _n).argtypes = _a
    getattr(_L, _n).restype = C.c_int
_L.atray_model_free.argtypes = [_vp]
_L.atray_model_free.restype = None
_L.atray_ctx_destroy.argtypes = [_vp]
_L.atray_ctx_destroy.restype = None
_L.atray_multi_destroy.argtypes = [_vp]
_L.atray_multi_destroy.restype = None
_L.atray_multi_ctx.argtypes = [_vp, _i]
_L.atray_multi_ctx.restype = _vp

EXPORTED_SYMBOLS = sorted(list(_SIGS) + ["atray_model_free", "atray_ctx_destroy", "atray_last_error", "atray_multi_destroy", "atray_multi_ctx"])


def _ck(rc):
    if rc < 0:
        raise AtrayError("atray_b200 error %d: %s" % (rc, (_L.atray_last_error() or b"").decode()))
    return rc


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f3(v):
    return np.asarray(v, dtype=np.float32).copy()


def make_material(m):
    return Material((C.c_float * 3)(*m[0]), (C.c_float * 3)(*m[1]), m[2])


class Model:
    """Host model: ModelData + surrounding_aabb + KD_Tree (model.h:30-71)."""

    def __init__(self, handle):
        self.h = handle

    def __del__(self):
        self.close()

    def close(self):
        if getattr(self, "h", None):
            _L.atray_model_free(self.h)
            self.h = None

    def counts(self):
        c = np.zeros(8, np.uint32)
        _ck(_L.atray_model_counts(self.h, _p(c)))
        return dict(zip(["vertices", "normals", "tex_coords", "faces", "nodes", "refs", "inner", "max_leaf"], map(int, c)))

    def dump(self):
        n = self.counts()
        v = np.zeros((n["vertices"], 3), np.float32)
        nr = np.zeros((n["normals"], 3), np.float32)
        t = np.zeros((n["tex_coords"], 3), np.float32)
        fv = np.zeros((n["faces"], 3), np.int32)
        fvt = np.zeros((n["faces"], 3), np.int32)
        fvn = np.zeros((n["faces"], 3), np.int32)
        box = np.zeros(6, np.float32)
        _ck(_L.atray_model_dump(self.h, _p(v), _p(nr), _p(t), _p(fv), _p(fvt), _p(fvn), _p(box)))
        return dict(vertices=v, normals=nr, tex_coords=t, face_v=fv, face_vt=fvt, face_vn=fvn, aabb=box)

    def tree(self):
        n = self.counts()
        boxes = np.zeros((n["nodes"], 6), np.float32)
        child = np.zeros(n["nodes"], np.int32)
        cnt = np.zeros(n["nodes"], np.uint32)
        first = np.zeros(n["nodes"], np.uint32)
        tri = np.zeros((n["refs"], 9), np.float32)
        face = np.zeros(n["refs"], np.uint32)
        _ck(_L.atray_model_tree_dump(self.h, _p(boxes), _p(child), _p(cnt), _p(first), _p(tri), _p(face)))
        return dict(boxes=boxes, children_start=child, prim_count=cnt, prim_first=first, tri=tri, face=face)

    def flatten(self):
        nn, ns = C.c_uint32(0), C.c_uint32(0)
        _ck(_L.atray_model_flatten(self.h, C.byref(nn), C.byref(ns), None, None))
        nodes = np.zeros((nn.value, 2, 4), np.float32)
        tris = np.zeros((ns.value // 8, 3, 8, 4), np.float32)
        _ck(_L.atray_model_flatten(self.h, None, None, _p(nodes), _p(tris)))
        return nodes, tris

    def film_rects(self, cam, hulls=False, px=False):
        """Host evaluation of the per-frame film rectangles (view_bounds.h): rects n_slots×4, owning leaf, position in
        the reference's leaf order, busy flag per 8x4 tile; hulls=True adds the n_slots×3×3 half-planes {a, b, c}, px=True the
        n_slots×4 inclusive pixel ranges {x0, x1, y0, y1} the kernels use."""
        nn, ns = C.c_uint32(0), C.c_uint32(0)
        _ck(_L.atray_model_flatten(self.h, C.byref(nn), C.byref(ns), None, None))
        rects = np.zeros((ns.value, 4), np.float32)
        leaf = np.zeros(ns.value, np.uint32)
        pos = np.zeros(ns.value, np.uint32)
        busy = np.zeros(((cam.height + 3) // 4, (cam.width + 7) // 8), np.uint8)
        hl = np.zeros((ns.value, 3, 3), np.float32) if hulls else None
        pr = np.zeros((ns.value, 4), np.int32) if px else None
        _ck(_L.atray_debug_film_rects(self.h, C.byref(cam), _p(rects), _p(leaf), _p(pos), _p(busy), _p(hl) if hulls else None,
                                      _p(pr) if px else None))
        out = [rects, leaf, pos, busy]
        if hulls:
            out.append(hl)
        if px:
            out.append(pr)
        return tuple(out)


# ---- reference-named host functions (app.cpp:62-156 call order) ----------------------------------------------------
def load_model_data(file_name, threads=1):
    """load_model_data(ModelData&, const char*, ThreadPool&) — OBJ_loader.cpp:278-361; `threads` = pool size."""
    h = _vp()
    _ck(_L.atray_model_load_obj(os.fsencode(file_name), threads, C.byref(h)))
    return Model(h)


def model_from_arrays(vertices, faces):
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
    f = np.ascontiguousarray(faces, np.int32).reshape(-1, 3)
    h = _vp()
    _ck(_L.atray_model_from_arrays(_p(v), v.shape[0], _p(f), f.shape[0], C.byref(h)))
    return Model(h)


def get_AABB(model):
    """model.surrounding_aabb = get_AABB(model.data) — model.h:41-61."""
    box = np.zeros(6, np.float32)
    _ck(_L.atray_model_compute_aabb(model.h, _p(box)))
    return box


def resize_scale(model, new_max_scale):
    _ck(_L.atray_model_resize_scale(model.h, new_max_scale))


def translate_to(model, new_center):
    c = _f3(new_center)
    _ck(_L.atray_model_translate_to(model.h, _p(c)))


def build_KD_tree(model, division_method=DIV_SAH, max_no_faces_per_node=300):
    _ck(_L.atray_model_build_tree(model.h, division_method, max_no_faces_per_node))


def set_camera(width, height, eye=DEFAULT_EYE, facing=DEFAULT_FACING, h_fov=DEFAULT_HFOV, anti_aliasing=0,
               samples_per_pixel=1, bounce_limit=1):
    """set_camera(Camera&, eye, facing, RenderSettings&, h_fov) — camera.h:36-45."""
    cam = CameraStruct()
    e, f = _f3(eye), _f3(facing)
    _ck(_L.atray_camera_set(C.byref(cam), width, height, anti_aliasing, samples_per_pixel, bounce_limit, _p(e), _p(f), h_fov))
    return cam


def update_camera_pos(cam, eye, facing):
    """update_camera_pos(Camera&, eye, facing_towards) — camera.h:23-34: a new position and orientation for an existing camera;
    resolution, samples, bounces and h_fov stay as set_camera left them. In place; returns the camera."""
    e, f = _f3(eye), _f3(facing)
    _ck(_L.atray_camera_set(C.byref(cam), cam.width, cam.height, cam.anti_aliasing, cam.samples_per_pixel, cam.bounce_limit, _p(e), _p(f),
                            cam.h_fov))
    return cam


def camera_as_array(cam):
    """19 floats in the reference Camera layout after render_settings (camera.h:11-20)."""
    return np.array([cam.aspect_ratio, *cam.camera_z, *cam.camera_x, *cam.camera_y, *cam.eye, *cam.frame_center,
                     cam.h_fov, cam.half_pixel_width, cam.half_pixel_height], dtype=np.float32)


def place_default(model, rescale=40.0, translate=(0.0, -15.0, -38.0)):
    """Placement used for every benchmark mesh: get_AABB → resize_scale(40) → translate_to(0,-15,-38);
    rescale <= 0 skips the resize as app.cpp:69-72 does for the dragon."""
    get_AABB(model)
    if rescale and rescale > 0:
        resize_scale(model, rescale)
    translate_to(model, translate)


def Write_To_File(bgrx, file_name):
    """Write_To_File(Texture&, const char*) — texture.cpp:66-115, 147-158: byte-identical BMP of a (H, W) uint32 BGRX bitmap
    into "<file_name>_<id>.bmp" (first id not on disk yet). Returns the name written."""
    a = np.ascontiguousarray(bgrx, np.uint32)
    H, W = a.shape
    out = C.create_string_buffer(1100)
    _ck(_L.atray_write_bmp(_p(a), W, H, os.fsencode(file_name), out, 1100))
    return os.fsdecode(out.value)


def device_count():
    n = _i(0)
    rc = _L.atray_device_count(C.byref(n))
    return n.value if rc == 0 else 0
