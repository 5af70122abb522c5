"""TEST INFRASTRUCTURE ONLY: one ctypes binding for both CPU checkers, which export the same C shapes:

* ``ref()``  — oracle/_ref/libatray_ref.so: the reference's own hot-path sources (/root/reference/Source, built by
  oracle/ref/build_ref.sh) behind oracle/ref/ref_harness.cpp (prefix ``ref_``);
* ``port()`` — oracle/liboracle.so: the plain-C restatement oracle/atray_oracle.c (prefix ``or_``).

Only tests/, __graft_entry__.smoke() and bench.py may import this module; the product (atray_b200/) never does."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libatray_ref.so")
ASSETS = os.path.join(HERE, "_ref", "assets")
MISS = 0xFFFFFFFF

DEFAULT_EYE = (0.1, 2.0, 0.0)       # app.cpp:88
DEFAULT_FACING = (-0.1, -0.5, -1.0)
DEFAULT_HFOV = 1.0


PORT_PATH = os.path.join(HERE, "liboracle.so")
NOTREE_PATH = os.path.join(HERE, "_ref", "libatray_ref_notree.so")


class Backend:
    def __init__(self, path, prefix):
        self.path, self.prefix = path, prefix
        L = C.CDLL(path)
        g = lambda n: getattr(L, prefix + n)
        self.scene_create = g("scene_create")
        self.scene_create.restype = C.c_void_p
        self.scene_create.argtypes = [C.c_char_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_uint]
        self.scene_destroy = g("scene_destroy"); self.scene_destroy.argtypes = [C.c_void_p]
        self.counts = g("counts"); self.counts.argtypes = [C.c_void_p, C.c_void_p]
        self.model_dump = g("model_dump"); self.model_dump.argtypes = [C.c_void_p] + [C.c_void_p] * 7
        self.tree_dump = g("tree_dump"); self.tree_dump.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        self.camera = g("camera"); self.camera.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        self.trace_primary = g("trace_primary")
        self.trace_primary.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_int,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        self.trace_rays = g("trace_rays")
        self.trace_rays.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        self.trace_rays_uv = g("trace_rays_uv")
        self.trace_rays_uv.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        self.render_jitter = g("render_jitter")
        self.render_jitter.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_int, C.c_ulonglong, C.c_void_p,
                                       C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        self.pcg32 = g("pcg32"); self.pcg32.argtypes = [C.c_ulonglong, C.c_ulonglong, C.c_int, C.c_void_p]
        self.scene_add_sphere = g("scene_add_sphere")
        self.scene_add_sphere.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        self.scene_add_plane = g("scene_add_plane")
        self.scene_add_plane.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        self.render_path = g("render_path")
        self.render_path.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_int, C.c_int, C.c_ulonglong, C.c_void_p,
                                     C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p]
        self.write_bmp = g("write_bmp"); self.write_bmp.restype = C.c_int
        self.write_bmp.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_char_p]
        self.render = None
        self.trace_primary_ex = None
        if prefix == "ref_":
            self.render = g("render")
            self.render.restype = C.c_double
            self.render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                    C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
        else:
            self.trace_primary_ex = g("trace_primary_ex")
            self.trace_primary_ex.argtypes = self.trace_primary.argtypes + [C.c_void_p]
            self.scene_use_kd_tree = g("scene_use_kd_tree"); self.scene_use_kd_tree.argtypes = [C.c_void_p, C.c_int]
            self.trace_c5 = g("trace_c5")
            self.trace_c5.argtypes = [C.c_void_p, C.c_longlong, C.c_longlong, C.c_uint, C.c_ulonglong, C.c_int, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p]
            self.accepted_hits = g("accepted_hits")
            self.accepted_hits.restype = C.c_longlong
            self.accepted_hits.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_int,
                                           C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p]


_backends = {}


def ref_available():
    return os.path.exists(LIB_PATH)


def port_available():
    return os.path.exists(PORT_PATH)


def ref():
    if "ref" not in _backends:
        _backends["ref"] = Backend(LIB_PATH, "ref_")
    return _backends["ref"]


def ref_notree():
    """The reference built without `#define USE_KD_TREE` (renderer.h:8): get_intersection_data's brute-force branch."""
    if "ref_notree" not in _backends:
        _backends["ref_notree"] = Backend(NOTREE_PATH, "ref_")
    return _backends["ref_notree"]


def ref_notree_available():
    return os.path.exists(NOTREE_PATH)


def port():
    if "port" not in _backends:
        _backends["port"] = Backend(PORT_PATH, "or_")
    return _backends["port"]


def _f3(v):
    return np.asarray(v, dtype=np.float32).copy()


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Scene:
    """One model loaded, placed and tree-built by a CPU checker backend (ref() or port())."""

    def __init__(self, backend, obj_path, loader_threads=1, rescale=40.0, translate=(0.0, -15.0, -38.0), method=1, cap=300):
        self.b = backend
        self.h = backend.scene_create(os.fsencode(obj_path), loader_threads, rescale, translate[0], translate[1],
                                        translate[2], method, cap)
        if not self.h:
            raise RuntimeError("oracle scene_create failed for %s (missing file or degenerate tree)" % obj_path)
        c = np.zeros(8, dtype=np.uint32)
        self.b.counts(self.h, _p(c))
        (self.n_vertices, self.n_normals, self.n_tex, self.n_faces, self.n_nodes, self.n_refs, self.n_inner,
         self.max_leaf) = [int(x) for x in c]

    def use_kd_tree(self, use):
        """port() only: 0 → the no-tree branch of get_intersection_data (renderer.cpp:58-82)."""
        self.b.scene_use_kd_tree(self.h, 1 if use else 0)

    def close(self):
        if self.h:
            self.b.scene_destroy(self.h)
            self.h = None

    def model(self):
        v = np.zeros((self.n_vertices, 3), np.float32)
        n = np.zeros((self.n_normals, 3), np.float32)
        t = np.zeros((self.n_tex, 3), np.float32)
        fv = np.zeros((self.n_faces, 3), np.int32)
        fvt = np.zeros((self.n_faces, 3), np.int32)
        fvn = np.zeros((self.n_faces, 3), np.int32)
        box = np.zeros(6, np.float32)
        self.b.model_dump(self.h, _p(v), _p(n), _p(t), _p(fv), _p(fvt), _p(fvn), _p(box))
        return dict(vertices=v, normals=n, tex_coords=t, face_v=fv, face_vt=fvt, face_vn=fvn, aabb=box)

    def tree(self):
        boxes = np.zeros((self.n_nodes, 6), np.float32)
        child = np.zeros(self.n_nodes, np.int32)
        cnt = np.zeros(self.n_nodes, np.uint32)
        first = np.zeros(self.n_nodes, np.uint32)
        tri = np.zeros((self.n_refs, 9), np.float32)
        face = np.zeros(self.n_refs, np.uint32)
        self.b.tree_dump(self.h, _p(boxes), _p(child), _p(cnt), _p(first), _p(tri), _p(face))
        return dict(boxes=boxes, children_start=child, prim_count=cnt, prim_first=first, tri=tri, face=face)

    def render(self, W, H, spp=1, aa=0, bounce=1, threads=1, eye=DEFAULT_EYE, facing=DEFAULT_FACING, hfov=DEFAULT_HFOV):
        out = np.zeros((H, W), np.uint32)
        casts = C.c_longlong(0)
        e, f = _f3(eye), _f3(facing)
        sec = self.b.render(self.h, W, H, spp, aa, bounce, threads, _p(e), _p(f), hfov, _p(out), C.byref(casts))
        return out, sec, casts.value

    def trace_primary(self, W, H, threads=1, eye=DEFAULT_EYE, facing=DEFAULT_FACING, hfov=DEFAULT_HFOV, want_uv=False):
        face = np.zeros((H, W), np.uint32)
        t = np.zeros((H, W), np.float32)
        uv = np.zeros((H, W, 2), np.float32) if want_uv else None
        img = np.zeros((H, W), np.uint32)
        e, f = _f3(eye), _f3(facing)
        self.b.trace_primary(self.h, W, H, _p(e), _p(f), hfov, threads, _p(face), _p(t), _p(uv), _p(img))
        return (face, t, img, uv) if want_uv else (face, t, img)

    def primary_counters(self, W, H, threads=1, eye=DEFAULT_EYE, facing=DEFAULT_FACING, hfov=DEFAULT_HFOV):
        """port() only: work counters of one primary-ray frame (box tests, triangle tests, ... per frame)."""
        ct = np.zeros(8, np.uint64)
        e, f = _f3(eye), _f3(facing)
        self.b.trace_primary_ex(self.h, W, H, _p(e), _p(f), hfov, threads, None, None, None, None, _p(ct))
        names = ["box_tests", "tri_tests", "pass_det", "pass_u", "pass_v", "inner_pops", "leaves_listed", "leaves_scanned"]
        return dict(zip(names, [int(x) for x in ct]))

    def accepted_hits(self, W, H, threads=1, eye=DEFAULT_EYE, facing=DEFAULT_FACING, hfov=DEFAULT_HFOV, cap=None):
        """port() only: every (pixel, leaf node, position in leaf) the triangle test accepts in any listed leaf."""
        cap = cap or 8 * W * H
        e, f = _f3(eye), _f3(facing)
        for _ in range(2):   # the call returns the number of hits even when they did not fit: second pass with that size
            px, node, pos = np.zeros(cap, np.uint32), np.zeros(cap, np.uint32), np.zeros(cap, np.uint32)
            n = self.b.accepted_hits(self.h, W, H, _p(e), _p(f), hfov, threads, cap, _p(px), _p(node), _p(pos))
            if n <= cap:
                break
            cap = int(n)
        assert n <= cap, "accepted_hits: capacity too small (%d > %d)" % (n, cap)
        return px[:n], node[:n], pos[:n]

    def trace_c5(self, p0, p1, spp, seed, threads=1, want_rays=False):
        """port() only: C5 (incoherent rays, `spp` per pixel index in [p0, p1)) — per ray face / t (and the rays), per pixel BGRX."""
        n = (p1 - p0) * spp
        face, t = np.zeros(n, np.uint32), np.zeros(n, np.float32)
        bgrx = np.zeros(p1 - p0, np.uint32)
        rays = np.zeros((n, 6), np.float32) if want_rays else None
        self.b.trace_c5(self.h, p0, p1, spp, seed, threads, _p(face), _p(t), _p(bgrx), _p(rays) if want_rays else None)
        return (face, t, bgrx, rays) if want_rays else (face, t, bgrx)

    def trace_rays(self, rays, threads=1):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = rays.shape[0]
        face = np.zeros(n, np.uint32)
        t = np.zeros(n, np.float32)
        self.b.trace_rays(self.h, n, _p(rays), threads, _p(face), _p(t))
        return face, t

    def trace_rays_uv(self, rays, threads=1):
        """trace_rays plus TriangleIntersectionData's (u, v) of the winning triangle; rows of misses stay NaN."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = rays.shape[0]
        face = np.zeros(n, np.uint32)
        t = np.zeros(n, np.float32)
        uv = np.full((n, 2), np.nan, np.float32)
        self.b.trace_rays_uv(self.h, n, _p(rays), threads, _p(face), _p(t), _p(uv))
        return face, t, uv

    def render_jitter(self, W, H, spp, seed, threads=1, bounce=1, eye=DEFAULT_EYE, facing=DEFAULT_FACING, hfov=DEFAULT_HFOV):
        img = np.zeros((H, W), np.uint32)
        hits = np.zeros((H, W), np.uint32)
        e, f = _f3(eye), _f3(facing)
        self.b.render_jitter(self.h, W, H, spp, bounce, seed, _p(e), _p(f), hfov, threads, _p(img), _p(hits))
        return img, hits


    # ---- multi-bounce path: spheres, planes, cast_ray at any bounce_limit -------------------
    def add_sphere(self, center, radius, material):
        """material = (emission rgb, reflection rgb, scatter) as material.h:4-9."""
        c = _f3(center)
        m = np.array([*material[0], *material[1], material[2]], np.float32)
        assert self.b.scene_add_sphere(self.h, _p(c), radius, _p(m)) == 0

    def add_plane(self, normal, distance, material):
        n = _f3(normal)
        m = np.array([*material[0], *material[1], material[2]], np.float32)
        assert self.b.scene_add_plane(self.h, _p(n), distance, _p(m)) == 0

    def render_path(self, W, H, spp, bounce, seed, aa=0, threads=1, eye=DEFAULT_EYE, facing=DEFAULT_FACING, hfov=DEFAULT_HFOV):
        """Pixel loop + cast_ray with the per-pixel stream {splitmix64(seed ^ p), 2p+1}. Returns bgrx, t, prim, kind
        (nearest hit of the last sample's primary ray; kind 0 sky / 1 triangle / 2 sphere / 3 plane) and the total
        ray_casts (sum of bounce iterations, renderer.cpp:260)."""
        img = np.zeros((H, W), np.uint32)
        t = np.zeros((H, W), np.float32)
        prim = np.zeros((H, W), np.uint32)
        kind = np.zeros((H, W), np.uint32)
        casts = C.c_longlong(0)
        e, f = _f3(eye), _f3(facing)
        self.b.render_path(self.h, W, H, spp, aa, bounce, seed, _p(e), _p(f), hfov, threads, _p(img), _p(t), _p(prim),
                           _p(kind), C.byref(casts))
        return img, t, prim, kind, casts.value


def camera(backend, W, H, eye=DEFAULT_EYE, facing=DEFAULT_FACING, hfov=DEFAULT_HFOV):
    out = np.zeros(19, np.float32)
    e, f = _f3(eye), _f3(facing)
    backend.camera(W, H, _p(e), _p(f), hfov, _p(out))
    return out


def write_bmp(backend, bgrx, file_name):
    """Write_To_File on a (H, W) uint32 bitmap: "<file_name>_<id>.bmp" (texture.cpp:66-115)."""
    a = np.ascontiguousarray(bgrx, np.uint32)
    return backend.write_bmp(_p(a), a.shape[1], a.shape[0], os.fsencode(file_name))


def pcg32(backend, state, stream, n):
    out = np.zeros(n, np.uint32)
    backend.pcg32(state, stream, n, _p(out))
    return out
