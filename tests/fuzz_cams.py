"""Seeded camera / mesh generators of the culling fuzz tests (tests only).

The film-space culling (atray_b200/csrc/view_bounds.h) is sound iff for every camera the culled frame equals the unculled
one bit for bit. These generators aim at the places where a float-error margin could be too small: an eye inside the model
box, on a leaf face, grazing a face, far away; h_fov from 0.01 to 3; facing almost parallel to `up` (camera.h:23-34 crosses
with (0,1,0)); frame sizes that are no multiples of the 8x4 tile; slivers, zero-area triangles and triangles that cross
the eye plane."""
import numpy as np


def sliver_mesh(seed=7, n=1600):
    """Triangles chosen to stress the bounds: ordinary, slivers (two nearly coincident vertices / nearly collinear), zero-area
    (repeated vertex, collinear), tiny, and huge ones (they cross the eye plane of any eye inside the box)."""
    rng = np.random.default_rng(seed)
    v, f = [], []

    def add(a, b, c):
        k = len(v)
        v.extend([a, b, c])
        f.append((k, k + 1, k + 2))

    for i in range(n):
        kind = i % 8
        c = rng.uniform(-1.0, 1.0, 3)
        if kind in (0, 1, 2):                       # ordinary, sizes over three decades
            s = 10.0 ** rng.uniform(-2.5, -0.5)
            add(c + s * rng.standard_normal(3), c + s * rng.standard_normal(3), c + s * rng.standard_normal(3))
        elif kind == 3:                             # sliver: third vertex almost on the first
            s = 10.0 ** rng.uniform(-2.0, -0.3)
            a, b = c + s * rng.standard_normal(3), c + s * rng.standard_normal(3)
            add(a, b, a + 1e-5 * s * rng.standard_normal(3))
        elif kind == 4:                             # sliver: nearly collinear
            s = 10.0 ** rng.uniform(-2.0, -0.3)
            d = rng.standard_normal(3)
            a = c + s * d
            add(a, c - s * d, c + 1e-4 * s * rng.standard_normal(3))
        elif kind == 5:                             # zero area
            a = c + 0.1 * rng.standard_normal(3)
            b = c + 0.1 * rng.standard_normal(3)
            if i % 16 == 5:
                add(a, a, b)
            else:
                add(a, b, a + 0.5 * (b - a))        # collinear (up to rounding)
        elif kind == 6:                             # huge: spans the whole box and beyond
            add(3.0 * rng.standard_normal(3), 3.0 * rng.standard_normal(3), 3.0 * rng.standard_normal(3))
        else:                                       # tiny
            s = 10.0 ** rng.uniform(-5.0, -3.0)
            add(c + s * rng.standard_normal(3), c + s * rng.standard_normal(3), c + s * rng.standard_normal(3))
    return np.array(v, np.float32), np.array(f, np.int32)


def fuzz_cameras(n, seed, aabb, leaf_boxes=None):
    """n camera keyword dicts (width, height, eye, facing, h_fov, anti_aliasing, samples_per_pixel) around the box `aabb`
    (6 floats); `leaf_boxes` (k×6) lets some eyes sit exactly on a leaf face."""
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(aabb[:3], np.float64), np.asarray(aabb[3:], np.float64)
    c, e = (lo + hi) / 2, np.maximum(hi - lo, 1e-3)
    diag = float(np.linalg.norm(e))
    out = []
    for i in range(n):
        kind = i % 6
        if kind == 0:                               # inside the box
            eye = lo + rng.uniform(0.02, 0.98, 3) * e
        elif kind == 1 and leaf_boxes is not None and len(leaf_boxes):   # exactly on a leaf face
            b = leaf_boxes[rng.integers(len(leaf_boxes))].astype(np.float64)
            eye = b[:3] + rng.uniform(0.0, 1.0, 3) * (b[3:] - b[:3])
            ax = rng.integers(3)
            eye[ax] = b[ax] if rng.integers(2) else b[3 + ax]
        elif kind == 2:                             # far away
            d = rng.standard_normal(3)
            eye = c + d / np.linalg.norm(d) * diag * 10.0 ** rng.uniform(0.3, 2.5)
        elif kind == 3:                             # grazing: just outside a face of the model box
            eye = lo + rng.uniform(0.0, 1.0, 3) * e
            ax = rng.integers(3)
            eye[ax] = (lo[ax] - 1e-3 * e[ax]) if rng.integers(2) else (hi[ax] + 1e-3 * e[ax])
        else:                                       # nearby
            d = rng.standard_normal(3)
            eye = c + d / np.linalg.norm(d) * diag * rng.uniform(0.6, 2.5)
        target = lo + rng.uniform(-0.2, 1.2, 3) * e
        facing = target - eye
        if np.linalg.norm(facing) < 1e-6 * diag:
            facing = np.array([0.3, 0.2, -1.0])
        facing /= np.linalg.norm(facing)
        if i % 11 == 0:                             # nearly parallel to `up`: camera_x = normalize(cross(up, -facing)) is tiny before normalising
            facing = np.array([rng.uniform(-1, 1) * 10.0 ** rng.uniform(-4, -1), 1.0 if rng.integers(2) else -1.0,
                               rng.uniform(-1, 1) * 10.0 ** rng.uniform(-4, -1)])
        if i % 13 == 0:                             # looking away from the model
            facing = -facing
        h_fov = float(10.0 ** rng.uniform(-2.0, np.log10(3.0)))
        W = int(rng.integers(9, 200))
        H = int(rng.integers(5, 120))
        if i % 7 == 0:
            W, H = int(rng.integers(3, 40)), int(rng.integers(60, 180))     # portrait: large jitter in x relative to a pixel
        aa = int(i % 3 == 0)
        out.append(dict(width=W, height=H, eye=tuple(float(x) for x in eye), facing=tuple(float(x) for x in facing),
                        h_fov=h_fov, anti_aliasing=aa, samples_per_pixel=2 if aa else 1))
    return out
