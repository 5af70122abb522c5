"""GPU: randomized soundness check of the film-space culling.

Every production frame depends on view_bounds.h's float-error margins: a ray the reference accepts but a rectangle
excludes would be a silently wrong pixel. ≥ 500 seeded cameras per scene (eye inside the model box, on a leaf face,
grazing a face, far away; h_fov 0.01 … 3; facing nearly parallel to `up`; portrait and odd frame sizes; AA on/off)
× {knot_mid SAH/300, knot_mid CENTER/32, Monkey, a sliver mesh with zero-area and eye-plane-crossing triangles}:
the culled frame must equal the unculled one bit for bit (the unculled kernel is pinned to the oracle by
test_gpu_parity.py). A counter-example, once fixed, stays in REGRESSION_CAMERAS."""
import numpy as np
import pytest

from fuzz_cams import fuzz_cameras, sliver_mesh
from helpers import SEED, same_bits, write_obj

pytestmark = pytest.mark.gpu

N_CAMERAS = 520

# (scene name, camera kwargs) of every counter-example found so far — none yet
REGRESSION_CAMERAS = []


@pytest.fixture(scope="module")
def A(built):
    import atray_b200 as A
    assert A.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    return A


def scene_for(A, name, meshes, tmp_path_factory):
    if name == "slivers":
        v, f = sliver_mesh()
        path = write_obj(str(tmp_path_factory.mktemp("fuzz") / "slivers.obj"), v, f)
        m = A.load_model_data(path, 2)
        A.place_default(m, 40.0)
        sc = A.Scene(m, division_method=A.DIV_SAH, max_no_faces_per_node=64)
    else:
        mesh, method, cap = {"knot_sah300": ("knot_mid", 1, 300), "knot_center32": ("knot_mid", 0, 32),
                             "monkey": ("Monkey", 1, 300)}[name]
        if mesh not in meshes:
            pytest.skip("mesh not staged")
        path, rescale = meshes[mesh]
        m = A.load_model_data(path, 4)
        A.place_default(m, rescale)
        sc = A.Scene(m, division_method=method, max_no_faces_per_node=cap)
    A.prep_scene(sc)
    return sc


def check_camera(A, sc, kw, seed):
    cam = A.set_camera(**kw)
    a = sc.ctx.render(cam, sc.ctx.opts(seed=seed))
    b = sc.ctx.render(cam, sc.ctx.opts(seed=seed, cull=False))
    ok = (a["prim"] == b["prim"]).all() and same_bits(a["t"], b["t"]) and (a["bgrx"] == b["bgrx"]).all()
    return ok, int((a["prim"] != b["prim"]).sum()), int((b["prim"] != A.MISS).sum())


@pytest.mark.parametrize("name", ["knot_sah300", "knot_center32", "monkey", "slivers"])
def test_culled_equals_unculled_on_random_cameras(name, A, meshes, tmp_path_factory):
    sc = scene_for(A, name, meshes, tmp_path_factory)
    tree = sc.model.tree()
    leaves = tree["boxes"][tree["children_start"] == 0]
    cams = fuzz_cameras(N_CAMERAS, 1000 + len(name), sc.model.dump()["aabb"], leaves)
    cams += [kw for (n, kw) in REGRESSION_CAMERAS if n == name]
    bad, hit_frames = [], 0
    for i, kw in enumerate(cams):
        ok, n_diff, n_hit = check_camera(A, sc, kw, SEED + i)
        hit_frames += n_hit > 0
        if not ok:
            bad.append((i, n_diff, kw))
    assert not bad, "culled != unculled for %d of %d cameras, first: %r" % (len(bad), len(cams), bad[:3])
    assert hit_frames > len(cams) // 4, "the fuzz cameras hardly ever see the model (%d of %d)" % (hit_frames, len(cams))
