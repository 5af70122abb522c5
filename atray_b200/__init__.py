"""atray_b200 — B200-native (sm_100a CUDA) implementation of ATRay's per-pixel render path behind a C ABI.
See DESIGN.md. Importing this package loads atray_b200/libatray_b200.so; there is no CPU fallback."""
from . import api  # noqa: F401  (raises ImportError when the library is not built)
from .api import (AtrayError, DIV_CENTER, DIV_SAH, FP_FAST, FP_STRICT, MISS, MAX_FLOAT, build_KD_tree, get_AABB,  # noqa: F401
                  load_model_data, model_from_arrays, place_default, resize_scale, set_camera, translate_to, update_camera_pos, device_count)
from .renderer import (Context, MultiContext, PinnedArray, RenderInfo, Scene, prep_scene, start_render_from_camera,  # noqa: F401
                       wait_for_render_from_camera_to_finish, clear_work_queue, activate_pool, wait_for_pool,
                       get_ray_kd_tree_intersection)
