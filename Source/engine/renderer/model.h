// Model / ModelData and the model transforms (engine/renderer/model.h:30-152) over the C ABI.
#pragma once
#include "aabb.h"
#include "kd_tree.h"
#include "material.h"
#include "atray_b200.h"

struct ModelData
{
    Material* material = nullptr;
    atray_model* handle = nullptr;  // loader output + tree (vertices/normals/faces stay host-side inside the library)
};
struct Model { KD_Tree kd_tree; ModelData data; AABB surrounding_aabb{}; };

// model.h:41-61 — also records the box inside the handle so resize_scale/translate_to can use it
inline AABB get_AABB(ModelData& data)
{
    AABB b{};
    atray_model_compute_aabb(data.handle, &b.min.x);
    return b;
}
// model.h:108-121
inline void resize_scale(Model& mdl, f32 new_max_scale)
{
    atray_model_resize_scale(mdl.data.handle, new_max_scale);
    atray_model_dump(mdl.data.handle, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, &mdl.surrounding_aabb.min.x);
}
// model.h:123-134: declared by the reference with an empty body (no rotation is implemented there either)
struct Bases { vec3f origin, x_axis, y_axis; };
inline void transform(Model& mdl, vec3f axis, f32 radian) { (void)mdl; (void)axis; (void)radian; }
// model.h:136-152
inline void translate_to(Model& mdl, vec3f new_center)
{
    atray_model_translate_to(mdl.data.handle, &new_center.x);
    atray_model_dump(mdl.data.handle, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, &mdl.surrounding_aabb.min.x);
}
// kd_tree.cpp:20-64
inline void build_KD_tree(ModelData mdl, KD_Tree& tree)
{
    if (atray_model_build_tree(mdl.handle, tree.division_method == KD_Division_Method::SAH ? ATRAY_DIV_SAH : ATRAY_DIV_CENTER,
                               tree.max_no_faces_per_node) != ATRAY_OK)
        return;
    uint32_t c[8];
    atray_model_counts(mdl.handle, c);
    tree.tree.length = (int32)c[4];
    tree.tree.front = mdl.handle;
}
