// KD_Tree parameters (engine/renderer/kd_tree.h:13-47). The node array lives inside the model handle
// (atray_model); `tree.length` mirrors the node count so prep_scene's "build only if empty" test keeps working.
#pragma once
#include "PL/PL_math.h"
#include "atray_b200.h"
#include "ray.h"
enum class KD_Division_Method { CENTER, SAH };
struct KD_NodeArrayView { int32 length = 0; const void* front = nullptr; };
struct KD_Node;                                            // kd_tree.h:23-36: opaque here (the nodes live in the library)
struct LeafNodePair { KD_Node* node; f32 distance_from_ray; };   // kd_tree.h:49-53: the caller's leaf stack, unused on the device
struct TriangleIntersectionData { uint32 face_index; f32 u, v; };   // model.h:18-22
struct KD_Tree
{
    uint32 max_no_faces_per_node = 300;
    KD_Division_Method division_method = KD_Division_Method::SAH;
    KD_NodeArrayView tree;
    atray_ctx* gpu = nullptr;   // device context holding the uploaded tree (set by prep_scene)
};

// get_ray_kd_tree_intersection (kd_tree.h:56, kd_tree.cpp:302-313): distance of the nearest accepted triangle along op_ray, or
// MAX_FLOAT; tri_data is written only on a hit. The two stack arguments stay in the signature and are ignored (the device kernel
// has its own). One ray per call is a round trip to the GPU: use atray_trace_rays_uv directly for batches.
inline f32 get_ray_kd_tree_intersection(Optimized_Ray& op_ray, KD_Tree& tree, TriangleIntersectionData& tri_data, KD_Node** hit_stack_front,
                                        LeafNodePair* leaf_stack_front)
{
    (void)hit_stack_front; (void)leaf_stack_front;
    const float ray[6] = {op_ray.ray.origin.x, op_ray.ray.origin.y, op_ray.ray.origin.z,
                          op_ray.ray.direction.x, op_ray.ray.direction.y, op_ray.ray.direction.z};
    uint32_t prim = ATRAY_MISS;
    float t = ATRAY_MAX_FLOAT, uv[2] = {tri_data.u, tri_data.v};
    if (!tree.gpu || atray_trace_rays_uv(tree.gpu, ray, 1, &prim, &t, uv) != ATRAY_OK) return ATRAY_MAX_FLOAT;
    if (prim != ATRAY_MISS) { tri_data.face_index = prim; tri_data.u = uv[0]; tri_data.v = uv[1]; }
    return t;
}
