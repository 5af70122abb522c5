// Host-side data model of the render path: mesh arrays, model box, the 8-way spatial tree and its flat GPU form.
// These are the ☆ rows of SURVEY.md §2 — they stay on the CPU and run once per scene; the renderer proper is CUDA.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace atray {

struct Vec3 { float x, y, z; };
struct Box { Vec3 lo, hi; };
struct Face { int32_t v[3]; int32_t vt[3]; int32_t vn[3]; };
struct PrimRef { Vec3 a, b, c; uint32_t face; };  // KD_Primitive: leaf-local copy of the triangle + original face index

struct TreeNode
{
    Box box;
    int32_t children_start = 0;      // 0 ⇔ leaf (reference keeps has_children in the same word)
    std::vector<PrimRef> prims;      // leaves only (inner nodes hand their list to the children)
};

struct HostModel
{
    std::vector<Vec3> vertices, normals, tex_coords;
    std::vector<Face> faces;
    Box aabb{};                      // Model::surrounding_aabb
    bool has_aabb = false;
    std::vector<TreeNode> tree;      // node order identical to the reference's DBuffer<KD_Node>
    uint32_t max_faces_per_node = 0;
    int division_method = 1;
};

// Flat, 16-byte-aligned GPU form.
struct FlatScene
{
    // node i = two float4: {lo.xyz, w0} {hi.xyz, w1}; inner: w0 = children_start (>0), w1 = 0;
    // leaf: w0 = ~first_slot (<0, first_slot multiple of 8), w1 = primitive count.
    std::vector<float> nodes;        // n_nodes * 8
    // triangle slots in packets of 8: packet p = 8×{a.xyz, face} | 8×{ab.xyz, position in the reference's leaf order}
    // | 8×{ac.xyz, 0}; within a leaf the slots are in Morton order of the triangle centroids
    std::vector<float> tris;         // n_slots/8 * 24 * 4
    std::vector<uint32_t> packet_leaf;  // n_slots/8: node index of the leaf that owns the packet
    // inner nodes grouped by depth, deepest level first (bottom-up union of the per-frame film rectangles):
    // level l = inner_order[level_start[l] .. level_start[l+1])
    std::vector<uint32_t> inner_order, level_start;
    // tables of the film-space rasteriser (render.cu: k_raster / k_resolve):
    // leaf_seq[node]  = rank of the leaf in the reference's insertion order when every node is visited (kd_tree.cpp:363-435:
    //                   LIFO over inner nodes, children examined in index order, leaves inserted when their parent is
    //                   processed) — a ray's own insertion order is this order restricted to the leaves it lists;
    // seq_leaf        = inverse of leaf_seq over the leaves; parent[node] = parent node (root: 0xffffffff);
    // pos_face[first_slot + pos] = face index of the leaf's primitive at position `pos` of the reference's leaf order.
    std::vector<uint32_t> leaf_seq, seq_leaf, parent, pos_face;
    // node_pk[2*node], node_pk[2*node+1] = packets [begin, end) of the node's subtree (slots are laid out depth-first)
    std::vector<uint32_t> node_pk;
    uint32_t n_nodes = 0, n_slots = 0, n_leaves = 0, max_leaf_prims = 0;
};

int load_obj(const char* path, int threads, HostModel& m, std::string& err);
void compute_aabb(HostModel& m);
void resize_scale(HostModel& m, float new_max_scale);
void translate_to(HostModel& m, Vec3 new_center);
// threads > 1: the subtrees below the root are built concurrently (same node array, see host_model.cpp)
int build_tree(HostModel& m, int method, uint32_t cap, std::string& err, int threads = 1);
void flatten(const HostModel& m, FlatScene& out);

struct CameraBasis { float aspect; Vec3 cz, cx, cy, eye, fc; float h_fov, hpw, hph; };
CameraBasis make_camera(int w, int h, Vec3 eye, Vec3 facing, float h_fov);

}  // namespace atray
