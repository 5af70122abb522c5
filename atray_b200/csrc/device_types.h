// Types shared between the CUDA kernels (render.cu) and the C-ABI host code (capi.cpp).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "view_bounds.h"

namespace atray {

struct DevScene
{
    const float4* nodes;   // 2 float4 per node
    const float4* tris;    // packets of 8 slots: 8×{a,face} | 8×{ab,pos} | 8×{ac,0}
    const uint32_t* inner_order;   // inner nodes, deepest level first (FlatScene::inner_order)
    const uint32_t* level_start;   // n_levels + 1 offsets into inner_order
    uint32_t n_levels;
    uint32_t n_nodes;
    uint32_t n_slots;
    uint32_t root_is_leaf;
    float sky[3];          // materials[0].emission_color (renderer.cpp:152, 226)
    float mdl[3];          // model material emission_color (renderer.cpp:258)
    // ---- multi-bounce path (k_path): full materials, per-face normal sources, analytic primitives -----------------
    float mdl_refl[3];     // model material reflection_color / scatter (material.h:4-9)
    float mdl_scatter;
    const float4* fnorm;   // smooth: 3 float4 per face = the face's vertex normals (renderer.cpp:128-137);
    uint32_t smooth;       // flat: 1 float4 per face = cross(v0 - v1, v0 - v2) precomputed in f32 (renderer.cpp:141-146)
    const float4* spheres; // 3 float4 per sphere: {center, radius} {emission, scatter} {reflection, 0}  (sphere.h:5-10)
    const float4* planes;  // 3 float4 per plane : {normal, distance} {emission, scatter} {reflection, 0} (plane.h:5-10)
    uint32_t n_spheres, n_planes;
    // ---- film-space rasteriser (k_raster / k_resolve): FlatScene::packet_leaf, leaf_seq, seq_leaf, parent, pos_face ----
    const uint32_t* packet_leaf;   // n_slots/8: leaf node that owns the packet
    const uint32_t* leaf_seq;      // n_nodes  : static insertion rank of a leaf (kd_tree.cpp:363-435 with every node visited)
    const uint32_t* seq_leaf;      // n_leaves : inverse of leaf_seq
    const uint32_t* parent;        // n_nodes  : parent node, 0xffffffff for the root
    const uint32_t* pos_face;      // n_slots  : [first_slot + position in the reference's leaf order] → face index
    const uint2* node_pk;          // n_nodes  : packets [x, y) of the node's subtree (leaf slots are laid out in depth-first order,
                                   //            so every subtree owns one contiguous packet range; FlatScene::node_pk)
    // ---- no-tree path (renderer.cpp:58-82, the #ifndef USE_KD_TREE branch): model box, then every face in file order ----
    const float4* faces;           // 3 float4 per face: {a, 0} {b - a, 0} {c - a, 0} from the vertex arrays
    uint32_t n_faces;
    uint32_t brute;                // per launch: non-zero → the model query is the no-tree branch
    float aabb_lo[3], aabb_hi[3];  // Model::surrounding_aabb (model.h:41-61 after the transforms)
    // ---- per context, not per scene: the cancel word every persistent loop polls (atray_render_cancel) ----
    const uint32_t* cancel;
};

struct DevCamera
{
    int32_t W, H;
    int32_t aa;
    uint32_t spp;
    float aspect, h_fov, hpw, hph;
    float eye[3], fc[3], cx[3], cy[3];
    int32_t bounce;        // RenderSettings.bounce_limit (settings.h:4-10)
};

struct DevPartition { int32_t rank, world, band_rows; };

struct DevOut { uint32_t* bgrx; float* t; uint32_t* prim; };

struct DevCounters  // mirrors AtrayCounters
{
    unsigned long long rays, box_tests, tri_tests, pass_det, pass_u, pass_v, leaves_listed, leaves_scanned, overflow_rays;
    unsigned long long rect_tests, tiles_sky, tiles_busy;
};

// Per-frame film-space bounds written by k_view_bounds / k_view_nodes and read by the culled render kernel (view_bounds.h).
// Every bound is an inclusive pixel range packed as uint2 {x0 | y0 << 16, x1 | y1 << 16} (film_rect_px: the pixels that
// can hold a ray whose film point lies in the float rectangle); empty = {0xffffffff, 0}.
struct DevView
{
    uint2* tri_px;         // n_slots
    uint2* pkt_px;         // n_slots/8  union over the packet's slots
    uint2* chunk_px;       // n_slots/256 union over 32 consecutive packets (one k_view_bounds block)
    uint2* node_px;        // n_nodes    union over the packets of the node's subtree
    float4* tri_rect;      // n_slots {xlo, ylo, xhi, yhi} float rectangles; only the rasteriser reads them (null otherwise)
    uint32_t* stamps;      // one per 8x4-pixel tile: == frame_id ⇔ some triangle's pixel range touches the tile
    uint32_t* tile_cost;   // one per tile: packet ranges that touch it this frame (k_view_nodes reads and clears it)
    uint32_t* all_flag;    // == frame_id ⇔ a triangle covers half the frame or more: every tile is busy
    uint32_t* busy_list;   // tiles of this rank's bands that must be traced (local tile numbering of the render kernel):
                           // kCostBuckets regions of bucket_stride entries, see below
    uint32_t bucket_stride; // entries per cost class (>= tiles of the frame)
    uint4* tile_log;       // verification aid (ATRAY_TILE_LOG=1), else null: per local tile {start ns, duration ns, sampled cost, class}
    int32_t cost_shift;    // class = floor(log2(ranges)) - cost_shift, clamped to [0, kCostBuckets)
    int32_t cost_sample;   // one packet in 2^cost_sample contributes to the tile costs (cost_shift = -cost_sample keeps the classes)
    // Bins of kBinW x kBinH pixels: the tree nodes whose pixel range touches the bin, as flat lists (k_view_nodes). A tile
    // (8x4 pixels, inside exactly one bin) filters its bin's lists instead of walking the tree.
    //   leaf entry  = 3 uint4: {px.lo, px.hi, leaf_seq, first_slot} {box lo.xyz, count} {box hi.xyz, node}
    //   inner entry = 2 uint4: {px.lo, px.hi, node, 0} {split point xyz (= hi corner of child 0, kd_tree.cpp:120-148), 0}
    uint2* big_list;       // kBigCap pixel ranges of triangles whose range covers more than 48 tiles: stamped by the node blocks of k_view_nodes, a warp per range
    uint4* wide_inner;     // kWideCap x 2 (same entry form): inner nodes near the root, whose range touches hundreds of bins
    uint32_t* bin_cnt;     // 2 per bin: leaf entries, inner entries (may exceed the capacities: the bin is then unusable)
    uint4* bin_leaf;       // bins x bin_leaf_cap x 3
    uint4* bin_inner;      // bins x bin_inner_cap x 2
    int32_t bins_x, bins_y;
    uint32_t bin_leaf_cap, bin_inner_cap;
    uint32_t* busy_count;  // control words, see kCtl* below (per-slab list lengths, cursors, finished-tile counters, ticket)
    uint32_t frame_id;
    int32_t tiles_x;
};
#ifndef ATRAY_BIN_SHIFT_X
#define ATRAY_BIN_SHIFT_X 6
#endif
#ifndef ATRAY_BIN_SHIFT_Y
#define ATRAY_BIN_SHIFT_Y 5
#endif
static constexpr int kBinShiftX = ATRAY_BIN_SHIFT_X, kBinShiftY = ATRAY_BIN_SHIFT_Y;   // 64 x 32 pixels = 8 x 8 tiles (A/B builds may override)
static constexpr int kWideCap = 32, kWideBins = 48;    // an inner node that touches more bins goes to the one wide list

// Where k_view_bounds stores what it computes. One GPU: n = 1 and the tables of `view`. N GPUs that share the per-frame view
// pass (atray_view_share_*): every rank computes the pixel ranges of 1/N of the triangle slots — blocks [block_lo, block_hi)
// of 256 slots — and stores them into the tables of ALL ranks through peer pointers (NVLink); a tile's stamp and cost go
// to the one rank that owns the tile's band. The pass is then followed by a flag per rank that the peers' streams wait on.
static constexpr int kMaxPeers = 8;
struct ViewPeers
{
    int32_t n;                        // ranks the results go to (1 = only this GPU)
    uint32_t block_lo, block_hi;      // this rank's share of the k_view_bounds blocks
    uint32_t debug_skip;              // verification aid (ATRAY_SHARE_SKIP): 1 no remote cost, 2 no remote stamps, 4 no remote ranges
    uint32_t inline_big;              // k_view_bounds stamps large ranges itself instead of queueing them for k_view_nodes
    uint32_t ranges_only;             // shared pass: peers receive the ranges only; each rank stamps its own tiles in k_view_stamp
    int32_t self;                     // this rank's index in the tables
    uint2* tri_px[kMaxPeers];
    uint2* pkt_px[kMaxPeers];
    uint2* chunk_px[kMaxPeers];
    uint32_t* stamps[kMaxPeers];
    uint32_t* tile_cost[kMaxPeers];
    uint32_t* all_flag[kMaxPeers];
    // shared pass only: the last block of k_view_bounds to finish raises flags[q][self] = seq on every rank q (after a
    // system-wide fence by every thread); k_view_nodes of rank q waits for the flags of all ranks before it reads the tables
    uint32_t* flags[kMaxPeers];
    uint32_t seq;
    uint32_t* ticket;                 // this GPU: blocks of the pass that have finished (reset by the last one)
};
struct ViewWait
{
    const uint32_t* my_flags;         // null: nothing to wait for
    int32_t n, rank;
    uint32_t seq;
    unsigned long long timeout_ns;    // a peer that never signals ends the wait with *timed_out = seq instead of hanging the GPU
    uint32_t* timed_out;
};

// Per-frame state of the film-space rasteriser (render.cu: k_film_tables, k_raster, k_raster_big, k_hit_filter, k_resolve).
// key1/key2 hold one 64-bit word per pixel each and are all-ones between frames (k_resolve restores what it consumes):
//   key1 = (entry distance bits << 32) | leaf_seq   minimum over the (leaf, triangle) pairs the reference would accept for the
//          pixel's ray = the first leaf of the ray's sorted leaf list that yields a hit (kd_tree.cpp:437-462); 0 = the pixel
//          must be traced by the exact per-ray walker (irregular ray, possible "at most 5 children" cut, hit list overflow);
//   key2 = (t bits << 32) | position in the reference's leaf order, minimum over the accepted triangles of that leaf.
struct DevRaster
{
    float4* tabx;                  // W: {frame_center + camera_x*film_x (xyz), film_x}      (renderer.cpp:329,350)
    float4* taby;                  // H: {camera_y*film_y (xyz), film_y}                      (renderer.cpp:317,350)
    unsigned long long* key1;
    unsigned long long* key2;
    uint4* hit_a;                  // accepted hits that were the best of their pixel when found: {pixel, dis bits, leaf_seq, t bits}
    uint32_t* hit_pos;             // … and the position of the triangle in the reference's leaf order
    uint32_t hit_cap;
    uint2* big;                    // {slot, pixel pairs} of the slots whose rectangle is too large for one warp pass
    uint32_t big_cap;
    uint32_t* ctl;                 // [0] hits appended, [1] big-list entries
};

// Slabs: the rank's tiles (band-major local numbering) split into up to kMaxSlabs contiguous ranges of whole bands. The
// render kernel finishes slab after slab and publishes a finished-tile counter per slab, so the host side can start the
// device→host copy of a slab's scanlines while later slabs are still being traced (capi.cpp: atray_render_begin).
// The busy tiles of a slab are listed per cost class (class = log2 of the number of triangle rectangles touching the tile,
// see tile_cost_class) and traced most expensive class first: the longest tiles start earliest, which is what bounds the
// kernel's tail. Region b of busy_list holds class b; inside it the segment of slab s starts at tile_begin[s].
static constexpr int kMaxSlabs = 16;
static constexpr int kCostBuckets = 8;
static constexpr int kCtlCount = 0;                              // [s * kCostBuckets + b] entries of slab s in class b
static constexpr int kCtlCursor = kMaxSlabs * kCostBuckets;      // [s] consumer cursor of slab s
static constexpr int kCtlDone = kCtlCursor + kMaxSlabs;          // [s] tiles of slab s whose stores are complete (sky + traced)
static constexpr int kCtlWide = kCtlDone + kMaxSlabs;            // inner nodes in the wide list (DevView::wide_inner)
static constexpr int kCtlBigDone = kCtlWide + 1;                 // node blocks of k_view_nodes that have stamped their share of the queued large ranges
static constexpr int kCtlTicket = kCtlBigDone + 1;               // end of the words k_view_bounds clears every frame
static constexpr int kCtlBig = kCtlTicket + 1;                   // entries of DevView::big_list (reset by the last node block of k_view_nodes to have used it)
static constexpr int kCtlWords = kCtlTicket + 4;
static constexpr uint32_t kBigCap = 1u << 16;
struct DevSlabs
{
    int32_t n;                              // 1 … kMaxSlabs
    uint32_t tile_begin[kMaxSlabs + 1];     // local tile index range of slab s = [tile_begin[s], tile_begin[s+1])
    uint32_t live_tiles[kMaxSlabs];         // tiles of slab s that lie inside the image = final value of ctl[kCtlDone + s]
};

// Per-warp spill area for rays whose leaf list / inner-node stack outgrow the per-lane arrays.
struct DevScratch
{
    uint2* leaves;        // warps × leaf_cap
    uint32_t* stack;      // warps × stack_cap
    uint32_t leaf_cap, stack_cap;
};

struct LaunchCfg { int grid, block; };

// host-callable launchers (render.cu)
cudaError_t launch_render(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                          const DevScratch& scratch, unsigned int* tile_counter, DevCounters* counters, int fp_fast,
                          unsigned long long jitter_seed, LaunchCfg cfg, cudaStream_t stream);
// culled frame = k_view_bounds (per-frame film rectangles + tile stamps) followed by k_render_culled
// k_view_bounds (per-slot pixel ranges, packet / chunk unions, tile stamps) + k_view_nodes (node unions, bins, compact list
// of this rank's busy tiles)
// `peers` = null: everything into `view` (one GPU, or a rank that computes the whole pass itself). `wait` (shared pass):
// k_view_nodes starts by waiting for the flags of the other ranks.
cudaError_t launch_view_bounds(const DevScene& sc, const ViewParams& vp, const DevView& view, const DevCamera& cam,
                               const DevPartition& part, const DevSlabs& slabs, cudaStream_t stream, cudaEvent_t after_reset,
                               bool finish = true, const ViewPeers* peers = nullptr, const ViewWait* wait = nullptr,
                               cudaEvent_t after_bounds = nullptr);   // (timing event between the two kernels)
// one warp: flags[p][rank] = seq for every peer p, after a system-wide fence (the peers' streams wait for the value)
cudaError_t launch_view_signal(uint32_t* const* peer_flags, int n, int rank, uint32_t seq, cudaStream_t stream);
// one warp: waits (on the stream) until every other rank's flag in my_flags has reached seq; *timed_out = seq after timeout_ns
cudaError_t launch_view_wait(const uint32_t* my_flags, int n, int rank, uint32_t seq, unsigned long long timeout_ns, uint32_t* timed_out,
                             cudaStream_t stream);
// rasterised frame (AA off, root is an inner node): k_film_tables → [k_view_bounds by the caller] → k_raster → k_raster_big →
// k_hit_filter → k_resolve (sky fill + per-pixel decode + exact per-ray fallback, slab by slab like the culled kernel)
cudaError_t launch_film_tables(const DevCamera& cam, const DevRaster& rs, cudaStream_t stream);
cudaError_t launch_raster(const DevScene& sc, const ViewParams& vp, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                          const DevScratch& scratch, const DevView& view, const DevRaster& rs, const DevSlabs& slabs,
                          DevCounters* counters, int fp_fast, LaunchCfg cfg, cudaStream_t stream);
cudaError_t launch_render_culled(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                                 const DevScratch& scratch, const DevView& view, const DevSlabs& slabs, DevCounters* counters,
                                 int fp_fast, unsigned long long jitter_seed, LaunchCfg cfg, cudaStream_t stream);
// local tile count of a rank and the slab table for `n_slabs` (clamped to the number of local bands)
DevSlabs make_slabs(const DevCamera& cam, const DevPartition& part, int n_slabs, const int* weights = nullptr);
cudaError_t launch_slabs_final(const DevView& view, const DevSlabs& slabs, cudaStream_t stream);
// K3: pixel loop + cast_ray at any bounce_limit with spheres/planes (renderer.cpp:34-160, 213-262, 314-365)
cudaError_t launch_path(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                        const DevScratch& scratch, unsigned int* tile_counter, unsigned long long* casts, int fp_fast,
                        unsigned long long seed, LaunchCfg cfg, cudaStream_t stream);
cudaError_t launch_trace(const DevScene& sc, const float* rays, unsigned long long n, uint32_t* prim, float* t,
                         const DevScratch& scratch, unsigned int* tile_counter, DevCounters* counters, int fp_fast,
                         LaunchCfg cfg, cudaStream_t stream);
cudaError_t launch_trace_uv(const DevScene& sc, const float* rays, unsigned long long n, uint32_t* prim, float* t, float* uv,
                            const DevScratch& scratch, unsigned int* tile_counter, LaunchCfg cfg, cudaStream_t stream);
// K2b: C5 — `spp` incoherent rays per pixel index in [p0, p1), generated on the device, traced, reduced to the pixel colour
cudaError_t launch_c5(const DevScene& sc, unsigned long long p0, unsigned long long p1, uint32_t spp, uint32_t chunk_px,
                      unsigned long long seed, const DevPartition& part, uint32_t* bgrx, uint32_t* prim, float* t, const DevScratch& scratch,
                      unsigned int* counter, DevCounters* counters, int fp_fast, LaunchCfg cfg, cudaStream_t stream);
// ray-list trace on the file-scope static scratch of render.cu (no allocation anywhere): see launch_trace_static there
static constexpr size_t kStaticScratchBytes = size_t(16) << 20;   // per GPU that loads the library, held until exit
size_t static_scratch_capacity();
size_t static_trace_bytes_per_block(uint32_t leaf_cap, uint32_t stack_cap);
int static_trace_blocks(uint32_t leaf_cap, uint32_t stack_cap, LaunchCfg cfg);
cudaError_t static_scratch_address(int device, void** out);
cudaError_t launch_trace_static(const DevScene& sc, const float* rays, unsigned long long n, uint32_t* prim, float* t, uint32_t leaf_cap,
                                uint32_t stack_cap, int fp_fast, LaunchCfg cfg, int device, cudaStream_t stream);
cudaError_t launch_microbench_fp32(float* sink, int iters, LaunchCfg cfg, cudaStream_t stream);
cudaError_t launch_microbench_l2(const float4* buf, size_t n_vec, int passes, float* sink, LaunchCfg cfg, cudaStream_t stream);
LaunchCfg render_launch_cfg(int sm_count);
int render_warps_in_flight(LaunchCfg cfg);

}  // namespace atray
