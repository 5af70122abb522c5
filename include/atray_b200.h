/* atray_b200 — C ABI of the B200-native (sm_100a) implementation of ATRay's per-pixel render path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types. Each entry point names the
 * reference interface it replaces (paths relative to the reference's Source/ directory). The reference has no
 * FFI of its own: its boundary is the C++ function surface app.cpp calls, which the host layer
 * Source/engine/ in this repo re-exposes (same names/signatures) on top of this ABI. See INTEGRATION.md.
 *
 * All functions return 0 (ATRAY_OK) on success or a negative ATRAY_E* code; atray_last_error() returns a
 * thread-local message. There is NO CPU fallback: every render/trace entry fails with ATRAY_ECUDA when no
 * sm_100-class device is usable.
 */
#ifndef ATRAY_B200_H
#define ATRAY_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ATRAY_OK 0
#define ATRAY_EINVAL (-1)   /* bad argument */
#define ATRAY_EIO (-2)      /* file could not be read */
#define ATRAY_ECUDA (-3)    /* CUDA error / no device */
#define ATRAY_ESTATE (-4)   /* call order violated (e.g. render before upload) */
#define ATRAY_ELIMIT (-5)   /* tree build exceeded the node/primitive guard (reference: no guard, kd_tree.cpp:74-286) */

#define ATRAY_MISS 0xffffffffu          /* primitive id written for rays that hit nothing */
#define ATRAY_MAX_FLOAT 3.402823466e+38F /* t written for misses (PL/PL_base_defs.h:70, renderer.cpp:36) */

#define ATRAY_DIV_CENTER 0 /* KD_Division_Method::CENTER, engine/renderer/kd_tree.h:13-16 */
#define ATRAY_DIV_SAH 1    /* KD_Division_Method::SAH */

#define ATRAY_FP_STRICT 0 /* reference arithmetic: no FMA contraction, IEEE division — bit-exact parity mode */
#define ATRAY_FP_FAST 1   /* FMA contraction allowed; ids/t inside north_star tolerances, not bit-exact */

/* AtrayRenderOpts.flags. Default (0): per-frame film-space culling — identical results, far less work.
 * NO_CULL renders with the unculled kernel that scans every listed leaf completely, as the reference does. */
#define ATRAY_RENDER_NO_CULL 1u
/* RASTER (opt-in, frames without anti-aliasing): the frame is rasterised triangle by triangle in film space instead of
 * traversed tile by tile — identical results again; measured slower on the headline workload. */
#define ATRAY_RENDER_RASTER 2u

typedef struct atray_model atray_model; /* host: ModelData + Model + KD_Tree (model.h:30-71, kd_tree.h:38-47) */
typedef struct atray_ctx atray_ctx;     /* one GPU: uploaded scene, streams, frame buffers */

/* Material, engine/renderer/material.h:4-9 (28 bytes, same layout) */
typedef struct AtrayMaterial { float emission_color[3]; float reflection_color[3]; float scatter; } AtrayMaterial;

/* Sphere / Plane with their material by value (engine/renderer/sphere.h:5-10, plane.h:5-10 hold a Material*). */
typedef struct AtraySphere { float center[3]; float radius; AtrayMaterial material; } AtraySphere;
typedef struct AtrayPlane { float normal[3]; float distance; AtrayMaterial material; } AtrayPlane;

/* Camera + RenderSettings as the kernel consumes them: the values set_camera/update_camera_pos compute
 * (engine/renderer/camera.h:9-45, settings.h:4-10). Filled by atray_camera_set on the host. */
typedef struct AtrayCamera
{
    int32_t width, height;            /* RenderSettings.resolution */
    int32_t anti_aliasing;            /* RenderSettings.anti_aliasing (b32) */
    uint32_t samples_per_pixel;       /* RenderSettings.samples_per_pixel */
    int32_t bounce_limit;             /* RenderSettings.bounce_limit: 1 = primary rays (film-space culled kernel);
                                         > 1 = cast_ray's bounce loop, renderer.cpp:213-262 (multi-bounce kernel) */
    float aspect_ratio;
    float camera_z[3], camera_x[3], camera_y[3];
    float eye[3], frame_center[3];
    float h_fov, half_pixel_width, half_pixel_height;
} AtrayCamera;

/* Which part of the image this context renders (replaces the tile work queue of renderer.cpp:403-445):
 * bands of band_rows scanlines, band b belongs to partition (b % world); world = 1 renders everything. */
typedef struct AtrayPartition { int32_t rank, world, band_rows; } AtrayPartition;

/* Per-frame options. */
typedef struct AtrayRenderOpts
{
    int32_t fp_mode;        /* ATRAY_FP_STRICT / ATRAY_FP_FAST */
    int32_t count_work;     /* non-zero: accumulate AtrayCounters (slower instrumented kernel) */
    uint64_t jitter_seed;   /* anti_aliasing: pixel p uses RNG_Stream{splitmix64(seed ^ p), 2p+1} */
    AtrayPartition part;
    uint32_t flags;         /* ATRAY_RENDER_* */
} AtrayRenderOpts;

/* Host frame the renderer fills (replaces Texture/BitmapBuffer, engine/tools/texture.h:10-23, plus the two
 * extra per-pixel planes north_star asks for). Row 0 = bottom scanline, index y*width + x (texture.h:27-38).
 * Any plane may be NULL. */
typedef struct AtrayFrame
{
    uint32_t* bgrx;   /* 0x00RRGGBB (B,G,R,0 in memory) — Set_Pixel packing */
    float* t;         /* hit distance, ATRAY_MAX_FLOAT on miss */
    uint32_t* prim;   /* face index (KD_Primitive.face_index), ATRAY_MISS on miss */
} AtrayFrame;

/* Work counters of the last counted frame/trace (SURVEY §8d roofline formula). */
typedef struct AtrayCounters
{
    uint64_t rays, box_tests, tri_tests, pass_det, pass_u, pass_v, leaves_listed, leaves_scanned, overflow_rays;
    uint64_t rect_tests;  /* culled kernel: film-rectangle tests (leaf, packet, triangle level) */
    uint64_t tiles_sky;   /* culled kernel: 8x4 tiles no triangle rectangle can touch (written as sky without tracing) */
    uint64_t tiles_busy;  /* culled kernel: tiles that were traced */
} AtrayCounters;

const char* atray_last_error(void);
int atray_version(void);

/* ---- host side: loader, transforms, tree (the ☆ rows of SURVEY §2) ---------------------------------------- */
/* load_model_data(ModelData&, const char*, ThreadPool&) — engine/tools/OBJ_loader.cpp:278-361 */
int atray_model_load_obj(const char* path, int loader_threads, atray_model** out);
/* builds a model from arrays (no reference counterpart; for embedders that already hold a mesh) */
int atray_model_from_arrays(const float* vertices, uint32_t n_vertices, const int32_t* face_vertex_indices,
                            uint32_t n_faces, atray_model** out);
void atray_model_free(atray_model* m);
/* model.surrounding_aabb = get_AABB(model.data) — engine/renderer/model.h:41-61; out6 = min.xyz,max.xyz (may be NULL) */
int atray_model_compute_aabb(atray_model* m, float* out6);
/* resize_scale(Model&, f32) — model.h:108-121 */
int atray_model_resize_scale(atray_model* m, float new_max_scale);
/* translate_to(Model&, vec3f) — model.h:136-152 */
int atray_model_translate_to(atray_model* m, const float* new_center3);
/* build_KD_tree(ModelData, KD_Tree&) with KD_Tree.division_method / max_no_faces_per_node — kd_tree.cpp:20-288 */
int atray_model_build_tree(atray_model* m, int division_method, uint32_t max_no_faces_per_node);
/* counts[8] = vertices, normals, tex_coords, faces, tree nodes, primitive refs, inner nodes, largest leaf */
int atray_model_counts(const atray_model* m, uint32_t* counts8);
/* copies of the loader output (any pointer may be NULL): vertices/normals/tex n×3 f32, face index triplets n_faces×3 i32 */
int atray_model_dump(const atray_model* m, float* vertices, float* normals, float* tex_coords, int32_t* face_v,
                     int32_t* face_vt, int32_t* face_vn, float* aabb6);
/* the tree as the reference holds it, one row per node in DBuffer order (for 1:1 comparison, north_star):
 * boxes n×6, children_start n (0 = leaf), prim_count n, prim_first n (offset into the concatenated per-leaf
 * primitive arrays), tri9 refs×9 (a,b,c), face_index refs */
int atray_model_tree_dump(const atray_model* m, float* boxes, int32_t* children_start, uint32_t* prim_count,
                          uint32_t* prim_first, float* tri9, uint32_t* face_index);
/* sizes and contents of the flattened GPU form: nodes = n_nodes×2 float4, tris = n_slots/8 packets of
 * 3×8 float4 {a.xyz,face}{ab.xyz,position in the reference's leaf order}{ac.xyz,0}, Morton-ordered inside each leaf.
 * Either output pointer may be NULL to query sizes only. */
int atray_model_flatten(const atray_model* m, uint32_t* n_nodes, uint32_t* n_slots, float* nodes, float* tris);

/* Host evaluation of the per-frame film-space bounds the culled kernel uses (atray_b200/csrc/view_bounds.h), for
 * verification without a GPU: rects = n_slots×4 {xlo,ylo,xhi,yhi} per triangle slot of the flat form, slot_leaf /
 * slot_pos = owning leaf node and position in the reference's leaf order (0xffffffff for padding slots), tile_busy =
 * one byte per 8x4-pixel tile (row-major, ceil(W/8) per row), hulls = n_slots×9 {a,b,c}×3: half-planes a*fx + b*fy <= c that
 * hold for every accepted ray of the slot, px_ranges = n_slots×4 {x0,x1,y0,y1}: the inclusive pixel range the kernels use
 * ({0,-1,0,-1} when no pixel can hold an accepted ray). Any output may be NULL. Returns ATRAY_ESTATE when the camera cannot
 * be used for culling (the renderer then takes the unculled kernel). */
int atray_debug_film_rects(const atray_model* m, const AtrayCamera* cam, float* rects, uint32_t* slot_leaf,
                           uint32_t* slot_pos, uint8_t* tile_busy, float* hulls, int32_t* px_ranges);

/* set_camera(Camera&, eye, facing, RenderSettings&, h_fov) — engine/renderer/camera.h:23-45 */
int atray_camera_set(AtrayCamera* cam, int32_t width, int32_t height, int32_t anti_aliasing, uint32_t samples_per_pixel,
                     int32_t bounce_limit, const float* eye3, const float* facing3, float h_fov);

/* ---- device side --------------------------------------------------------------------------------------------- */
int atray_device_count(int* n);
int atray_ctx_create(int device, atray_ctx** out);
void atray_ctx_destroy(atray_ctx* ctx);
/* prep_scene(Scene&, uint32&) tail: flatten + upload once — renderer.cpp:264-291. materials[0] = sky
 * (renderer.cpp:152), materials[1] = the model's material (model.data.material). */
int atray_scene_upload(atray_ctx* ctx, const atray_model* m, const AtrayMaterial* sky, const AtrayMaterial* model_mat);

/* renderer.h:8 `#define USE_KD_TREE` as a run-time switch of the uploaded scene (default 1). With 0 every model query
 * of frames, paths and ray lists is the reference's no-tree branch (renderer.cpp:58-82): get_ray_AABB_intersection on
 * Model::surrounding_aabb, then every face in file order, accepting t > tolerance && t < best. O(faces) per ray — a
 * geometry-exact cross-check for the tree, not a fast path. */
int atray_scene_use_kd_tree(atray_ctx* ctx, int use_kd_tree);

/* scene.spheres / scene.planes (engine/renderer/scene.h:13-24; app.cpp:111-129,139-142): replaces the current sets
 * (n = 0 removes them). Plane normals are normalised as prep_scene does (renderer.cpp:267-270). With primitives
 * present, or bounce_limit > 1, frames run get_intersection_data in full (plane > sphere > model > sky,
 * renderer.cpp:86-157) and cast_ray's bounce loop; random draws come from the per-pixel stream
 * RNG_Stream{splitmix64(jitter_seed ^ p), 2p+1} (the reference seeds per thread from rdtsc, renderer.cpp:376-378).
 * The t / prim planes then describe the nearest hit of the last sample's primary ray (prim = ATRAY_MISS unless it is
 * a triangle). Call after atray_scene_upload (which clears them). */
int atray_scene_set_primitives(atray_ctx* ctx, const AtraySphere* spheres, uint32_t n_spheres, const AtrayPlane* planes,
                               uint32_t n_planes);
/* info.total_ray_casts of the last multi-bounce frame: bounce iterations that hit an object (renderer.cpp:260) */
int atray_last_ray_casts(atray_ctx* ctx, uint64_t* ray_casts);
/* Verification aid, no reference counterpart: with ATRAY_TILE_LOG=1 in the environment the culled render kernel records, per
 * local 8x4-pixel tile of the last frame, {start ns, duration ns, sampled cost, cost class} (4 x uint32 per tile; sky tiles
 * stay zero). tools/tile_log_probe.py reads it to check the heaviest-first schedule. */
int atray_debug_tile_log(atray_ctx* ctx, uint32_t* out, size_t n_tiles);

/* start_render_from_camera(RenderInfo&, ThreadPool&) — renderer.cpp:403-455: returns immediately; kernels and the
 * device→host copies of the frame planes are queued on the context's stream. */
int atray_render_begin(atray_ctx* ctx, const AtrayCamera* cam, const AtrayRenderOpts* opts, const AtrayFrame* host_out);
/* wait_for_render_from_camera_to_finish(RenderInfo&, ThreadPool&, ms) — renderer.cpp:457-471:
 * returns 1 while still running (timed out), 0 when finished, negative on error. */
int atray_render_poll(atray_ctx* ctx, uint32_t timeout_ms);
/* info.twq.jobs_done / info.twq.jobs.size as app.cpp:176-185 polls them for "Tiles k/n" and the live blit: scanlines
 * [0, rows_done) of the host frame are final. Large frames are rendered slab by slab (bottom first) and each slab is
 * copied out while later slabs are still being traced, so rows_done advances during the frame; small or partitioned
 * frames jump from 0 to rows_total on completion. */
int atray_render_progress(atray_ctx* ctx, int32_t* rows_done, int32_t* rows_total);
/* info.twq.clear() + wait_for_pool(UINT32MAX), the cancel path — app.cpp:165-172 (workers finish the tile they are on and find
 * the queue empty). Raises the context's cancel word from a second stream while the kernel runs; every persistent loop (primary,
 * culled, path, ray-list and C5 kernels) polls it next to its work cursor and stops taking tiles / ray groups; the call returns
 * when the streams have drained. Pixels / rays not yet traced keep what the output buffers held; the pending host copy-out is
 * dropped (atray_render_poll then reports "nothing in flight"). Acts on whatever the context stream holds: begin/poll frames and
 * the asynchronous *_device entries. Returns 1 when work was on the stream and has been cut short, 0 when the context was idle.
 * Call it from the thread that owns the context (the reference's app thread calls twq.clear() itself). */
int atray_render_cancel(atray_ctx* ctx);
/* Same frame, outputs left resident in device memory (planes are device pointers, may live on a peer GPU);
 * asynchronous on the context stream; pair with atray_ctx_sync. Used for the HBM-resident bench leg. */
int atray_render_device(atray_ctx* ctx, const AtrayCamera* cam, const AtrayRenderOpts* opts, const AtrayFrame* device_out);
int atray_ctx_sync(atray_ctx* ctx);
/* device time of the last completed atray_render_device / atray_render_begin kernel(s), CUDA events, ms */
int atray_last_kernel_ms(atray_ctx* ctx, float* ms);
/* Same frame, the render kernel alone (k_render_culled: CUDA events after the per-frame view pass and after the kernel). */
int atray_last_render_kernel_ms(atray_ctx* ctx, float* ms);
/* the view pass of the last culled frame split at the kernel boundary: k_view_bounds (with, in a shared pass, its peer
 * stores) and k_view_nodes (with its wait for the other ranks' flags) */
int atray_last_view_split_ms(atray_ctx* ctx, float* bounds_ms, float* nodes_ms);

/* get_intersection_data model branch for an arbitrary ray list — renderer.cpp:34-57 + kd_tree.cpp:302-313.
 * rays = n × {ox,oy,oz,dx,dy,dz} (direction used as given). Host pointers; synchronous. */
int atray_trace_rays(atray_ctx* ctx, const float* rays, uint64_t n, int32_t fp_mode, int32_t count_work, uint32_t* prim,
                     float* t);
/* BASELINE.json configs[4] ("C5", divergence stress): `spp` incoherent random-direction rays per pixel index p in [p0, p1),
 * generated ON THE DEVICE from the per-ray PCG32 stream (engine/PL/PL_math.h:506-541) {state = splitmix64(seed ^ i), stream = 2i+1},
 * i = p*spp + s: six rand_bi draws give an origin on a sphere of radius 1.5*mag(extent) around the centre of
 * Model::surrounding_aabb and a target inside the box; SetRay (ray.h:24-29); traced through the model query of
 * get_intersection_data (renderer.cpp:41-57, kd_tree.cpp:302-465). Outputs: bgrx[p - p0] = the pixel's flat colour (sequential
 * f32 sum of the samples' sky / model emissions / spp, clamp, truncate, pack: renderer.cpp:358-365); prim / t per ray at
 * (p - p0)*spp + s (either may be NULL). `part` (NULL = everything): the pixel range is cut into chunks of chunk_px pixels dealt
 * round-robin to the ranks (chunk k -> rank k % world); a rank traces its chunks and leaves the others' entries untouched.
 * The reference has no such entry point: its rays come from render_tile_from_camera only (renderer.cpp:294-369). */
int atray_trace_incoherent_device(atray_ctx* ctx, uint64_t p0, uint64_t p1, uint32_t spp, uint64_t seed, const AtrayPartition* part,
                                  uint32_t chunk_px, int32_t fp_mode, int32_t count_work, uint32_t* bgrx_dev, uint32_t* prim_dev,
                                  float* t_dev);
/* same with host outputs (synchronous; device buffers allocated for the call) */
int atray_trace_incoherent(atray_ctx* ctx, uint64_t p0, uint64_t p1, uint32_t spp, uint64_t seed, int32_t fp_mode, int32_t count_work,
                           uint32_t* bgrx, uint32_t* prim, float* t);

/* same with device pointers, asynchronous on the context stream */
int atray_trace_rays_device(atray_ctx* ctx, const float* rays_dev, uint64_t n, int32_t fp_mode, int32_t count_work,
                            uint32_t* prim_dev, float* t_dev);
/* get_ray_kd_tree_intersection — kd_tree.h:56, kd_tree.cpp:302-313 (+ the scan loop's tri_data update, kd_tree.cpp:436-462):
 * per ray the distance (MAX_FLOAT on a miss), TriangleIntersectionData.face_index (ATRAY_MISS) and its u, v (uv[2i], uv[2i+1];
 * untouched on a miss, like tri_data). Host pointers, synchronous, strict arithmetic. n = 1 is the reference's call shape. */
int atray_trace_rays_uv(atray_ctx* ctx, const float* rays, uint64_t n, uint32_t* prim, float* t, float* uv);
/* The same query as atray_trace_rays_device (device pointers, asynchronous on the context stream, same outputs bit for bit) for
 * callers that cannot allocate: this entry calls no cudaMalloc / cudaFree / allocating routine on any path. The kernel's scratch
 * (the work cursor of its warps, the per-warp overflow areas of the leaf list and the node stack) lives in one static device array
 * of ATRAY_STATIC_SCRATCH_BYTES inside the library, which every GPU that loads the library holds until exit. Limits: a scene whose
 * tree needs more than that for a single thread block (8 warps x ((leaves + 2) x 8 + (nodes + 8) x 4) bytes) is refused with
 * ATRAY_ELIMIT and a message naming both sizes; n <= 2^36 as for the other entry; no work counters. Smaller scenes run with as
 * many thread blocks as fit. All calls on one GPU share the array (contexts included) and it keeps the last call's contents (the
 * cursor is cleared on the stream by every call): calls that overlap on the GPU race, so issue them on one context / in stream order.
 * The first call per device resolves the array's address (cudaGetSymbolAddress); make it before any stream capture.
 * No counterpart in the reference. */
#define ATRAY_STATIC_SCRATCH_BYTES (16u << 20)
int atray_trace_rays_static(atray_ctx* ctx, const float* rays_dev, uint64_t n, int32_t fp_mode, uint32_t* prim_dev, float* t_dev);
int atray_get_counters(atray_ctx* ctx, AtrayCounters* out);

/* device memory helpers for embedders without a CUDA runtime binding of their own */
int atray_dev_alloc(atray_ctx* ctx, size_t bytes, void** out);
int atray_dev_free(atray_ctx* ctx, void* p);
int atray_dev_copy_to_host(atray_ctx* ctx, void* host, const void* dev, size_t bytes);
int atray_dev_copy_from_host(atray_ctx* ctx, void* dev, const void* host, size_t bytes);
/* asynchronous memset on the context stream (bench.py uses it to flush L2 between timed frames) */
int atray_dev_memset(atray_ctx* ctx, void* dev, int value, size_t bytes);
/* Write_To_File(Texture&, const char*) — engine/tools/texture.cpp:66-115, 147-158: byte-identical 32-bit BMP (14-byte
 * file header, 56-byte DIB header with BI_BITFIELDS masks, rows bottom-up as stored) into "<file_name>_<id>.bmp" for the
 * first id whose file does not exist. `written_name` (optional, `cap` bytes) receives the name used. Host only. */
int atray_write_bmp(const uint32_t* bgrx, uint32_t width, uint32_t height, const char* file_name, char* written_name, size_t cap);
int atray_host_alloc_pinned(size_t bytes, void** out);
int atray_host_free_pinned(void* p);
/* Page-locks memory the caller already owns (e.g. a bitmap in shared memory that several rank processes render into:
 * with AtrayRenderOpts.part.world > 1, atray_render_begin writes exactly this rank's scanlines of the host frame, so all
 * ranks can be handed the same bitmap and their copies run in parallel over their own PCIe links). */
int atray_host_register(void* p, size_t bytes);
int atray_host_unregister(void* p);
/* cross-process gather target (multi-GPU, one process per GPU): export a device allocation of this context and
 * open it in a peer process; the opened pointer can be used as AtrayFrame planes of atray_render_device
 * (the kernel stores straight into GPU 0's frame over NVLink). handle64 = 64 bytes. */
int atray_ipc_export(atray_ctx* ctx, void* dev, unsigned char* handle64);
int atray_ipc_open(atray_ctx* ctx, const unsigned char* handle64, void** out);
int atray_ipc_close(atray_ctx* ctx, void* p);

/* Shared per-frame view pass for N GPUs (N <= 8) that render one frame together, scene replicated on every GPU: each rank
 * computes the pixel ranges of 1/N of the triangle slots and stores them into the tables of ALL ranks with plain peer stores
 * over NVLink (k_view_bounds), then raises a flag per rank that a one-warp kernel on each peer's stream waits for —
 * the part of a frame that does not shrink with N when every rank computes it alone. No counterpart in the reference (its
 * workers share one address space, renderer.cpp:371-399).
 *   atray_view_share_create : after atray_scene_upload; allocates this context's block for frames up to max_width x
 *                             max_height and returns its device pointer (export it with atray_ipc_export for other processes);
 *   atray_view_share_connect: peer_blocks[world] = the blocks of all ranks as pointers valid in THIS process (own block at
 *                             [rank]; atray_ipc_open of a peer's handle, or the peer context's pointer after
 *                             atray_ctx_enable_peer in a single process). From then on frames rendered with
 *                             opts.part = {rank, world, …} take the shared pass; all ranks must render the same frames.
 *   atray_view_share_off    : back to every rank computing the whole pass itself. */
int atray_view_share_create(atray_ctx* ctx, int32_t max_width, int32_t max_height, void** block, size_t* bytes);
int atray_view_share_connect(atray_ctx* ctx, int32_t rank, int32_t world, void* const* peer_blocks);
int atray_view_share_off(atray_ctx* ctx);
/* queues "every rank has reached this point" on the context stream (a flag per rank + stream waits, no host round trip):
 * frames queued after it start together on all GPUs. No-op when the shared pass is off. */
int atray_view_share_rendezvous(atray_ctx* ctx);
/* drains the stream and reports ATRAY_ECUDA if a wait for a peer's flag ever timed out (2 s): a rank died or rendered other frames */
int atray_view_share_check(atray_ctx* ctx);
/* single process, several contexts: lets ctx's device store into other's device memory (cudaDeviceEnablePeerAccess) */
int atray_ctx_enable_peer(atray_ctx* ctx, atray_ctx* other);
/* verification aid: run the view pass of every later frame as `shards` shard launches on this GPU alone (0 = off) */
int atray_debug_view_shards(atray_ctx* ctx, int32_t shards);

/* Several GPUs of one node from ONE process — the shape of the reference's tile pool (one address space, one bitmap, N workers:
 * renderer.cpp:371-455, thread_pool.h:14-26) with GPUs as the workers: a context per device, the scene replicated, band b of
 * band_rows scanlines rendered by context b % n, every context computing the per-frame view pass of its own bands
 * (ATRAY_MULTI_SHARE=1 shares that pass between the contexts over NVLink instead, atray_view_share_* — measured slower on 2, 4 and 8
 * B200s, so it is opt-in).
 *   atray_ctx_create_multi   : devices[n] (NULL: spread over the visible devices), 1 <= n <= 8;
 *   atray_multi_ctx          : context i, for the per-context calls (counters, primitives, kernel times …);
 *   atray_multi_scene_upload : atray_scene_upload on every context (one host thread per device);
 *   atray_multi_render_begin / _poll / _progress / _cancel: as the atray_render_* calls; opts->part.rank / world are overwritten.
 *                              host_out: EVERY context copies exactly its scanlines into the same host planes (the planes that are
 *                              not NULL: bgrx only, or all three), each over its own GPU's PCIe link; the frame is complete when
 *                              poll returns 0. Progress counts the scanlines of the contexts that have finished;
 *   atray_multi_render_device: asynchronous; device_out planes live on ONE device (normally context 0's, atray_dev_alloc) and the
 *                              other contexts' kernels store their bands into them over NVLink; atray_multi_sync drains all. */
typedef struct atray_multi atray_multi;
int atray_ctx_create_multi(const int32_t* devices, int32_t n, atray_multi** out);
void atray_multi_destroy(atray_multi* m);
int atray_multi_count(const atray_multi* m);
atray_ctx* atray_multi_ctx(atray_multi* m, int32_t i);
int atray_multi_device(const atray_multi* m, int32_t i);
int atray_multi_scene_upload(atray_multi* m, const atray_model* model, const AtrayMaterial* sky, const AtrayMaterial* model_mat);
int atray_multi_scene_set_primitives(atray_multi* m, const AtraySphere* spheres, uint32_t n_spheres, const AtrayPlane* planes,
                                     uint32_t n_planes);
int atray_multi_scene_use_kd_tree(atray_multi* m, int use_kd_tree);
int atray_multi_render_begin(atray_multi* m, const AtrayCamera* cam, const AtrayRenderOpts* opts, const AtrayFrame* host_out);
int atray_multi_render_poll(atray_multi* m, uint32_t timeout_ms);
int atray_multi_render_progress(atray_multi* m, int32_t* rows_done, int32_t* rows_total);
int atray_multi_render_cancel(atray_multi* m);
int atray_multi_render_device(atray_multi* m, const AtrayCamera* cam, const AtrayRenderOpts* opts, const AtrayFrame* device_out);
int atray_multi_sync(atray_multi* m);

/* K0 microbenchmarks for the roofline denominators: FP32 issue slots/s and L2 read GB/s. */
int atray_microbench(atray_ctx* ctx, double* fp32_lane_ops_per_s, double* l2_read_gbs);

#ifdef __cplusplus
}
#endif
#endif
