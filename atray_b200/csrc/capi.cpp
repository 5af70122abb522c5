// C ABI (include/atray_b200.h) over the host data model (host_model.cpp) and the sm_100a kernels (render.cu).
// No CPU fallback: every device entry point fails with ATRAY_ECUDA if CUDA or the device is unavailable.
#include "../../include/atray_b200.h"
#include "device_types.h"
#include "host_model.h"

#include <cerrno>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

using namespace atray;

struct atray_model { HostModel m; };

struct atray_ctx
{
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    LaunchCfg cfg{};
    float4* d_nodes = nullptr;
    float4* d_tris = nullptr;
    uint32_t* d_inner_order = nullptr;
    uint32_t* d_level_start = nullptr;
    DevScene scene{};
    bool has_scene = false;
    DevScratch scratch{};
    DevView view{};              // per-frame film rectangles + tile stamps (k_view_bounds)
    size_t stamp_cap = 0;        // tiles the stamp buffer holds
    size_t bin_cap = 0;          // bins the bin lists hold
    // Shared view pass (atray_view_share_*): one device block per rank holding two copies (even / odd frames) of the tables
    // k_view_bounds fills — slot / packet / chunk ranges, tile stamps, tile costs, the all-busy flag — and a flag word per
    // rank. Every rank computes 1/world of the slots and stores them into every peer's block over NVLink.
    struct Share
    {
        bool on = false;
        int rank = 0, world = 1;
        char* block = nullptr;
        size_t bytes = 0;
        char* peer[kMaxPeers] = {};
        size_t off_tri = 0, off_pkt = 0, off_chunk = 0, off_stamps = 0, off_cost = 0, off_all = 0, buf_stride = 0, off_flags = 0;
        size_t tiles_cap = 0;
        uint32_t n_slots = 0;
        uint32_t seq = 0;        // frames rendered through the shared pass (the stamp value and the flag value of a frame)
        uint32_t rdv = 0;        // rendezvous count (atray_view_share_rendezvous)
        int debug_shards = 0;    // verification aid: run the pass as this many shard launches on this GPU alone
    } share;
    uint2* d_node_pk = nullptr;  // FlatScene::node_pk
    uint32_t frame_seq = 0;      // stamp value of the last frame (0 = buffer freshly cleared)
    // film-space rasteriser (raster.cuh): scene tables, per-frame tables / keys / lists
    uint32_t* d_packet_leaf = nullptr; uint32_t* d_leaf_seq = nullptr; uint32_t* d_seq_leaf = nullptr;
    uint32_t* d_parent = nullptr; uint32_t* d_pos_face = nullptr;
    DevRaster rs{};
    size_t key_cap_px = 0;           // pixels the key planes hold (all-ones between frames)
    int tab_w = 0, tab_h = 0;        // capacity of the film tables
    int use_raster = 0;              // ATRAY_RASTER=1 rasterises every AA-off frame (otherwise per frame: ATRAY_RENDER_RASTER)
    float4* d_faces = nullptr;       // face table of the no-tree branch (renderer.cpp:58-82)
    float4* d_fnorm = nullptr;       // per-face normal sources of the multi-bounce path
    float4* d_spheres = nullptr;     // analytic primitives (atray_scene_set_primitives)
    float4* d_planes = nullptr;
    unsigned long long* d_casts = nullptr;   // ray_casts of the last path frame (renderer.cpp:260)
    bool casts_valid = false;
    unsigned int* d_tile_counter = nullptr;
    // cancel: a device word the persistent loops poll (DevScene::cancel), raised by a 4-byte copy from pinned memory on a stream
    // of its own so that it overtakes the running kernel (atray_render_cancel)
    uint32_t* d_cancel = nullptr;
    uint32_t* h_cancel_one = nullptr;
    cudaStream_t ctl_stream = nullptr;
    uint64_t cancels = 0;            // cancel requests that found work on the stream
    DevCounters* d_counters = nullptr;
    bool counters_valid = false;
    // frame planes owned by the context (atray_render_begin)
    uint32_t* d_bgrx = nullptr; float* d_t = nullptr; uint32_t* d_prim = nullptr;
    size_t frame_cap_px = 0;
    // pinned staging for callers that pass pageable host memory
    uint32_t* h_bgrx = nullptr; float* h_t = nullptr; uint32_t* h_prim = nullptr;
    size_t stage_cap_px = 0;
    AtrayFrame pending_user{};   // where staged planes still have to be copied on completion
    size_t pending_px = 0;
    AtrayPartition pending_part{0, 1, 8};
    int32_t pending_w = 0;
    bool stage_bgrx = false, stage_t = false, stage_prim = false;
    bool in_flight = false;
    cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr, ev_done = nullptr;
    cudaEvent_t ev_kr = nullptr;     // between the per-frame view pass and the render kernel (atray_last_render_kernel_ms)
    cudaEvent_t ev_vb = nullptr;     // between k_view_bounds and k_view_nodes (atray_last_view_split_ms)
    bool have_split = false;
    bool have_kernel_time = false;
    // slab-pipelined copy-out (atray_render_begin): the copy stream waits on the render kernel's per-slab counters
    cudaStream_t copy_stream = nullptr;
    cudaStream_t copy_stream2 = nullptr;         // odd slabs of a pipelined frame when copy_streams == 2
    int copy_streams = 2;                        // slabs alternate between two copy streams (ATRAY_COPY_STREAMS=1: one)
    cudaEvent_t ev_reset = nullptr;              // recorded after k_view_bounds cleared the counters of the previous frame
    cudaEvent_t ev_slab[kMaxSlabs] = {};         // slab s has landed in the host frame
    int slabs_in_flight = 0;                     // slabs of the frame in flight (0 = unpipelined frame)
    int32_t slab_row_end[kMaxSlabs] = {};        // first scanline after slab s (host rows landed once ev_slab[s] fired)
    int32_t pending_rows = 0;                    // height of the frame in flight
    // Slab plan of a pipelined frame: a short first slab lets the copy-out start right after the view pass (frames bound by
    // the PCIe copy), a short last slab leaves little to copy once the kernel has retired (frames bound by tracing).
    // ATRAY_SLABS=n: n equal slabs (1 = no pipelining); ATRAY_SLAB_WEIGHTS="1,4,8,8,8,4,2": relative heights. tools/slab_sweep.sh.
    int slab_copy = 7;
    int slab_w[kMaxSlabs] = {1, 4, 8, 8, 8, 4, 2};   // (r02p, host frame of C4: 696 us against 723 us for 2,6,8,8,6,2; 709 for 1,3,6,8,8,6,3,1)
    typedef int (*WaitValue32)(cudaStream_t, unsigned long long, uint32_t, unsigned int);
    WaitValue32 wait_value32 = nullptr;          // cuStreamWaitValue32 through cudaGetDriverEntryPoint
    // ray-list buffers (atray_trace_rays)
    float* d_rays = nullptr; uint32_t* d_rprim = nullptr; float* d_rt = nullptr;
    size_t ray_cap = 0;
};

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
static int cuda_fail(cudaError_t e, const char* what)
{
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return ATRAY_ECUDA;
}
// ATRAY_SHARE_MODE=0: the first version of the shared view pass (stamps and costs stored to the owning rank by the computing rank)
static uint32_t share_ranges_only()
{
    const char* env = std::getenv("ATRAY_SHARE_MODE");
    return (env && std::atoi(env) == 0) ? 0u : 1u;
}
namespace atray { int set_error(int code, const char* msg) { return fail(code, msg ? msg : ""); } }   // for multi_ctx.cpp
#define CU(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return cuda_fail(e__, #call); } while (0)

extern "C" {

const char* atray_last_error(void) { return g_err.c_str(); }
int atray_version(void) { return 100; }

// ---- host side --------------------------------------------------------------------------------------------------
int atray_model_load_obj(const char* path, int loader_threads, atray_model** out)
{
    if (!path || !out) return fail(ATRAY_EINVAL, "null argument");
    atray_model* m = new atray_model();
    std::string err;
    int rc = load_obj(path, loader_threads, m->m, err);
    if (rc != 0) { delete m; return fail(rc == -2 ? ATRAY_EIO : ATRAY_EINVAL, err); }
    *out = m;
    return ATRAY_OK;
}

int atray_model_from_arrays(const float* vertices, uint32_t n_vertices, const int32_t* fvi, uint32_t n_faces, atray_model** out)
{
    if ((!vertices && n_vertices) || (!fvi && n_faces) || !out) return fail(ATRAY_EINVAL, "null argument");
    atray_model* m = new atray_model();
    m->m.vertices.resize(n_vertices);
    if (n_vertices) std::memcpy(m->m.vertices.data(), vertices, size_t(n_vertices) * 12);
    m->m.faces.resize(n_faces);
    for (uint32_t i = 0; i < n_faces; ++i)
        for (int k = 0; k < 3; ++k)
        {
            int32_t v = fvi[3 * size_t(i) + k];
            if (v < 0 || uint32_t(v) >= n_vertices) { delete m; return fail(ATRAY_EINVAL, "face index out of range"); }
            m->m.faces[i].v[k] = v;
            m->m.faces[i].vt[k] = -1;
            m->m.faces[i].vn[k] = -1;
        }
    *out = m;
    return ATRAY_OK;
}

void atray_model_free(atray_model* m) { delete m; }

int atray_model_compute_aabb(atray_model* m, float* out6)
{
    if (!m) return fail(ATRAY_EINVAL, "null model");
    compute_aabb(m->m);
    if (out6) std::memcpy(out6, &m->m.aabb, 24);
    return ATRAY_OK;
}
int atray_model_resize_scale(atray_model* m, float s)
{
    if (!m) return fail(ATRAY_EINVAL, "null model");
    if (!m->m.has_aabb) return fail(ATRAY_ESTATE, "resize_scale before compute_aabb");
    resize_scale(m->m, s);
    return ATRAY_OK;
}
int atray_model_translate_to(atray_model* m, const float* c)
{
    if (!m || !c) return fail(ATRAY_EINVAL, "null argument");
    if (!m->m.has_aabb) return fail(ATRAY_ESTATE, "translate_to before compute_aabb");
    translate_to(m->m, Vec3{c[0], c[1], c[2]});
    return ATRAY_OK;
}
int atray_model_build_tree(atray_model* m, int method, uint32_t cap)
{
    if (!m) return fail(ATRAY_EINVAL, "null model");
    if (method != ATRAY_DIV_CENTER && method != ATRAY_DIV_SAH) return fail(ATRAY_EINVAL, "unknown division method");
    std::string err;
    // the subtrees below the root are built by up to 8 host threads (same node array as the reference's sequential builder);
    // ATRAY_BUILD_THREADS=1 builds sequentially
    int threads = int(std::thread::hardware_concurrency());
    if (threads > 8) threads = 8;
    if (const char* env = std::getenv("ATRAY_BUILD_THREADS")) threads = std::atoi(env);
    int rc = build_tree(m->m, method, cap, err, threads < 1 ? 1 : threads);
    if (rc != 0) return fail(ATRAY_ELIMIT, err);
    return ATRAY_OK;
}
int atray_model_counts(const atray_model* m, uint32_t* c)
{
    if (!m || !c) return fail(ATRAY_EINVAL, "null argument");
    c[0] = uint32_t(m->m.vertices.size()); c[1] = uint32_t(m->m.normals.size()); c[2] = uint32_t(m->m.tex_coords.size());
    c[3] = uint32_t(m->m.faces.size()); c[4] = uint32_t(m->m.tree.size());
    uint32_t refs = 0, inner = 0, big = 0;
    for (const TreeNode& n : m->m.tree)
    {
        if (n.children_start) { ++inner; continue; }
        refs += uint32_t(n.prims.size());
        if (n.prims.size() > big) big = uint32_t(n.prims.size());
    }
    c[5] = refs; c[6] = inner; c[7] = big;
    return ATRAY_OK;
}
int atray_model_dump(const atray_model* m, float* v, float* n, float* t, int32_t* fv, int32_t* fvt, int32_t* fvn, float* aabb6)
{
    if (!m) return fail(ATRAY_EINVAL, "null model");
    const HostModel& h = m->m;
    if (v && !h.vertices.empty()) std::memcpy(v, h.vertices.data(), h.vertices.size() * 12);
    if (n && !h.normals.empty()) std::memcpy(n, h.normals.data(), h.normals.size() * 12);
    if (t && !h.tex_coords.empty()) std::memcpy(t, h.tex_coords.data(), h.tex_coords.size() * 12);
    for (size_t i = 0; i < h.faces.size(); ++i)
        for (int k = 0; k < 3; ++k)
        {
            if (fv) fv[3 * i + k] = h.faces[i].v[k];
            if (fvt) fvt[3 * i + k] = h.faces[i].vt[k];
            if (fvn) fvn[3 * i + k] = h.faces[i].vn[k];
        }
    if (aabb6) std::memcpy(aabb6, &h.aabb, 24);
    return ATRAY_OK;
}
int atray_model_tree_dump(const atray_model* m, float* boxes, int32_t* child, uint32_t* cnt, uint32_t* first, float* tri9, uint32_t* face)
{
    if (!m) return fail(ATRAY_EINVAL, "null model");
    uint32_t off = 0;
    for (size_t i = 0; i < m->m.tree.size(); ++i)
    {
        const TreeNode& n = m->m.tree[i];
        if (boxes) std::memcpy(boxes + 6 * i, &n.box, 24);
        if (child) child[i] = n.children_start;
        const uint32_t c = n.children_start ? 0u : uint32_t(n.prims.size());
        if (cnt) cnt[i] = c;
        if (first) first[i] = off;
        for (uint32_t k = 0; k < c; ++k)
        {
            if (tri9) std::memcpy(tri9 + 9 * size_t(off + k), &n.prims[k], 36);
            if (face) face[off + k] = n.prims[k].face;
        }
        off += c;
    }
    return ATRAY_OK;
}
int atray_model_flatten(const atray_model* m, uint32_t* n_nodes, uint32_t* n_slots, float* nodes, float* tris)
{
    if (!m) return fail(ATRAY_EINVAL, "null model");
    if (m->m.tree.empty()) return fail(ATRAY_ESTATE, "flatten before build_tree");
    FlatScene f;
    flatten(m->m, f);
    if (n_nodes) *n_nodes = f.n_nodes;
    if (n_slots) *n_slots = f.n_slots;
    if (nodes) std::memcpy(nodes, f.nodes.data(), f.nodes.size() * 4);
    if (tris && !f.tris.empty()) std::memcpy(tris, f.tris.data(), f.tris.size() * 4);
    return ATRAY_OK;
}

int atray_camera_set(AtrayCamera* cam, int32_t w, int32_t h, int32_t aa, uint32_t spp, int32_t bounce, const float* eye,
                     const float* facing, float h_fov)
{
    if (!cam || !eye || !facing) return fail(ATRAY_EINVAL, "null argument");
    if (w <= 0 || h <= 0 || spp == 0) return fail(ATRAY_EINVAL, "bad resolution or sample count");
    CameraBasis b = make_camera(w, h, Vec3{eye[0], eye[1], eye[2]}, Vec3{facing[0], facing[1], facing[2]}, h_fov);
    cam->width = w; cam->height = h; cam->anti_aliasing = aa; cam->samples_per_pixel = spp; cam->bounce_limit = bounce;
    cam->aspect_ratio = b.aspect;
    std::memcpy(cam->camera_z, &b.cz, 12); std::memcpy(cam->camera_x, &b.cx, 12); std::memcpy(cam->camera_y, &b.cy, 12);
    std::memcpy(cam->eye, &b.eye, 12); std::memcpy(cam->frame_center, &b.fc, 12);
    cam->h_fov = b.h_fov; cam->half_pixel_width = b.hpw; cam->half_pixel_height = b.hph;
    return ATRAY_OK;
}

// ---- device side ------------------------------------------------------------------------------------------------
int atray_device_count(int* n)
{
    if (!n) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaGetDeviceCount(n));
    return ATRAY_OK;
}

int atray_ctx_create(int device, atray_ctx** out)
{
    if (!out) return fail(ATRAY_EINVAL, "null argument");
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(ATRAY_ECUDA, "no such CUDA device");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(ATRAY_ECUDA, std::string("device is not sm_100-class: ") + prop.name);
    atray_ctx* c = new atray_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->cfg = render_launch_cfg(c->sm_count);
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_tile_counter, sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMalloc(&c->d_counters, sizeof(DevCounters));
    if (e == cudaSuccess) e = cudaMalloc(&c->d_casts, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&c->d_cancel, 256);
    if (e == cudaSuccess) e = cudaMemset(c->d_cancel, 0, 256);
    if (e == cudaSuccess) e = cudaMallocHost(&c->h_cancel_one, sizeof(uint32_t));
    if (e == cudaSuccess) { *c->h_cancel_one = 1u; c->scene.cancel = c->d_cancel; }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->ctl_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev_k0);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev_k1);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev_kr);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev_vb);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream2, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_reset, cudaEventDisableTiming);
    for (int k = 0; k < kMaxSlabs && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&c->ev_slab[k], cudaEventDisableTiming);
    if (e != cudaSuccess) { delete c; return cuda_fail(e, "context setup"); }
    {
        // stream memory operations (driver API) without linking libcuda: the copy stream waits on device counters
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess && fn)
            c->wait_value32 = reinterpret_cast<atray_ctx::WaitValue32>(fn);
        else
            cudaGetLastError();
        if (const char* env = std::getenv("ATRAY_RASTER")) c->use_raster = std::atoi(env) != 0;
        if (const char* env = std::getenv("ATRAY_SLABS"))
        {
            const int v = std::atoi(env);
            c->slab_copy = v < 1 ? 1 : (v > kMaxSlabs ? kMaxSlabs : v);
            for (int k = 0; k < kMaxSlabs; ++k) c->slab_w[k] = 0;
        }
        if (const char* env = std::getenv("ATRAY_COPY_STREAMS")) c->copy_streams = std::atoi(env) == 2 ? 2 : 1;
        if (const char* env = std::getenv("ATRAY_SLAB_WEIGHTS"))
        {
            int n = 0;
            for (int k = 0; k < kMaxSlabs; ++k) c->slab_w[k] = 0;
            for (const char* p = env; *p && n < kMaxSlabs;)
            {
                char* end = nullptr;
                const long v = std::strtol(p, &end, 10);
                if (end == p) break;
                c->slab_w[n++] = v < 1 ? 1 : int(v);
                p = *end == ',' ? end + 1 : end;
            }
            if (n > 0) c->slab_copy = n;
        }
    }
    *out = c;
    return ATRAY_OK;
}

static void free_scene(atray_ctx* c)
{
    cudaFree(c->d_nodes); cudaFree(c->d_tris); cudaFree(c->scratch.leaves); cudaFree(c->scratch.stack);
    cudaFree(c->view.tri_px); cudaFree(c->view.pkt_px); cudaFree(c->view.chunk_px); cudaFree(c->view.node_px);
    cudaFree(c->view.tri_rect); cudaFree(c->d_node_pk);
    cudaFree(c->d_inner_order); cudaFree(c->d_level_start);
    c->view.tri_px = c->view.pkt_px = c->view.chunk_px = c->view.node_px = nullptr;
    c->view.tri_rect = nullptr;
    c->d_node_pk = nullptr;
    c->d_inner_order = c->d_level_start = nullptr;
    c->d_nodes = c->d_tris = nullptr;
    cudaFree(c->d_fnorm); cudaFree(c->d_spheres); cudaFree(c->d_planes);
    c->d_fnorm = c->d_spheres = c->d_planes = nullptr;
    cudaFree(c->d_packet_leaf); cudaFree(c->d_leaf_seq); cudaFree(c->d_seq_leaf); cudaFree(c->d_parent); cudaFree(c->d_pos_face);
    c->d_packet_leaf = c->d_leaf_seq = c->d_seq_leaf = c->d_parent = c->d_pos_face = nullptr;
    cudaFree(c->rs.big);
    c->rs.big = nullptr; c->rs.big_cap = 0;
    cudaFree(c->d_faces);
    c->d_faces = nullptr;
    c->scratch = DevScratch{};
    c->scene = DevScene{};
    c->scene.cancel = c->d_cancel;
    c->has_scene = false;
    cudaFree(c->share.block);   // sized for the scene: a new scene needs a new atray_view_share_create
    c->share = atray_ctx::Share{};
}

void atray_ctx_destroy(atray_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    free_scene(c);
    cudaFree(c->d_tile_counter); cudaFree(c->d_counters); cudaFree(c->d_casts);
    cudaFree(c->view.stamps); cudaFree(c->view.all_flag); cudaFree(c->view.busy_list); cudaFree(c->view.busy_count);
    cudaFree(c->view.tile_cost); cudaFree(c->view.tile_log); cudaFree(c->view.wide_inner); cudaFree(c->view.big_list);
    cudaFree(c->view.bin_cnt); cudaFree(c->view.bin_leaf); cudaFree(c->view.bin_inner);
    cudaFree(c->share.block);
    cudaFree(c->rs.tabx); cudaFree(c->rs.taby); cudaFree(c->rs.key1); cudaFree(c->rs.key2); cudaFree(c->rs.hit_a);
    cudaFree(c->rs.hit_pos); cudaFree(c->rs.ctl);
    cudaFree(c->d_bgrx); cudaFree(c->d_t); cudaFree(c->d_prim);
    cudaFreeHost(c->h_bgrx); cudaFreeHost(c->h_t); cudaFreeHost(c->h_prim);
    cudaFree(c->d_rays); cudaFree(c->d_rprim); cudaFree(c->d_rt);
    if (c->ev_k0) cudaEventDestroy(c->ev_k0);
    if (c->ev_k1) cudaEventDestroy(c->ev_k1);
    if (c->ev_kr) cudaEventDestroy(c->ev_kr);
    if (c->ev_vb) cudaEventDestroy(c->ev_vb);
    if (c->ev_done) cudaEventDestroy(c->ev_done);
    if (c->ev_reset) cudaEventDestroy(c->ev_reset);
    for (int k = 0; k < kMaxSlabs; ++k)
        if (c->ev_slab[k]) cudaEventDestroy(c->ev_slab[k]);
    if (c->ctl_stream) cudaStreamDestroy(c->ctl_stream);
    cudaFree(c->d_cancel);
    cudaFreeHost(c->h_cancel_one);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->copy_stream2) cudaStreamDestroy(c->copy_stream2);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int atray_scene_upload(atray_ctx* c, const atray_model* m, const AtrayMaterial* sky, const AtrayMaterial* mat)
{
    if (!c || !m || !sky || !mat) return fail(ATRAY_EINVAL, "null argument");
    if (m->m.tree.empty()) return fail(ATRAY_ESTATE, "scene_upload before build_tree");
    if (c->in_flight) return fail(ATRAY_ESTATE, "render in flight");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    free_scene(c);
    FlatScene f;
    flatten(m->m, f);
    CU(cudaMalloc(&c->d_nodes, f.nodes.size() * 4));
    CU(cudaMalloc(&c->d_tris, f.tris.size() * 4 + 16));
    CU(cudaMemcpy(c->d_nodes, f.nodes.data(), f.nodes.size() * 4, cudaMemcpyHostToDevice));
    if (!f.tris.empty()) CU(cudaMemcpy(c->d_tris, f.tris.data(), f.tris.size() * 4, cudaMemcpyHostToDevice));
    const int warps = render_warps_in_flight(c->cfg);
    c->scratch.leaf_cap = f.n_leaves + 2;
    c->scratch.stack_cap = f.n_nodes + 8;
    CU(cudaMalloc(&c->scratch.leaves, size_t(warps) * c->scratch.leaf_cap * sizeof(uint2)));
    CU(cudaMalloc(&c->scratch.stack, size_t(warps) * c->scratch.stack_cap * sizeof(uint32_t)));
    CU(cudaMalloc(&c->view.tri_px, size_t(f.n_slots ? f.n_slots : 1) * sizeof(uint2)));
    CU(cudaMalloc(&c->view.pkt_px, size_t(f.n_slots / 8 + 1) * sizeof(uint2)));
    CU(cudaMalloc(&c->view.chunk_px, size_t(f.n_slots / 256 + 1) * sizeof(uint2)));
    CU(cudaMalloc(&c->view.node_px, size_t(f.n_nodes) * sizeof(uint2)));
    CU(cudaMalloc(&c->d_node_pk, size_t(f.n_nodes) * sizeof(uint2)));
    CU(cudaMemcpy(c->d_node_pk, f.node_pk.data(), size_t(f.n_nodes) * sizeof(uint2), cudaMemcpyHostToDevice));
    c->scene.node_pk = c->d_node_pk;
    CU(cudaMalloc(&c->d_inner_order, (f.inner_order.size() + 1) * sizeof(uint32_t)));
    CU(cudaMalloc(&c->d_level_start, f.level_start.size() * sizeof(uint32_t)));
    if (!f.inner_order.empty())
        CU(cudaMemcpy(c->d_inner_order, f.inner_order.data(), f.inner_order.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_level_start, f.level_start.data(), f.level_start.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    {
        auto up = [](uint32_t** dst, const std::vector<uint32_t>& v) -> cudaError_t {
            cudaError_t e = cudaMalloc(dst, (v.size() + 1) * sizeof(uint32_t));
            if (e == cudaSuccess && !v.empty()) e = cudaMemcpy(*dst, v.data(), v.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
            return e;
        };
        CU(up(&c->d_packet_leaf, f.packet_leaf)); CU(up(&c->d_leaf_seq, f.leaf_seq)); CU(up(&c->d_seq_leaf, f.seq_leaf));
        CU(up(&c->d_parent, f.parent)); CU(up(&c->d_pos_face, f.pos_face));
        c->scene.packet_leaf = c->d_packet_leaf; c->scene.leaf_seq = c->d_leaf_seq; c->scene.seq_leaf = c->d_seq_leaf;
        c->scene.parent = c->d_parent; c->scene.pos_face = c->d_pos_face;
        c->rs.big_cap = f.n_slots ? f.n_slots : 1u;
        CU(cudaMalloc(&c->rs.big, size_t(c->rs.big_cap) * sizeof(uint2)));
    }
    c->scene.inner_order = c->d_inner_order;
    c->scene.level_start = c->d_level_start;
    c->scene.n_levels = uint32_t(f.level_start.size() - 1);
    c->scene.nodes = c->d_nodes;
    c->scene.tris = c->d_tris;
    c->scene.n_nodes = f.n_nodes;
    c->scene.n_slots = f.n_slots;
    c->scene.root_is_leaf = m->m.tree[0].children_start == 0 ? 1u : 0u;
    std::memcpy(c->scene.sky, sky->emission_color, 12);
    std::memcpy(c->scene.mdl, mat->emission_color, 12);
    std::memcpy(c->scene.mdl_refl, mat->reflection_color, 12);
    c->scene.mdl_scatter = mat->scatter;
    {
        // normal sources of get_intersection_data (renderer.cpp:124-148). With vertex normals: the face's three normals
        // (interpolated per hit). Without: cross(v0 - v1, v0 - v2), the same f32 operations the reference runs per hit.
        const HostModel& hm = m->m;
        const bool smooth = !hm.normals.empty();
        std::vector<float> fn(size_t(hm.faces.size()) * (smooth ? 12 : 4) + 4, 0.f);
        for (size_t i = 0; i < hm.faces.size(); ++i)
        {
            const Face& f = hm.faces[i];
            if (smooth)
                for (int k = 0; k < 3; ++k)
                {
                    const int32_t ni = f.vn[k];
                    if (ni < 0 || size_t(ni) >= hm.normals.size()) continue;   // the reference would read out of bounds
                    float* q = &fn[(3 * i + k) * 4];
                    q[0] = hm.normals[ni].x; q[1] = hm.normals[ni].y; q[2] = hm.normals[ni].z;
                }
            else
            {
                const Vec3 a = hm.vertices[f.v[0]], b = hm.vertices[f.v[1]], cc = hm.vertices[f.v[2]];
                const Vec3 ab = {a.x - b.x, a.y - b.y, a.z - b.z}, ac = {a.x - cc.x, a.y - cc.y, a.z - cc.z};
                float* q = &fn[i * 4];
                q[0] = ab.y * ac.z - ab.z * ac.y; q[1] = ab.z * ac.x - ab.x * ac.z; q[2] = ab.x * ac.y - ab.y * ac.x;
            }
        }
        CU(cudaMalloc(&c->d_fnorm, fn.size() * 4));
        CU(cudaMemcpy(c->d_fnorm, fn.data(), fn.size() * 4, cudaMemcpyHostToDevice));
        c->scene.fnorm = c->d_fnorm;
        c->scene.smooth = smooth ? 1u : 0u;
    }
    {
        // face table of the no-tree branch: the three vertices of face j as {a, b - a, c - a}; the two subtractions are the
        // first operations of get_triangle_ray_intersection_culled (model.h:78-79), done here once in f32
        const HostModel& hm = m->m;
        std::vector<float> ft(size_t(hm.faces.size()) * 12 + 4, 0.f);
        for (size_t i = 0; i < hm.faces.size(); ++i)
        {
            const Face& fc = hm.faces[i];
            const Vec3 a = hm.vertices[fc.v[0]], b = hm.vertices[fc.v[1]], cc = hm.vertices[fc.v[2]];
            float* q = &ft[i * 12];
            q[0] = a.x; q[1] = a.y; q[2] = a.z;
            q[4] = b.x - a.x; q[5] = b.y - a.y; q[6] = b.z - a.z;
            q[8] = cc.x - a.x; q[9] = cc.y - a.y; q[10] = cc.z - a.z;
        }
        CU(cudaMalloc(&c->d_faces, ft.size() * 4));
        CU(cudaMemcpy(c->d_faces, ft.data(), ft.size() * 4, cudaMemcpyHostToDevice));
        c->scene.faces = c->d_faces;
        c->scene.n_faces = uint32_t(hm.faces.size());
        c->scene.brute = 0;
        const Box bb = hm.aabb;
        c->scene.aabb_lo[0] = bb.lo.x; c->scene.aabb_lo[1] = bb.lo.y; c->scene.aabb_lo[2] = bb.lo.z;
        c->scene.aabb_hi[0] = bb.hi.x; c->scene.aabb_hi[1] = bb.hi.y; c->scene.aabb_hi[2] = bb.hi.z;
    }
    c->has_scene = true;
    return ATRAY_OK;
}

int atray_scene_use_kd_tree(atray_ctx* c, int use_kd_tree)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "use_kd_tree before scene_upload");
    if (c->in_flight) return fail(ATRAY_ESTATE, "render in flight");
    c->scene.brute = use_kd_tree ? 0u : 1u;
    return ATRAY_OK;
}

int atray_scene_set_primitives(atray_ctx* c, const AtraySphere* spheres, uint32_t n_spheres, const AtrayPlane* planes,
                               uint32_t n_planes)
{
    if (!c || (n_spheres && !spheres) || (n_planes && !planes)) return fail(ATRAY_EINVAL, "null argument");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "set_primitives before scene_upload");
    if (c->in_flight) return fail(ATRAY_ESTATE, "render in flight");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(c->d_spheres); cudaFree(c->d_planes);
    c->d_spheres = c->d_planes = nullptr;
    c->scene.spheres = c->scene.planes = nullptr;
    c->scene.n_spheres = c->scene.n_planes = 0;
    auto pack = [](std::vector<float>& v, const float* head3, float w, const AtrayMaterial& m) {
        const float rec[12] = {head3[0], head3[1], head3[2], w, m.emission_color[0], m.emission_color[1], m.emission_color[2],
                               m.scatter, m.reflection_color[0], m.reflection_color[1], m.reflection_color[2], 0.f};
        v.insert(v.end(), rec, rec + 12);
    };
    if (n_spheres)
    {
        std::vector<float> v;
        for (uint32_t i = 0; i < n_spheres; ++i) pack(v, spheres[i].center, spheres[i].radius, spheres[i].material);
        CU(cudaMalloc(&c->d_spheres, v.size() * 4));
        CU(cudaMemcpy(c->d_spheres, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    }
    if (n_planes)
    {
        std::vector<float> v;
        for (uint32_t i = 0; i < n_planes; ++i)
        {
            // prep_scene normalises every plane normal (renderer.cpp:267-270): n * (1/sqrt((x*x + y*y) + z*z))
            const float* n = planes[i].normal;
            const float inv = 1.0f / std::sqrt((n[0] * n[0]) + (n[1] * n[1]) + (n[2] * n[2]));
            const float nn[3] = {n[0] * inv, n[1] * inv, n[2] * inv};
            pack(v, nn, planes[i].distance, planes[i].material);
        }
        CU(cudaMalloc(&c->d_planes, v.size() * 4));
        CU(cudaMemcpy(c->d_planes, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    }
    c->scene.spheres = c->d_spheres; c->scene.planes = c->d_planes;
    c->scene.n_spheres = n_spheres; c->scene.n_planes = n_planes;
    return ATRAY_OK;
}

int atray_last_ray_casts(atray_ctx* c, uint64_t* ray_casts)
{
    if (!c || !ray_casts) return fail(ATRAY_EINVAL, "null argument");
    if (!c->casts_valid) return fail(ATRAY_ESTATE, "no multi-bounce frame has been rendered");
    CU(cudaSetDevice(c->device));
    unsigned long long v = 0;
    CU(cudaMemcpyAsync(&v, c->d_casts, sizeof v, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *ray_casts = v;
    return ATRAY_OK;
}

static int check_render_args(atray_ctx* c, const AtrayCamera* cam, const AtrayRenderOpts* o, const AtrayFrame* out)
{
    if (!c || !cam || !o || !out) return fail(ATRAY_EINVAL, "null argument");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "render before scene_upload");
    if (cam->width <= 0 || cam->height <= 0 || cam->samples_per_pixel == 0) return fail(ATRAY_EINVAL, "bad camera");
    if (cam->bounce_limit < 1 || cam->bounce_limit > 64) return fail(ATRAY_EINVAL, "bounce_limit must be in 1..64");
    if (o->part.world < 1 || o->part.rank < 0 || o->part.rank >= o->part.world) return fail(ATRAY_EINVAL, "bad partition");
    if (o->part.band_rows < 4 || (o->part.band_rows & 3)) return fail(ATRAY_EINVAL, "band_rows must be a positive multiple of 4");
    return ATRAY_OK;
}

static DevCamera to_dev(const AtrayCamera* cam)
{
    DevCamera d;
    d.W = cam->width; d.H = cam->height; d.aa = cam->anti_aliasing; d.spp = cam->samples_per_pixel;
    d.aspect = cam->aspect_ratio; d.h_fov = cam->h_fov; d.hpw = cam->half_pixel_width; d.hph = cam->half_pixel_height;
    std::memcpy(d.eye, cam->eye, 12); std::memcpy(d.fc, cam->frame_center, 12);
    std::memcpy(d.cx, cam->camera_x, 12); std::memcpy(d.cy, cam->camera_y, 12);
    d.bounce = cam->bounce_limit;
    return d;
}

// Per-frame view parameters of the film-space culling (view_bounds.h), computed in double from the camera the kernel
// uses. The ideal ray of film point (fx, fy) is eye + s*(F + cx*fx + cy*fy) with F = frame_center - eye and the float
// camera vectors taken as exact reals; minv inverts [cx cy F]. `ang` bounds the angle between that ideal line and the
// float ray camera_ray() produces: pixel_pos has a componentwise error <= u(2|fc_i| + 3|cx_i fx| + 3|cy_i fy|), the
// subtraction of the eye adds u|d_i|, normalisation (three products, one division, three products) <= 3u relative; the
// ideal direction is at least 1/||minv||_F long. The bound is doubled.
static ViewParams make_view_params(const AtrayCamera* cam)
{
    ViewParams vp{};
    const double U = 5.9604644775390625e-8;
    double M[3][3];
    for (int i = 0; i < 3; ++i)
    {
        vp.eye[i] = cam->eye[i];
        M[i][0] = cam->camera_x[i];
        M[i][1] = cam->camera_y[i];
        M[i][2] = double(cam->frame_center[i]) - double(cam->eye[i]);
    }
    const double det = M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) +
                       M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
    vp.W = cam->width; vp.H = cam->height;
    vp.tiles_x = (cam->width + 7) >> 3; vp.tiles_y = (cam->height + 3) >> 2;
    vp.kx = double(cam->h_fov) * double(cam->aspect_ratio);
    vp.valid = 0;
    if (!(std::fabs(det) > 1e-3) || !std::isfinite(det) || !std::isfinite(vp.kx) || !(std::fabs(vp.kx) > 1e-9)) return vp;
    const double id = 1.0 / det;
    vp.minv[0][0] = (M[1][1] * M[2][2] - M[1][2] * M[2][1]) * id; vp.minv[0][1] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) * id;
    vp.minv[0][2] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) * id; vp.minv[1][0] = (M[1][2] * M[2][0] - M[1][0] * M[2][2]) * id;
    vp.minv[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) * id; vp.minv[1][2] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) * id;
    vp.minv[2][0] = (M[1][0] * M[2][1] - M[1][1] * M[2][0]) * id; vp.minv[2][1] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) * id;
    vp.minv[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) * id;
    double fro = 0.0;
    for (int r = 0; r < 3; ++r)
    {
        double n2 = 0.0;
        for (int k = 0; k < 3; ++k) n2 += vp.minv[r][k] * vp.minv[r][k];
        vp.row[r] = std::sqrt(n2) * (1.0 + 1e-9);
        fro += n2;
    }
    const double aa_x = cam->anti_aliasing ? std::fabs(double(cam->half_pixel_width)) : 0.0;
    const double aa_y = cam->anti_aliasing ? std::fabs(double(cam->half_pixel_height)) : 0.0;
    const double fx_max = std::fabs(vp.kx) * (1.0 + 1e-6) + aa_x, fy_max = 1.0 + 1e-6 + aa_y;
    vp.fx_max = fx_max; vp.fy_max = fy_max;
    double e2 = 0.0, eye_inf = 0.0;
    for (int i = 0; i < 3; ++i)
    {
        const double e = U * (2.0 * std::fabs(double(cam->frame_center[i])) + 3.0 * std::fabs(M[i][0]) * fx_max +
                              3.0 * std::fabs(M[i][1]) * fy_max);
        e2 += e * e;
        eye_inf = std::fmax(eye_inf, std::fabs(vp.eye[i]));
    }
    const double dmin = 1.0 / std::sqrt(fro);
    vp.ang = 2.0 * (std::sqrt(e2) / dmin + 4.0 * U);
    // film→pixel inversion slack (view_bounds.h: film_rect_px): the float film coordinate of a pixel is within 2.6u*W pixels
    // of the real one; a jittered sample (renderer.cpp:340-341, |rand_bi| <= 1) moves the film point by at most
    // |half_pixel_width| = that many film units = |hpw|*W/(2|kx|) pixels, |half_pixel_height|*H/2 pixels in y — taken from
    // the camera the kernel uses, whatever the caller put there
    vp.inv_kx = 1.0 / vp.kx;
    vp.px_slack_x = 2e-6 * vp.W + 1e-9 + (cam->anti_aliasing ? aa_x * vp.W * 0.5 * std::fabs(vp.inv_kx) * (1.0 + 1e-6) : 0.0);
    vp.px_slack_y = 2e-6 * vp.H + 1e-9 + (cam->anti_aliasing ? aa_y * vp.H * 0.5 * (1.0 + 1e-6) : 0.0);
    bool fin = std::isfinite(vp.ang) && std::isfinite(eye_inf) && eye_inf < 1073741824.0 && std::isfinite(vp.px_slack_x) &&
               std::isfinite(vp.px_slack_y) && cam->width < 65535 && cam->height < 65535;   // pixel ranges are packed in 16 bits
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < 3; ++k) fin = fin && std::isfinite(vp.minv[r][k]);
    vp.valid = fin ? 1 : 0;
    return vp;
}

// Per-frame buffers of the rasteriser: film tables for W and H, two 64-bit key planes (all-ones between frames; k_resolve
// restores what a frame consumed), the hit list (4 entries per pixel, at least 1 Mi) and three counters.
static int ensure_raster_buffers(atray_ctx* c, const AtrayCamera* cam)
{
    const size_t px = size_t(cam->width) * size_t(cam->height);
    if (!c->rs.ctl) CU(cudaMalloc(&c->rs.ctl, 4 * sizeof(uint32_t)));
    if (cam->width > c->tab_w)
    {
        cudaFree(c->rs.tabx); c->rs.tabx = nullptr; c->tab_w = 0;
        CU(cudaMalloc(&c->rs.tabx, size_t(cam->width) * sizeof(float4)));
        c->tab_w = cam->width;
    }
    if (cam->height > c->tab_h)
    {
        cudaFree(c->rs.taby); c->rs.taby = nullptr; c->tab_h = 0;
        CU(cudaMalloc(&c->rs.taby, size_t(cam->height) * sizeof(float4)));
        c->tab_h = cam->height;
    }
    if (px > c->key_cap_px)
    {
        cudaFree(c->rs.key1); cudaFree(c->rs.key2); cudaFree(c->rs.hit_a); cudaFree(c->rs.hit_pos);
        c->rs.key1 = c->rs.key2 = nullptr; c->rs.hit_a = nullptr; c->rs.hit_pos = nullptr; c->key_cap_px = 0;
        size_t cap = px * 4;
        if (cap < (size_t(1) << 20)) cap = size_t(1) << 20;
        if (cap > (size_t(1) << 28)) cap = size_t(1) << 28;
        if (const char* env = std::getenv("ATRAY_HIT_CAP")) cap = size_t(std::atoll(env)) ? size_t(std::atoll(env)) : cap;   // tests: force the overflow path
        CU(cudaMalloc(&c->rs.key1, px * 8)); CU(cudaMalloc(&c->rs.key2, px * 8));
        CU(cudaMalloc(&c->rs.hit_a, cap * sizeof(uint4))); CU(cudaMalloc(&c->rs.hit_pos, cap * sizeof(uint32_t)));
        CU(cudaMemsetAsync(c->rs.key1, 0xff, px * 8, c->stream)); CU(cudaMemsetAsync(c->rs.key2, 0xff, px * 8, c->stream));
        c->rs.hit_cap = uint32_t(cap);
        c->key_cap_px = px;
    }
    return ATRAY_OK;
}

// Queues one frame on the context stream. `want_slabs` > 1 asks for a slab table (culled kernel only); *slabs_out tells
// the caller what it got (n = 0: unculled kernel, no per-slab counters).
static int enqueue_render(atray_ctx* c, const AtrayCamera* cam, const AtrayRenderOpts* o, const DevOut& out, int want_slabs = 1,
                          DevSlabs* slabs_out = nullptr)
{
    if (slabs_out) slabs_out->n = 0;
    DevCounters* ctr = nullptr;
    if (o->count_work)
    {
        CU(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
        ctr = c->d_counters;
        c->counters_valid = true;
    }
    DevPartition p{o->part.rank, o->part.world, o->part.band_rows};
    if (cam->bounce_limit != 1 || c->scene.n_spheres || c->scene.n_planes)   // (the path kernel honours scene.brute too)
    {
        // multi-bounce / analytic primitives: the full get_intersection_data + cast_ray kernel
        CU(cudaEventRecord(c->ev_k0, c->stream));
        CU(launch_path(c->scene, to_dev(cam), p, out, c->scratch, c->d_tile_counter, c->d_casts, o->fp_mode == ATRAY_FP_FAST,
                       o->jitter_seed, c->cfg, c->stream));
        CU(cudaEventRecord(c->ev_k1, c->stream));
        c->have_kernel_time = true;
        c->have_split = false;
        c->casts_valid = true;
        return ATRAY_OK;
    }
    ViewParams vp = make_view_params(cam);
    const bool cull = !(o->flags & ATRAY_RENDER_NO_CULL) && vp.valid && !c->scene.brute;   // no tree, no film-space culling
    if (cull)
    {
        const size_t tiles = size_t(vp.tiles_x) * size_t(vp.tiles_y);
        if (tiles > c->stamp_cap || c->frame_seq == 0xffffffffu)
        {
            if (tiles > c->stamp_cap)
            {
                cudaFree(c->view.stamps); cudaFree(c->view.busy_list); cudaFree(c->view.tile_cost);
                c->view.stamps = nullptr; c->view.busy_list = nullptr; c->view.tile_cost = nullptr; c->stamp_cap = 0;
                CU(cudaMalloc(&c->view.stamps, tiles * sizeof(uint32_t)));
                CU(cudaMalloc(&c->view.tile_cost, tiles * sizeof(uint32_t)));
                CU(cudaMalloc(&c->view.busy_list, tiles * kCostBuckets * sizeof(uint32_t)));
                CU(cudaMemsetAsync(c->view.tile_cost, 0, tiles * sizeof(uint32_t), c->stream));
                c->stamp_cap = tiles;
                c->view.bucket_stride = uint32_t(tiles);
                cudaFree(c->view.tile_log);
                c->view.tile_log = nullptr;
                if (const char* env = std::getenv("ATRAY_TILE_LOG"))
                    if (std::atoi(env) != 0)
                    {
                        CU(cudaMalloc(&c->view.tile_log, tiles * sizeof(uint4)));
                        CU(cudaMemsetAsync(c->view.tile_log, 0, tiles * sizeof(uint4), c->stream));
                    }
                c->view.cost_sample = 3;
                if (const char* env = std::getenv("ATRAY_COST_SAMPLE")) c->view.cost_sample = std::atoi(env) < 0 ? 0 : (std::atoi(env) > 5 ? 5 : std::atoi(env));
                c->view.cost_shift = -c->view.cost_sample;
                if (const char* env = std::getenv("ATRAY_COST_SHIFT")) c->view.cost_shift = std::atoi(env);
            }
            if (!c->view.all_flag) CU(cudaMalloc(&c->view.all_flag, sizeof(uint32_t)));
            if (!c->view.wide_inner) CU(cudaMalloc(&c->view.wide_inner, size_t(kWideCap) * 2 * sizeof(uint4)));
            if (!c->view.big_list) CU(cudaMalloc(&c->view.big_list, size_t(kBigCap) * sizeof(uint2)));
            if (!c->view.busy_count)
            {
                CU(cudaMalloc(&c->view.busy_count, kCtlWords * sizeof(uint32_t)));
                CU(cudaMemsetAsync(c->view.busy_count, 0, kCtlWords * sizeof(uint32_t), c->stream));
            }
            CU(cudaMemsetAsync(c->view.stamps, 0, c->stamp_cap * sizeof(uint32_t), c->stream));
            CU(cudaMemsetAsync(c->view.all_flag, 0, sizeof(uint32_t), c->stream));
            c->frame_seq = 0;
        }
        vp.frame_id = ++c->frame_seq;
        c->view.frame_id = vp.frame_id;
        c->view.tiles_x = vp.tiles_x;
        // bins of 64 x 32 pixels with flat node lists (k_view_nodes); a bin whose list outgrows its capacity sends its tiles
        // through the unculled core, so the capacities only bound memory: 192 leaves + 64 inner nodes = 11 KB per bin
        c->view.bins_x = (cam->width + (1 << kBinShiftX) - 1) >> kBinShiftX;
        c->view.bins_y = (cam->height + (1 << kBinShiftY) - 1) >> kBinShiftY;
        const size_t bins = size_t(c->view.bins_x) * size_t(c->view.bins_y);
        if (bins > c->bin_cap)
        {
            cudaFree(c->view.bin_cnt); cudaFree(c->view.bin_leaf); cudaFree(c->view.bin_inner);
            c->view.bin_cnt = nullptr; c->view.bin_leaf = nullptr; c->view.bin_inner = nullptr; c->bin_cap = 0;
            c->view.bin_leaf_cap = 192; c->view.bin_inner_cap = 64;
            if (const char* env = std::getenv("ATRAY_BIN_LEAF_CAP")) c->view.bin_leaf_cap = uint32_t(std::atoi(env) > 0 ? std::atoi(env) : 1);
            if (const char* env = std::getenv("ATRAY_BIN_INNER_CAP")) c->view.bin_inner_cap = uint32_t(std::atoi(env) > 0 ? std::atoi(env) : 1);
            CU(cudaMalloc(&c->view.bin_cnt, bins * 2 * sizeof(uint32_t)));
            CU(cudaMalloc(&c->view.bin_leaf, bins * c->view.bin_leaf_cap * 3 * sizeof(uint4)));
            CU(cudaMalloc(&c->view.bin_inner, bins * c->view.bin_inner_cap * 2 * sizeof(uint4)));
            c->bin_cap = bins;
        }
    }
    // opt-in, AA off and a real tree: the frame is rasterised triangle by triangle (raster.cuh) — same bits, no per-tile traversal
    const bool raster = cull && (c->use_raster || (o->flags & ATRAY_RENDER_RASTER)) && !cam->anti_aliasing && !c->scene.root_is_leaf;
    if (raster)
    {
        int rc = ensure_raster_buffers(c, cam);
        if (rc) return rc;
        if (!c->view.tri_rect) CU(cudaMalloc(&c->view.tri_rect, size_t(c->scene.n_slots ? c->scene.n_slots : 1) * sizeof(float4)));
    }
    CU(cudaEventRecord(c->ev_k0, c->stream));
    if (raster)
    {
        const DevSlabs slabs = make_slabs(to_dev(cam), p, want_slabs, want_slabs > 1 ? c->slab_w : nullptr);
        CU(launch_film_tables(to_dev(cam), c->rs, c->stream));
        CU(launch_view_bounds(c->scene, vp, c->view, to_dev(cam), p, slabs, c->stream, want_slabs > 1 ? c->ev_reset : nullptr, false));
        CU(launch_raster(c->scene, vp, to_dev(cam), p, out, c->scratch, c->view, c->rs, slabs, ctr, o->fp_mode == ATRAY_FP_FAST, c->cfg,
                         c->stream));
        if (slabs_out) *slabs_out = slabs;
    }
    else if (cull)
    {
        // the film rectangles depend on the camera: recomputed every frame, inside the timed region
        const DevSlabs slabs = make_slabs(to_dev(cam), p, want_slabs, want_slabs > 1 ? c->slab_w : nullptr);
        DevView vw = c->view;
        atray_ctx::Share& sh = c->share;
        const uint32_t n_blocks = (c->scene.n_slots + 255u) / 256u;
        const size_t tiles = size_t(vp.tiles_x) * size_t(vp.tiles_y);
        const bool shared = sh.on && o->part.world == sh.world && o->part.rank == sh.rank && sh.world > 1 && tiles <= sh.tiles_cap &&
                            sh.n_slots == c->scene.n_slots;
        if (shared)
        {
            // every rank computes blocks [lo, hi) of the pass and stores them into all ranks' tables of this frame's parity;
            // then a flag per rank, and this rank's stream waits for the flags of the others before k_view_nodes reads the tables
            const uint32_t seq = ++sh.seq;
            const size_t bo = size_t(seq & 1u) * sh.buf_stride;
            vp.frame_id = seq;
            vw.frame_id = seq;
            ViewPeers pe{};
            pe.n = sh.world;
            pe.self = sh.rank;
            if (const char* env = std::getenv("ATRAY_SHARE_SKIP")) pe.debug_skip = uint32_t(std::atoi(env));
            pe.ranges_only = share_ranges_only();
            pe.block_lo = uint32_t(uint64_t(n_blocks) * uint64_t(sh.rank) / uint64_t(sh.world));
            pe.block_hi = uint32_t(uint64_t(n_blocks) * uint64_t(sh.rank + 1) / uint64_t(sh.world));
            for (int q = 0; q < sh.world; ++q)
            {
                char* b = sh.peer[q] + bo;
                pe.tri_px[q] = reinterpret_cast<uint2*>(b + sh.off_tri); pe.pkt_px[q] = reinterpret_cast<uint2*>(b + sh.off_pkt);
                pe.chunk_px[q] = reinterpret_cast<uint2*>(b + sh.off_chunk); pe.stamps[q] = reinterpret_cast<uint32_t*>(b + sh.off_stamps);
                pe.tile_cost[q] = reinterpret_cast<uint32_t*>(b + sh.off_cost); pe.all_flag[q] = reinterpret_cast<uint32_t*>(b + sh.off_all);
            }
            vw.tri_px = pe.tri_px[sh.rank]; vw.pkt_px = pe.pkt_px[sh.rank]; vw.chunk_px = pe.chunk_px[sh.rank];
            vw.stamps = pe.stamps[sh.rank]; vw.tile_cost = pe.tile_cost[sh.rank]; vw.all_flag = pe.all_flag[sh.rank];
            vw.tri_rect = nullptr;
            uint32_t* row = reinterpret_cast<uint32_t*>(sh.block + sh.off_flags);
            for (int q = 0; q < sh.world; ++q) pe.flags[q] = reinterpret_cast<uint32_t*>(sh.peer[q] + sh.off_flags);
            pe.seq = seq;
            pe.ticket = row + 48;
            ViewWait wt{};
            wt.my_flags = row; wt.n = sh.world; wt.rank = sh.rank; wt.seq = seq; wt.timeout_ns = 2000000000ull; wt.timed_out = row + 32;
            CU(launch_view_bounds(c->scene, vp, vw, to_dev(cam), p, slabs, c->stream, want_slabs > 1 ? c->ev_reset : nullptr, true, &pe, &wt, c->ev_vb));
        }
        else if (sh.debug_shards > 1)
        {
            // verification aid: the same pass as `debug_shards` shard launches on this GPU alone (every "peer" is this GPU)
            ViewPeers pe{};
            pe.n = sh.debug_shards > o->part.world ? sh.debug_shards : o->part.world;
            if (pe.n > kMaxPeers) pe.n = kMaxPeers;
            pe.ranges_only = share_ranges_only();
            if (const char* env = std::getenv("ATRAY_SHARE_SKIP")) pe.debug_skip = uint32_t(std::atoi(env));
            for (int q = 0; q < pe.n; ++q)
            {
                pe.tri_px[q] = vw.tri_px; pe.pkt_px[q] = vw.pkt_px; pe.chunk_px[q] = vw.chunk_px; pe.stamps[q] = vw.stamps;
                pe.tile_cost[q] = vw.tile_cost; pe.all_flag[q] = vw.all_flag;
            }
            for (int sdx = 0; sdx < sh.debug_shards; ++sdx)
            {
                pe.block_lo = uint32_t(uint64_t(n_blocks) * uint64_t(sdx) / uint64_t(sh.debug_shards));
                pe.block_hi = uint32_t(uint64_t(n_blocks) * uint64_t(sdx + 1) / uint64_t(sh.debug_shards));
                const bool last = sdx == sh.debug_shards - 1;
                CU(launch_view_bounds(c->scene, vp, vw, to_dev(cam), p, slabs, c->stream, last && want_slabs > 1 ? c->ev_reset : nullptr, last, &pe,
                                      nullptr, last ? c->ev_vb : nullptr));
            }
        }
        else
            CU(launch_view_bounds(c->scene, vp, vw, to_dev(cam), p, slabs, c->stream, want_slabs > 1 ? c->ev_reset : nullptr, true, nullptr,
                                  nullptr, c->ev_vb));
        CU(cudaEventRecord(c->ev_kr, c->stream));
        CU(launch_render_culled(c->scene, to_dev(cam), p, out, c->scratch, vw, slabs, ctr, o->fp_mode == ATRAY_FP_FAST,
                                o->jitter_seed, c->cfg, c->stream));
        if (slabs_out) *slabs_out = slabs;
    }
    else
        CU(launch_render(c->scene, to_dev(cam), p, out, c->scratch, c->d_tile_counter, ctr, o->fp_mode == ATRAY_FP_FAST,
                         o->jitter_seed, c->cfg, c->stream));
    CU(cudaEventRecord(c->ev_k1, c->stream));
    c->have_kernel_time = true;
    c->have_split = cull && !raster;
    return ATRAY_OK;
}

int atray_render_device(atray_ctx* c, const AtrayCamera* cam, const AtrayRenderOpts* o, const AtrayFrame* dev_out)
{
    int rc = check_render_args(c, cam, o, dev_out);
    if (rc) return rc;
    CU(cudaSetDevice(c->device));
    DevOut out{dev_out->bgrx, dev_out->t, dev_out->prim};
    return enqueue_render(c, cam, o, out);
}

static bool host_ptr_is_pinned(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// Device→host copy of the scanlines of this rank's local bands [lb0, lb1) (local band lb = global band lb*world + rank,
// band_rows scanlines each): the whole bands are one strided 2-D copy (pitch = world bands; a plain copy when world = 1),
// a band clipped by the image's top edge is one more plain copy.
static cudaError_t copy_local_bands(void* dst, const void* src, int W, int H, const AtrayPartition& part, int lb0, int lb1,
                                    cudaMemcpyKind kind, cudaStream_t stream)
{
    const size_t band_bytes = size_t(part.band_rows) * size_t(W) * 4;
    const int n_bands = (H + part.band_rows - 1) / part.band_rows;
    const int mine = part.rank < n_bands ? (n_bands - part.rank + part.world - 1) / part.world : 0;
    if (lb1 > mine) lb1 = mine;
    if (lb0 >= lb1) return cudaSuccess;
    const int last_band = part.rank + (lb1 - 1) * part.world;
    const bool last_clipped = (last_band + 1) * part.band_rows > H;
    const int whole = (last_clipped ? lb1 - 1 : lb1) - lb0;
    const size_t off = size_t(part.rank + lb0 * part.world) * band_bytes;
    cudaError_t e = cudaSuccess;
    if (whole > 0 && part.world == 1)
        e = cudaMemcpyAsync(static_cast<char*>(dst) + off, static_cast<const char*>(src) + off, size_t(whole) * band_bytes, kind, stream);
    else if (whole > 0)
        e = cudaMemcpy2DAsync(static_cast<char*>(dst) + off, size_t(part.world) * band_bytes, static_cast<const char*>(src) + off,
                              size_t(part.world) * band_bytes, band_bytes, size_t(whole), kind, stream);
    if (e == cudaSuccess && last_clipped)
    {
        const size_t o2 = size_t(last_band) * band_bytes, rows = size_t(H - last_band * part.band_rows);
        e = cudaMemcpyAsync(static_cast<char*>(dst) + o2, static_cast<const char*>(src) + o2, rows * size_t(W) * 4, kind, stream);
    }
    return e;
}
static cudaError_t copy_rank_bands(void* dst, const void* src, int W, int H, const AtrayPartition& part, cudaMemcpyKind kind,
                                   cudaStream_t stream)
{
    return copy_local_bands(dst, src, W, H, part, 0, 0x7fffffff, kind, stream);
}

int atray_render_begin(atray_ctx* c, const AtrayCamera* cam, const AtrayRenderOpts* o, const AtrayFrame* host)
{
    int rc = check_render_args(c, cam, o, host);
    if (rc) return rc;
    if (c->in_flight) return fail(ATRAY_ESTATE, "a render is already in flight");
    CU(cudaSetDevice(c->device));
    const size_t px = size_t(cam->width) * size_t(cam->height);
    if (px > c->frame_cap_px)
    {
        cudaFree(c->d_bgrx); cudaFree(c->d_t); cudaFree(c->d_prim);
        c->d_bgrx = nullptr; c->d_t = nullptr; c->d_prim = nullptr; c->frame_cap_px = 0;
        CU(cudaMalloc(&c->d_bgrx, px * 4)); CU(cudaMalloc(&c->d_t, px * 4)); CU(cudaMalloc(&c->d_prim, px * 4));
        c->frame_cap_px = px;
    }
    c->stage_bgrx = host->bgrx && !host_ptr_is_pinned(host->bgrx);
    c->stage_t = host->t && !host_ptr_is_pinned(host->t);
    c->stage_prim = host->prim && !host_ptr_is_pinned(host->prim);
    if ((c->stage_bgrx || c->stage_t || c->stage_prim) && px > c->stage_cap_px)
    {
        cudaFreeHost(c->h_bgrx); cudaFreeHost(c->h_t); cudaFreeHost(c->h_prim);
        c->h_bgrx = nullptr; c->h_t = nullptr; c->h_prim = nullptr; c->stage_cap_px = 0;
        CU(cudaMallocHost(&c->h_bgrx, px * 4)); CU(cudaMallocHost(&c->h_t, px * 4)); CU(cudaMallocHost(&c->h_prim, px * 4));
        c->stage_cap_px = px;
    }
    DevOut out{host->bgrx ? c->d_bgrx : nullptr, host->t ? c->d_t : nullptr, host->prim ? c->d_prim : nullptr};
    // Whole image on this context and a frame large enough to matter: render slab by slab and copy each slab's scanlines
    // out while the kernel is still tracing the later ones (the copy stream waits on the kernel's per-slab counters).
    const bool pipelined = c->wait_value32 && c->slab_copy > 1 && px >= (size_t(1) << 20);
    DevSlabs slabs{};
    rc = enqueue_render(c, cam, o, out, pipelined ? c->slab_copy : 1, &slabs);
    if (rc) return rc;
    uint32_t* hb = c->stage_bgrx ? c->h_bgrx : host->bgrx;
    float* ht = c->stage_t ? c->h_t : host->t;
    uint32_t* hp = c->stage_prim ? c->h_prim : host->prim;
    c->slabs_in_flight = 0;
    if (pipelined && slabs.n > 1)
    {
        CU(launch_slabs_final(c->view, slabs, c->stream));   // safety net: after the kernel every slab is final
        CU(cudaStreamWaitEvent(c->copy_stream, c->ev_reset, 0));
        if (c->copy_streams == 2) CU(cudaStreamWaitEvent(c->copy_stream2, c->ev_reset, 0));
        const uint32_t tiles_per_band = uint32_t(((cam->width + 7) >> 3) * (o->part.band_rows >> 2));
        for (int s = 0; s < slabs.n; ++s)
        {
            // local bands of the slab; with a partition they are every world-th band of the image (strided copy)
            const int lb0 = int(slabs.tile_begin[s] / tiles_per_band), lb1 = int(slabs.tile_begin[s + 1] / tiles_per_band);
            int y1 = ((lb1 - 1) * o->part.world + o->part.rank + 1) * o->part.band_rows;
            if (y1 > cam->height) y1 = cam->height;
            cudaStream_t cs = (c->copy_streams == 2 && (s & 1)) ? c->copy_stream2 : c->copy_stream;
            const unsigned long long addr = (unsigned long long)(uintptr_t)(c->view.busy_count + kCtlDone + s);
            if (c->wait_value32(cs, addr, slabs.live_tiles[s], 0u /* CU_STREAM_WAIT_VALUE_GEQ */) != 0)
                return fail(ATRAY_ECUDA, "cuStreamWaitValue32 failed");
            if (hb) CU(copy_local_bands(hb, c->d_bgrx, cam->width, cam->height, o->part, lb0, lb1, cudaMemcpyDeviceToHost, cs));
            if (ht) CU(copy_local_bands(ht, c->d_t, cam->width, cam->height, o->part, lb0, lb1, cudaMemcpyDeviceToHost, cs));
            if (hp) CU(copy_local_bands(hp, c->d_prim, cam->width, cam->height, o->part, lb0, lb1, cudaMemcpyDeviceToHost, cs));
            CU(cudaEventRecord(c->ev_slab[s], cs));
            c->slab_row_end[s] = int32_t(y1);   // this rank's scanlines below y1 have landed
        }
        if (c->copy_streams == 2 && slabs.n > 1)   // the last odd slab's event orders stream 2 before the frame's end
            CU(cudaStreamWaitEvent(c->copy_stream, c->ev_slab[((slabs.n - 1) & 1) ? slabs.n - 1 : slabs.n - 2], 0));
        CU(cudaStreamWaitEvent(c->copy_stream, c->ev_k1, 0));   // the frame is done only when the kernel has retired too
        CU(cudaEventRecord(c->ev_done, c->copy_stream));
        CU(cudaStreamWaitEvent(c->stream, c->ev_done, 0));      // later work on the context stream must not overtake the copies
        c->slabs_in_flight = slabs.n;
    }
    else if (o->part.world > 1)
    {
        // A partition writes exactly this rank's scanlines of the host frame, as a worker of the reference's tile pool
        // writes only its tiles (renderer.cpp:294-369): every rank of a multi-GPU job can be handed the SAME host bitmap
        // (shared pinned memory) and the copies of all ranks run in parallel, each over its own GPU's PCIe link.
        if (hb) CU(copy_rank_bands(hb, c->d_bgrx, cam->width, cam->height, o->part, cudaMemcpyDeviceToHost, c->stream));
        if (ht) CU(copy_rank_bands(ht, c->d_t, cam->width, cam->height, o->part, cudaMemcpyDeviceToHost, c->stream));
        if (hp) CU(copy_rank_bands(hp, c->d_prim, cam->width, cam->height, o->part, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaEventRecord(c->ev_done, c->stream));
    }
    else
    {
        if (hb) CU(cudaMemcpyAsync(hb, c->d_bgrx, px * 4, cudaMemcpyDeviceToHost, c->stream));
        if (ht) CU(cudaMemcpyAsync(ht, c->d_t, px * 4, cudaMemcpyDeviceToHost, c->stream));
        if (hp) CU(cudaMemcpyAsync(hp, c->d_prim, px * 4, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaEventRecord(c->ev_done, c->stream));
    }
    c->pending_user = *host;
    c->pending_part = o->part;
    c->pending_w = cam->width;
    c->pending_px = px;
    c->pending_rows = cam->height;
    c->in_flight = true;
    return ATRAY_OK;
}

int atray_render_poll(atray_ctx* c, uint32_t timeout_ms)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    if (!c->in_flight) return 0;
    CU(cudaSetDevice(c->device));
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
    for (;;)
    {
        cudaError_t e = cudaEventQuery(c->ev_done);
        if (e == cudaSuccess) break;
        if (e != cudaErrorNotReady) { c->in_flight = false; return cuda_fail(e, "render"); }
        if (timeout_ms != 0xffffffffu && std::chrono::steady_clock::now() >= deadline) return 1;  // still running
        std::this_thread::yield();
    }
    // staged planes (caller passed pageable memory): whole frame, or this rank's scanlines only when partitioned
    auto unstage = [&](uint32_t* dst, const uint32_t* src) {
        const AtrayPartition& pt = c->pending_part;
        if (pt.world <= 1) { std::memcpy(dst, src, c->pending_px * 4); return; }
        const size_t row = size_t(c->pending_w);
        for (int y0 = pt.rank * pt.band_rows; y0 < c->pending_rows; y0 += pt.world * pt.band_rows)
        {
            const int y1 = y0 + pt.band_rows < c->pending_rows ? y0 + pt.band_rows : c->pending_rows;
            std::memcpy(dst + size_t(y0) * row, src + size_t(y0) * row, size_t(y1 - y0) * row * 4);
        }
    };
    if (c->stage_bgrx) unstage(c->pending_user.bgrx, c->h_bgrx);
    if (c->stage_t) unstage(reinterpret_cast<uint32_t*>(c->pending_user.t), reinterpret_cast<const uint32_t*>(c->h_t));
    if (c->stage_prim) unstage(c->pending_user.prim, c->h_prim);
    c->in_flight = false;
    return 0;
}

int atray_render_progress(atray_ctx* c, int32_t* rows_done, int32_t* rows_total)
{
    if (!c || !rows_done || !rows_total) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaSetDevice(c->device));
    *rows_total = c->pending_rows;
    if (!c->in_flight) { *rows_done = c->pending_rows; return ATRAY_OK; }
    int32_t rows = 0;
    for (int s = 0; s < c->slabs_in_flight; ++s)   // slabs land in order (one copy stream)
    {
        cudaError_t e = cudaEventQuery(c->ev_slab[s]);
        if (e == cudaErrorNotReady) break;
        if (e != cudaSuccess) return cuda_fail(e, "render");
        rows = c->slab_row_end[s];
    }
    *rows_done = rows;
    return ATRAY_OK;
}

int atray_render_cancel(atray_ctx* c)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    // WorkQueue::clear() + wait_for_pool of the reference's cancel path (app.cpp:165-172: the workers finish the tile they are on
    // and find the queue empty). Here: raise the context's cancel word from a stream of its own, so the 4-byte copy overtakes the
    // running kernel; every persistent loop (k_render*, k_path, k_trace, k_c5) reads the word next to its work cursor and stops
    // taking tiles / ray groups; then drain the streams (the slab counters the copy streams wait on are finalised by the kernel
    // behind the render kernel, so they drain too), drop the pending host copy-out and lower the word again. Pixels / rays not yet
    // traced keep whatever the output buffers held. Applies to whatever work the context stream holds, begin/poll frames and the
    // asynchronous *_device entries alike.
    CU(cudaSetDevice(c->device));
    const bool busy = cudaStreamQuery(c->stream) == cudaErrorNotReady;
    cudaGetLastError();
    if (busy)
    {
        CU(cudaMemcpyAsync(c->d_cancel, c->h_cancel_one, sizeof(uint32_t), cudaMemcpyHostToDevice, c->ctl_stream));
        CU(cudaStreamSynchronize(c->ctl_stream));
        c->cancels += 1;
    }
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaStreamSynchronize(c->copy_stream));
    CU(cudaStreamSynchronize(c->copy_stream2));
    if (busy)
    {
        CU(cudaMemsetAsync(c->d_cancel, 0, sizeof(uint32_t), c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    c->in_flight = false;
    return busy ? 1 : ATRAY_OK;
}

int atray_ctx_sync(atray_ctx* c)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    return ATRAY_OK;
}

int atray_last_kernel_ms(atray_ctx* c, float* ms)
{
    if (!c || !ms) return fail(ATRAY_EINVAL, "null argument");
    if (!c->have_kernel_time) return fail(ATRAY_ESTATE, "no kernel has been launched");
    CU(cudaSetDevice(c->device));
    CU(cudaEventSynchronize(c->ev_k1));
    CU(cudaEventElapsedTime(ms, c->ev_k0, c->ev_k1));
    return ATRAY_OK;
}

int atray_last_render_kernel_ms(atray_ctx* c, float* ms)
{
    if (!c || !ms) return fail(ATRAY_EINVAL, "null argument");
    if (!c->have_kernel_time || !c->have_split) return fail(ATRAY_ESTATE, "the last launch was not a culled frame");
    CU(cudaSetDevice(c->device));
    CU(cudaEventSynchronize(c->ev_k1));
    CU(cudaEventElapsedTime(ms, c->ev_kr, c->ev_k1));
    return ATRAY_OK;
}

int atray_last_view_split_ms(atray_ctx* c, float* bounds_ms, float* nodes_ms)
{
    if (!c || !bounds_ms || !nodes_ms) return fail(ATRAY_EINVAL, "null argument");
    if (!c->have_kernel_time || !c->have_split) return fail(ATRAY_ESTATE, "the last launch was not a culled frame");
    CU(cudaSetDevice(c->device));
    CU(cudaEventSynchronize(c->ev_k1));
    CU(cudaEventElapsedTime(bounds_ms, c->ev_k0, c->ev_vb));
    CU(cudaEventElapsedTime(nodes_ms, c->ev_vb, c->ev_kr));
    return ATRAY_OK;
}

// ---- ray lists ------------------------------------------------------------------------------------------------------
int atray_trace_rays_device(atray_ctx* c, const float* rays_dev, uint64_t n, int32_t fp_mode, int32_t count_work,
                            uint32_t* prim_dev, float* t_dev)
{
    if (!c || (!rays_dev && n) || (!prim_dev && n) || (!t_dev && n)) return fail(ATRAY_EINVAL, "null argument");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "trace before scene_upload");
    if (n == 0) return ATRAY_OK;
    if (n > (uint64_t(1) << 36)) return fail(ATRAY_EINVAL, "ray list too long for one call (split it)");
    CU(cudaSetDevice(c->device));
    DevCounters* ctr = nullptr;
    if (count_work)
    {
        CU(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
        ctr = c->d_counters;
        c->counters_valid = true;
    }
    CU(cudaEventRecord(c->ev_k0, c->stream));
    CU(launch_trace(c->scene, rays_dev, n, prim_dev, t_dev, c->scratch, c->d_tile_counter, ctr, fp_mode == ATRAY_FP_FAST, c->cfg,
                    c->stream));
    CU(cudaEventRecord(c->ev_k1, c->stream));
    c->have_kernel_time = true;
    c->have_split = false;
    return ATRAY_OK;
}

// Ray list with TriangleIntersectionData (face_index, u, v) per ray: what get_ray_kd_tree_intersection returns for ONE ray
// (kd_tree.cpp:302-313; u, v of the winning triangle from its scan loop, kd_tree.cpp:436-462). Synchronous, host pointers; a miss
// leaves uv[2i], uv[2i+1] untouched and reports t = MAX_FLOAT. n = 1 is the reference's call shape (a 32-lane kernel for one
// ray is a latency-bound round trip of two small copies: the entry exists for drop-in completeness and for batches).
int atray_trace_rays_uv(atray_ctx* c, const float* rays, uint64_t n, uint32_t* prim, float* t, float* uv)
{
    if (!c || (!rays && n) || (!prim && n) || (!t && n) || (!uv && n)) return fail(ATRAY_EINVAL, "null argument");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "trace before scene_upload");
    if (n == 0) return ATRAY_OK;
    CU(cudaSetDevice(c->device));
    const uint64_t need = n * 2;   // d_rays doubles as {rays | uv}: 6 + 2 floats per ray → capacity in units of 24-byte rays
    if (need > c->ray_cap)
    {
        cudaFree(c->d_rays); cudaFree(c->d_rprim); cudaFree(c->d_rt);
        c->d_rays = nullptr; c->d_rprim = nullptr; c->d_rt = nullptr; c->ray_cap = 0;
        CU(cudaMalloc(&c->d_rays, need * 24)); CU(cudaMalloc(&c->d_rprim, need * 4)); CU(cudaMalloc(&c->d_rt, need * 4));
        c->ray_cap = need;
    }
    float* d_uv = c->d_rays + n * 6;
    CU(cudaMemcpyAsync(c->d_rays, rays, n * 24, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_uv, uv, n * 8, cudaMemcpyHostToDevice, c->stream));   // entries of misses keep the caller's values
    CU(cudaEventRecord(c->ev_k0, c->stream));
    CU(launch_trace_uv(c->scene, c->d_rays, n, c->d_rprim, c->d_rt, d_uv, c->scratch, c->d_tile_counter, c->cfg, c->stream));
    CU(cudaEventRecord(c->ev_k1, c->stream));
    c->have_kernel_time = true;
    c->have_split = false;
    CU(cudaMemcpyAsync(prim, c->d_rprim, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(t, c->d_rt, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(uv, d_uv, n * 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return ATRAY_OK;
}

// Same query as atray_trace_rays_device, on the library's static device scratch: this entry allocates nothing (no cudaMalloc /
// cudaFree / allocating routine on any path), see launch_trace_static in render.cu.
int atray_trace_rays_static(atray_ctx* c, const float* rays_dev, uint64_t n, int32_t fp_mode, uint32_t* prim_dev, float* t_dev)
{
    if (!c || (!rays_dev && n) || (!prim_dev && n) || (!t_dev && n)) return fail(ATRAY_EINVAL, "null argument");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "trace before scene_upload");
    if (n == 0) return ATRAY_OK;
    if (n > (uint64_t(1) << 36)) return fail(ATRAY_EINVAL, "ray list too long for one call (split it)");
    static_assert(ATRAY_STATIC_SCRATCH_BYTES == kStaticScratchBytes, "header and kernel disagree on the static scratch size");
    if (static_trace_blocks(c->scratch.leaf_cap, c->scratch.stack_cap, c->cfg) < 1)
    {
        char msg[256];
        std::snprintf(msg, sizeof msg,
                      "atray_trace_rays_static: the scene needs %zu bytes of traversal scratch per thread block (%u leaves, %u nodes), the "
                      "static scratch holds %zu: use atray_trace_rays_device",
                      static_trace_bytes_per_block(c->scratch.leaf_cap, c->scratch.stack_cap), c->scratch.leaf_cap - 2,
                      c->scratch.stack_cap - 8, static_scratch_capacity());
        return fail(ATRAY_ELIMIT, msg);
    }
    CU(cudaSetDevice(c->device));
    CU(cudaEventRecord(c->ev_k0, c->stream));
    CU(launch_trace_static(c->scene, rays_dev, n, prim_dev, t_dev, c->scratch.leaf_cap, c->scratch.stack_cap, fp_mode == ATRAY_FP_FAST,
                           c->cfg, c->device, c->stream));
    CU(cudaEventRecord(c->ev_k1, c->stream));
    c->have_kernel_time = true;
    c->have_split = false;
    return ATRAY_OK;
}

int atray_trace_rays(atray_ctx* c, const float* rays, uint64_t n, int32_t fp_mode, int32_t count_work, uint32_t* prim, float* t)
{
    if (!c || (!rays && n) || (!prim && n) || (!t && n)) return fail(ATRAY_EINVAL, "null argument");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "trace before scene_upload");
    if (n == 0) return ATRAY_OK;
    CU(cudaSetDevice(c->device));
    if (n > c->ray_cap)
    {
        cudaFree(c->d_rays); cudaFree(c->d_rprim); cudaFree(c->d_rt);
        c->d_rays = nullptr; c->d_rprim = nullptr; c->d_rt = nullptr; c->ray_cap = 0;
        CU(cudaMalloc(&c->d_rays, n * 24)); CU(cudaMalloc(&c->d_rprim, n * 4)); CU(cudaMalloc(&c->d_rt, n * 4));
        c->ray_cap = n;
    }
    CU(cudaMemcpyAsync(c->d_rays, rays, n * 24, cudaMemcpyHostToDevice, c->stream));
    int rc = atray_trace_rays_device(c, c->d_rays, n, fp_mode, count_work, c->d_rprim, c->d_rt);
    if (rc) return rc;
    CU(cudaMemcpyAsync(prim, c->d_rprim, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(t, c->d_rt, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return ATRAY_OK;
}

// ---- C5: incoherent rays generated on the device (BASELINE.json configs[4]) --------------------------------------------------
int atray_trace_incoherent_device(atray_ctx* c, uint64_t p0, uint64_t p1, uint32_t spp, uint64_t seed, const AtrayPartition* part,
                                  uint32_t chunk_px, int32_t fp_mode, int32_t count_work, uint32_t* bgrx_dev, uint32_t* prim_dev, float* t_dev)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "trace before scene_upload");
    if (p1 < p0 || spp == 0 || chunk_px == 0) return fail(ATRAY_EINVAL, "bad pixel range, sample count or chunk size");
    if ((p1 - p0) / chunk_px + 2 > (uint64_t(1) << 31) / chunk_px) return fail(ATRAY_EINVAL, "pixel range too long for one call (split it)");
    AtrayPartition pt{0, 1, 8};
    if (part) pt = *part;
    if (pt.world < 1 || pt.rank < 0 || pt.rank >= pt.world) return fail(ATRAY_EINVAL, "bad partition");
    if (p1 == p0) return ATRAY_OK;
    CU(cudaSetDevice(c->device));
    DevCounters* ctr = nullptr;
    if (count_work)
    {
        CU(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
        ctr = c->d_counters;
        c->counters_valid = true;
    }
    CU(cudaEventRecord(c->ev_k0, c->stream));
    CU(launch_c5(c->scene, p0, p1, spp, chunk_px, seed, DevPartition{pt.rank, pt.world, pt.band_rows}, bgrx_dev, prim_dev, t_dev, c->scratch,
                 c->d_tile_counter, ctr, fp_mode == ATRAY_FP_FAST, c->cfg, c->stream));
    CU(cudaEventRecord(c->ev_k1, c->stream));
    c->have_kernel_time = true;
    c->have_split = false;
    return ATRAY_OK;
}

int atray_trace_incoherent(atray_ctx* c, uint64_t p0, uint64_t p1, uint32_t spp, uint64_t seed, int32_t fp_mode, int32_t count_work,
                           uint32_t* bgrx, uint32_t* prim, float* t)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    if (p1 < p0) return fail(ATRAY_EINVAL, "bad pixel range");
    if (p1 == p0) return ATRAY_OK;
    CU(cudaSetDevice(c->device));
    const uint64_t npx = p1 - p0, nrays = npx * spp;
    uint32_t *d_b = nullptr, *d_p = nullptr;
    float* d_t = nullptr;
    if (bgrx) CU(cudaMalloc(&d_b, npx * 4));
    if (prim) CU(cudaMalloc(&d_p, nrays * 4));
    if (t) CU(cudaMalloc(&d_t, nrays * 4));
    int rc = atray_trace_incoherent_device(c, p0, p1, spp, seed, nullptr, 4096, fp_mode, count_work, d_b, d_p, d_t);
    if (rc == ATRAY_OK)
    {
        cudaError_t e = cudaSuccess;
        if (bgrx) e = cudaMemcpyAsync(bgrx, d_b, npx * 4, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess && prim) e = cudaMemcpyAsync(prim, d_p, nrays * 4, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess && t) e = cudaMemcpyAsync(t, d_t, nrays * 4, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "incoherent trace");
    }
    cudaFree(d_b); cudaFree(d_p); cudaFree(d_t);
    return rc;
}

int atray_get_counters(atray_ctx* c, AtrayCounters* out)
{
    if (!c || !out) return fail(ATRAY_EINVAL, "null argument");
    if (!c->counters_valid) return fail(ATRAY_ESTATE, "no counted frame yet (set count_work)");
    static_assert(sizeof(AtrayCounters) == sizeof(DevCounters), "counter layouts differ");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaMemcpy(out, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost));
    return ATRAY_OK;
}

// ---- verification aid: the film rectangles of view_bounds.h evaluated on the host ------------------------------------
int atray_debug_film_rects(const atray_model* m, const AtrayCamera* cam, float* rects, uint32_t* slot_leaf, uint32_t* slot_pos,
                           uint8_t* tile_busy, float* hulls, int32_t* px_ranges)
{
    if (!m || !cam) return fail(ATRAY_EINVAL, "null argument");
    if (m->m.tree.empty()) return fail(ATRAY_ESTATE, "film_rects before build_tree");
    ViewParams vp = make_view_params(cam);
    if (!vp.valid) return fail(ATRAY_ESTATE, "camera not usable for culling");
    FlatScene f;
    flatten(m->m, f);
    const size_t tiles = size_t(vp.tiles_x) * size_t(vp.tiles_y);
    if (tile_busy) std::memset(tile_busy, 0, tiles);
    for (uint32_t s = 0; s < f.n_slots; ++s)
    {
        const float* pkt = &f.tris[size_t(s / 8) * 24 * 4];
        const float* A = pkt + size_t(s % 8) * 4;
        const float* AB = A + 32;
        const float* AC = A + 64;
        uint32_t face, pos;
        std::memcpy(&face, A + 3, 4);
        std::memcpy(&pos, AB + 3, 4);
        FilmRect r = film_rect_empty();
        FilmHull hl = film_hull_all();
        if (face != 0xffffffffu) r = tri_film_rect(vp, A, AB, AC, &hl);
        if (hulls)
            for (int k = 0; k < 3; ++k) { hulls[9 * size_t(s) + 3 * k] = hl.a[k]; hulls[9 * size_t(s) + 3 * k + 1] = hl.b[k]; hulls[9 * size_t(s) + 3 * k + 2] = hl.c[k]; }
        if (rects) { rects[4 * size_t(s)] = r.xlo; rects[4 * size_t(s) + 1] = r.ylo; rects[4 * size_t(s) + 2] = r.xhi; rects[4 * size_t(s) + 3] = r.yhi; }
        if (slot_leaf) slot_leaf[s] = f.packet_leaf[s / 8];
        if (slot_pos) slot_pos[s] = face != 0xffffffffu ? pos : 0xffffffffu;
        PxRect pr;
        const bool any = film_rect_px(vp, r, pr);   // pr is {0, -1, 0, -1} when no pixel can hold such a ray
        if (px_ranges) { px_ranges[4 * size_t(s)] = pr.x0; px_ranges[4 * size_t(s) + 1] = pr.x1; px_ranges[4 * size_t(s) + 2] = pr.y0; px_ranges[4 * size_t(s) + 3] = pr.y1; }
        if (tile_busy && any)
        {
            const int tx0 = pr.x0 >> 3, tx1 = pr.x1 >> 3, ty0 = pr.y0 >> 2, ty1 = pr.y1 >> 2;
            if (2 * (tx1 - tx0 + 1) * (ty1 - ty0 + 1) >= int(tiles)) std::memset(tile_busy, 1, tiles);
            else
                for (int ty = ty0; ty <= ty1; ++ty)
                    for (int tx = tx0; tx <= tx1; ++tx) tile_busy[size_t(ty) * vp.tiles_x + tx] = 1;
        }
    }
    return ATRAY_OK;
}

// Verification aid (ATRAY_TILE_LOG=1 when the context renders its first culled frame): per local tile of the last culled
// frame {start ns (low word of the global timer), duration ns, sampled cost, cost class}; zeros for sky tiles.
int atray_debug_tile_log(atray_ctx* c, uint32_t* out, size_t n_tiles)
{
    if (!c || !out) return fail(ATRAY_EINVAL, "null argument");
    if (!c->view.tile_log) return fail(ATRAY_ESTATE, "tile log not enabled (ATRAY_TILE_LOG=1)");
    if (n_tiles > c->stamp_cap) return fail(ATRAY_EINVAL, "more tiles than the last frame had");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaMemcpy(out, c->view.tile_log, n_tiles * sizeof(uint4), cudaMemcpyDeviceToHost));
    return ATRAY_OK;
}

// ---- memory helpers ----------------------------------------------------------------------------------------------
int atray_dev_alloc(atray_ctx* c, size_t bytes, void** out)
{
    if (!c || !out) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaSetDevice(c->device));
    CU(cudaMalloc(out, bytes ? bytes : 1));
    return ATRAY_OK;
}
int atray_dev_free(atray_ctx* c, void* p)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    CU(cudaSetDevice(c->device));
    CU(cudaFree(p));
    return ATRAY_OK;
}
int atray_dev_copy_to_host(atray_ctx* c, void* host, const void* dev, size_t bytes)
{
    if (!c || !host || !dev) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return ATRAY_OK;
}
int atray_dev_copy_from_host(atray_ctx* c, void* dev, const void* host, size_t bytes)
{
    if (!c || !host || !dev) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return ATRAY_OK;
}
int atray_dev_memset(atray_ctx* c, void* dev, int value, size_t bytes)
{
    if (!c || !dev) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaSetDevice(c->device));
    CU(cudaMemsetAsync(dev, value, bytes, c->stream));
    return ATRAY_OK;
}
// ---- bitmap file (texture.cpp:66-115): the one artefact the reference application writes ------------------------------
int atray_write_bmp(const uint32_t* bgrx, uint32_t width, uint32_t height, const char* file_name, char* written_name, size_t cap)
{
    if (!bgrx || !file_name) return fail(ATRAY_EINVAL, "null argument");
    const size_t len = std::strlen(file_name);
    if (len + 8 > 1024) return fail(ATRAY_EINVAL, "file name too long (reference buffer: 1024)");
#pragma pack(push, 1)
    struct FileHeader { uint8_t type[2]; int32_t total_size; int16_t r1, r2; uint32_t offset_bits; };
#pragma pack(pop)
    struct DibHeader
    {
        uint32_t size_header, width, height; int16_t planes, bits_per_pixel;
        uint32_t compression, buffer_size, ppm_x, ppm_y, clr_used, clr_important, red, green, blue, alpha;
    };
    static_assert(sizeof(FileHeader) == 14 && sizeof(DibHeader) == 56, "BMP header layout");
    const uint32_t bytes = 4u * width * height;
    DibHeader dib{uint32_t(sizeof(DibHeader)), width, height, 1, 32, 3u, bytes, 197u, 39u, 0u, 0u, 0x00FF0000u, 0x0000FF00u, 0x000000FFu, 0u};
    FileHeader fh{{'B', 'M'}, int32_t(14u + sizeof(DibHeader) + bytes), 0, 0, uint32_t(14u + sizeof(DibHeader))};
    char name[1024];
    FILE* f = nullptr;
    for (uint32_t id = 0; !f; ++id)   // CREATE_NEW semantics: the first "<name>_<id>.bmp" that does not exist yet
    {
        std::snprintf(name, len + 8, "%s_%u.bmp", file_name, id);
        f = std::fopen(name, "wbx");
        if (!f && errno != EEXIST) return fail(ATRAY_EIO, std::string("cannot create ") + name);
    }
    const bool ok = std::fwrite(&fh, 1, 14, f) == 14 && std::fwrite(&dib, 1, sizeof dib, f) == sizeof dib &&
                    std::fwrite(bgrx, 1, bytes, f) == bytes;
    std::fclose(f);
    if (written_name && cap) std::snprintf(written_name, cap, "%s", name);
    return ok ? ATRAY_OK : fail(ATRAY_EIO, "short write");
}

int atray_host_alloc_pinned(size_t bytes, void** out)
{
    if (!out) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaMallocHost(out, bytes ? bytes : 1));
    return ATRAY_OK;
}
int atray_host_register(void* p, size_t bytes)
{
    if (!p || !bytes) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return ATRAY_OK;
}

int atray_host_unregister(void* p)
{
    if (!p) return ATRAY_OK;
    CU(cudaHostUnregister(p));
    return ATRAY_OK;
}

int atray_host_free_pinned(void* p)
{
    CU(cudaFreeHost(p));
    return ATRAY_OK;
}
int atray_ipc_export(atray_ctx* c, void* dev, unsigned char* handle64)
{
    if (!c || !dev || !handle64) return fail(ATRAY_EINVAL, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    CU(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, dev));
    std::memcpy(handle64, &h, 64);
    return ATRAY_OK;
}
int atray_ipc_open(atray_ctx* c, const unsigned char* handle64, void** out)
{
    if (!c || !handle64 || !out) return fail(ATRAY_EINVAL, "null argument");
    CU(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    CU(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return ATRAY_OK;
}
int atray_ipc_close(atray_ctx* c, void* p)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    CU(cudaSetDevice(c->device));
    CU(cudaIpcCloseMemHandle(p));
    return ATRAY_OK;
}

// ---- shared view pass (multi-GPU) ------------------------------------------------------------------------------------
int atray_view_share_create(atray_ctx* c, int32_t max_width, int32_t max_height, void** block, size_t* bytes)
{
    if (!c || !block || max_width <= 0 || max_height <= 0) return fail(ATRAY_EINVAL, "bad argument");
    if (!c->has_scene) return fail(ATRAY_ESTATE, "view_share_create before scene_upload");
    if (c->in_flight) return fail(ATRAY_ESTATE, "render in flight");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    atray_ctx::Share& sh = c->share;
    cudaFree(sh.block);
    sh = atray_ctx::Share{};
    auto up = [](size_t v) { return (v + 255) & ~size_t(255); };
    const size_t tiles = size_t((max_width + 7) >> 3) * size_t((max_height + 3) >> 2);
    const size_t ns = c->scene.n_slots;
    size_t at = 0;
    sh.off_tri = at; at += up((ns ? ns : 1) * sizeof(uint2));
    sh.off_pkt = at; at += up((ns / 8 + 1) * sizeof(uint2));
    sh.off_chunk = at; at += up((ns / 256 + 1) * sizeof(uint2));
    sh.off_stamps = at; at += up(tiles * sizeof(uint32_t));
    sh.off_cost = at; at += up(tiles * sizeof(uint32_t));
    sh.off_all = at; at += 256;
    sh.buf_stride = at;
    sh.off_flags = 2 * at;
    sh.bytes = 2 * at + 256;
    sh.tiles_cap = tiles;
    sh.n_slots = c->scene.n_slots;
    CU(cudaMalloc(&sh.block, sh.bytes));
    CU(cudaMemset(sh.block, 0, sh.bytes));
    *block = sh.block;
    if (bytes) *bytes = sh.bytes;
    return ATRAY_OK;
}

int atray_view_share_connect(atray_ctx* c, int32_t rank, int32_t world, void* const* peer_blocks)
{
    if (!c || !peer_blocks) return fail(ATRAY_EINVAL, "null argument");
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return fail(ATRAY_EINVAL, "bad rank / world (at most 8 ranks)");
    if (!c->share.block) return fail(ATRAY_ESTATE, "view_share_connect before view_share_create");
    if (c->in_flight) return fail(ATRAY_ESTATE, "render in flight");
    for (int q = 0; q < world; ++q)
    {
        if (!peer_blocks[q]) return fail(ATRAY_EINVAL, "null peer block");
        c->share.peer[q] = static_cast<char*>(peer_blocks[q]);
    }
    if (c->share.peer[rank] != c->share.block) return fail(ATRAY_EINVAL, "peer_blocks[rank] must be this context's own block");
    c->share.rank = rank;
    c->share.world = world;
    c->share.seq = 0;
    c->share.rdv = 0;
    c->share.on = world > 1;
    return ATRAY_OK;
}

int atray_view_share_rendezvous(atray_ctx* c)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    atray_ctx::Share& sh = c->share;
    if (!sh.on) return ATRAY_OK;
    CU(cudaSetDevice(c->device));
    // second flag row of the block: "rank r has reached rendezvous k" — same signal / wait pair as the view pass
    const uint32_t k = ++sh.rdv;
    uint32_t* flags[kMaxPeers];
    for (int q = 0; q < sh.world; ++q) flags[q] = reinterpret_cast<uint32_t*>(sh.peer[q] + sh.off_flags) + 16;
    CU(launch_view_signal(flags, sh.world, sh.rank, k, c->stream));
    uint32_t* row = reinterpret_cast<uint32_t*>(sh.block + sh.off_flags);
    CU(launch_view_wait(row + 16, sh.world, sh.rank, k, 2000000000ull, row + 33, c->stream));
    return ATRAY_OK;
}

int atray_view_share_check(atray_ctx* c)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    atray_ctx::Share& sh = c->share;
    if (!sh.block) return ATRAY_OK;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    uint32_t w[2] = {0u, 0u};
    CU(cudaMemcpy(w, reinterpret_cast<uint32_t*>(sh.block + sh.off_flags) + 32, sizeof w, cudaMemcpyDeviceToHost));
    if (w[0]) return fail(ATRAY_ECUDA, "shared view pass: a peer did not signal frame " + std::to_string(w[0]) + " within 2 s (the frame is not valid)");
    if (w[1]) return fail(ATRAY_ECUDA, "shared view pass: a peer did not reach rendezvous " + std::to_string(w[1]) + " within 2 s");
    return ATRAY_OK;
}

int atray_view_share_off(atray_ctx* c)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    if (c->in_flight) return fail(ATRAY_ESTATE, "render in flight");
    c->share.on = false;
    return ATRAY_OK;
}

int atray_ctx_enable_peer(atray_ctx* c, atray_ctx* other)
{
    if (!c || !other) return fail(ATRAY_EINVAL, "null context");
    if (c->device == other->device) return ATRAY_OK;
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, c->device, other->device));
    if (!can) return fail(ATRAY_ECUDA, "no peer access between the two devices");
    CU(cudaSetDevice(c->device));
    cudaError_t e = cudaDeviceEnablePeerAccess(other->device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
    cudaGetLastError();
    return ATRAY_OK;
}

int atray_debug_view_shards(atray_ctx* c, int32_t shards)
{
    if (!c || shards < 0 || shards > 64) return fail(ATRAY_EINVAL, "bad argument");
    c->share.debug_shards = shards;
    return ATRAY_OK;
}

// ---- K0 ------------------------------------------------------------------------------------------------------------
int atray_microbench(atray_ctx* c, double* fp32_lane_ops_per_s, double* l2_read_gbs)
{
    if (!c) return fail(ATRAY_EINVAL, "null context");
    CU(cudaSetDevice(c->device));
    float* sink = nullptr;
    CU(cudaMalloc(&sink, 16));
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
    float ms = 0.f;
    if (fp32_lane_ops_per_s)
    {
        LaunchCfg cfg{c->sm_count * 8, 256};
        const int iters = 4096;
        CU(launch_microbench_fp32(sink, 64, cfg, c->stream));  // warm-up
        double best = 0.0;
        for (int rep = 0; rep < 3; ++rep)
        {
            CU(cudaEventRecord(a, c->stream));
            CU(launch_microbench_fp32(sink, iters, cfg, c->stream));
            CU(cudaEventRecord(b, c->stream));
            CU(cudaEventSynchronize(b));
            CU(cudaEventElapsedTime(&ms, a, b));
            const double ops = double(cfg.grid) * cfg.block * double(iters) * 64.0;  // one FFMA = one issue slot per lane
            if (ops / (ms * 1e-3) > best) best = ops / (ms * 1e-3);
        }
        *fp32_lane_ops_per_s = best;
    }
    if (l2_read_gbs)
    {
        const size_t bytes = size_t(32) << 20;  // 32 MiB: L2-resident on B200 (126 MB), larger than any L1
        float4* buf = nullptr;
        CU(cudaMalloc(&buf, bytes));
        CU(cudaMemsetAsync(buf, 0, bytes, c->stream));
        LaunchCfg cfg{c->sm_count * 8, 256};
        const int passes = 16;
        CU(launch_microbench_l2(buf, bytes / 16, 2, sink, cfg, c->stream));
        double best = 0.0;
        for (int rep = 0; rep < 3; ++rep)
        {
            CU(cudaEventRecord(a, c->stream));
            CU(launch_microbench_l2(buf, bytes / 16, passes, sink, cfg, c->stream));
            CU(cudaEventRecord(b, c->stream));
            CU(cudaEventSynchronize(b));
            CU(cudaEventElapsedTime(&ms, a, b));
            const double gbs = double(bytes) * passes / (ms * 1e-3) / 1e9;
            if (gbs > best) best = gbs;
        }
        *l2_read_gbs = best;
        CU(cudaFree(buf));
    }
    CU(cudaEventDestroy(a)); CU(cudaEventDestroy(b));
    CU(cudaFree(sink));
    return ATRAY_OK;
}

}  // extern "C"
