// Render entry points with the reference's names and semantics (engine/renderer/renderer.h:11-35,
// renderer.cpp:264-291, 403-471) on top of the C ABI. The ThreadPool parameters stay in the signatures and are ignored:
// the tile thread pool is replaced by the sm_100a kernels of libatray_b200.so.
#pragma once
#include <cstdlib>
#include <cstring>
#include "PL/PL_math.h"
#include "engine/tools/texture.h"
#include "engine/tools/work_queue.h"
#include "scene.h"
#include "camera.h"

#ifndef ATRAY_NO_KD_TREE
#define USE_KD_TREE   // renderer.h:8; compile with -DATRAY_NO_KD_TREE for the reference's no-tree branch (renderer.cpp:58-82)
#endif

struct RenderTile { Tile tile; int64 ray_casts; };

struct RenderInfo
{
    WorkQueue<RenderTile> twq;
    uint32 hit_stack_capacity = 0;   // kept for source compatibility; traversal scratch is sized by the library
    uint32 leaf_stack_capacity = 0;
    Camera* camera = nullptr;
    Scene* scene = nullptr;
    Texture* camera_tex = nullptr;
    int64 total_ray_casts = 0;
    // additions (optional outputs north_star asks for; leave null for the reference behaviour)
    f32* hit_t = nullptr;            // W*H, ATRAY_MAX_FLOAT on miss
    uint32* hit_prim = nullptr;      // W*H, ATRAY_MISS on miss
    int32 fp_mode = ATRAY_FP_STRICT;
    uint64 jitter_seed = 0x853c49e6748fea9bULL;
    int last_error = ATRAY_OK;
};

// renderer.cpp:264-291: builds each model's tree if it has none and reports the largest node count; additionally
// flattens and uploads the scene once. One model per scene on this path (the reference's default scene, app.cpp:138).
inline void prep_scene(Scene& scene, uint32& kd_tree_max_nodes)
{
    kd_tree_max_nodes = 0;
    for (int32 i = 0; i < scene.models.length; i++)
    {
        Model& m = scene.models[i];
        if (m.kd_tree.tree.front == 0) build_KD_tree(m.data, m.kd_tree);
        if (m.kd_tree.tree.length > (int32)kd_tree_max_nodes) kd_tree_max_nodes = m.kd_tree.tree.length;
    }
    if (scene.models.length < 1 || scene.materials.length < 1) return;
    if (!scene.gpus)
    {
        const int32 one = scene.gpu_device;
        const int32 n = scene.gpu_devices.length > 0 ? scene.gpu_devices.length : 1;
        if (atray_ctx_create_multi(scene.gpu_devices.length > 0 ? scene.gpu_devices.front : &one, n, &scene.gpus) != ATRAY_OK) return;
        scene.gpu = atray_multi_ctx(scene.gpus, 0);
    }
    Model& m = scene.models[0];
    AtrayMaterial sky, mat;
    static_assert(sizeof(Material) == sizeof(AtrayMaterial), "Material layout");
    std::memcpy(&sky, &scene.materials[0], sizeof sky);
    std::memcpy(&mat, m.data.material ? m.data.material : &scene.materials[0], sizeof mat);
    if (atray_multi_scene_upload(scene.gpus, m.data.handle, &sky, &mat) != ATRAY_OK)
    {
        atray_multi_destroy(scene.gpus);   // callers test scene.gpu after prep_scene
        scene.gpus = nullptr;
        scene.gpu = nullptr;
        return;
    }
    for (int32 i = 0; i < scene.models.length; i++) scene.models[i].kd_tree.gpu = scene.gpu;   // get_ray_kd_tree_intersection
    // spheres and planes (scene.h:13-24; prep_scene normalises the plane normals, renderer.cpp:267-270 — done by the library)
    if (scene.spheres.length > 0 || scene.planes.length > 0)
    {
        AtraySphere* sp = (AtraySphere*)std::calloc(scene.spheres.length > 0 ? scene.spheres.length : 1, sizeof(AtraySphere));
        AtrayPlane* pl = (AtrayPlane*)std::calloc(scene.planes.length > 0 ? scene.planes.length : 1, sizeof(AtrayPlane));
        for (int32 i = 0; i < scene.spheres.length; i++)
        {
            std::memcpy(sp[i].center, &scene.spheres[i].center, 12);
            sp[i].radius = scene.spheres[i].radius;
            std::memcpy(&sp[i].material, scene.spheres[i].material ? scene.spheres[i].material : &scene.materials[0], sizeof(AtrayMaterial));
        }
        for (int32 i = 0; i < scene.planes.length; i++)
        {
            std::memcpy(pl[i].normal, &scene.planes[i].normal, 12);
            pl[i].distance = scene.planes[i].distance;
            std::memcpy(&pl[i].material, scene.planes[i].material ? scene.planes[i].material : &scene.materials[0], sizeof(AtrayMaterial));
        }
        atray_multi_scene_set_primitives(scene.gpus, sp, (uint32_t)scene.spheres.length, pl, (uint32_t)scene.planes.length);
        std::free(sp);
        std::free(pl);
    }
#if !defined(USE_KD_TREE)
    atray_multi_scene_use_kd_tree(scene.gpus, 0);   // get_intersection_data's #else branch: model box, then every face
#endif
}

// renderer.cpp:403-455: returns immediately. twq.jobs is filled with one band-shaped job per 8 scanlines (inclusive
// rectangles as the reference's tiles) so callers that draw "Tiles k/n" keep working; jobs_done follows the scanlines
// that have landed in camera_tex (large frames are copied out slab by slab while the kernel is still tracing).
inline void cancel_render_from_queue(void* scene_gpus) { atray_multi_render_cancel((atray_multi*)scene_gpus); }   // WorkQueue::clear()

inline void start_render_from_camera(RenderInfo& info, ThreadPool& tpool)
{
    (void)tpool;
    // a frame of this scene still in flight (the caller did not wait for it): the new one replaces it, as clearing the queue would
    if (info.scene->gpus && atray_multi_render_poll(info.scene->gpus, 0) == 1) atray_multi_render_cancel(info.scene->gpus);
    info.twq.on_clear = nullptr;
    const int32 W = info.camera_tex->bmb.width, H = info.camera_tex->bmb.height;
    const int32 band = 8, n = (H + band - 1) / band;
    if (info.twq.jobs.front) info.twq.jobs.clear();
    RenderTile* t = info.twq.jobs.allocate(n);
    for (int32 b = 0; b < n; ++b)
    {
        t[b].tile.left_bottom = {0, b * band};
        t[b].tile.right_top = {W - 1, (b + 1) * band - 1 < H - 1 ? (b + 1) * band - 1 : H - 1};
        t[b].ray_casts = 0;
    }
    info.twq.jobs_done = 0;
    AtrayCamera cam;
    camera_to_abi(*info.camera, cam);
    AtrayRenderOpts o{};
    o.fp_mode = info.fp_mode; o.count_work = 0; o.jitter_seed = info.jitter_seed;
    o.part.rank = 0; o.part.world = 1; o.part.band_rows = band;
    AtrayFrame f{(uint32_t*)info.camera_tex->bmb.buffer_memory, info.hit_t, info.hit_prim};
    info.last_error = info.scene->gpus ? atray_multi_render_begin(info.scene->gpus, &cam, &o, &f) : ATRAY_ESTATE;
    if (info.last_error == ATRAY_OK)
    {
        info.twq.on_clear = cancel_render_from_queue;   // info.twq.clear() reaches the kernels (app.cpp:165-172)
        info.twq.on_clear_arg = info.scene->gpus;
    }
}

// renderer.cpp:457-471: TRUE while still running (timed out), FALSE when finished; totals are accumulated on completion.
inline b32 wait_for_render_from_camera_to_finish(RenderInfo& info, ThreadPool& tpool, uint32 ms_to_wait_for)
{
    (void)tpool;
    if (info.last_error != ATRAY_OK || !info.scene->gpus) return FALSE;
    if (info.twq.jobs.size == 0 && info.twq.jobs.front == nullptr) return FALSE;   // the queue was cleared: nothing is running
    int r = atray_multi_render_poll(info.scene->gpus, ms_to_wait_for);
    if (r == 1)
    {
        int32_t rows = 0, total = 0;
        if (atray_multi_render_progress(info.scene->gpus, &rows, &total) == ATRAY_OK) info.twq.jobs_done = rows / 8;
        return TRUE;
    }
    info.twq.on_clear = nullptr;
    if (r < 0) { info.last_error = r; return FALSE; }
    info.twq.jobs_done = info.twq.jobs.size;
    if (info.camera->render_settings.bounce_limit != 1 || info.scene->spheres.length > 0 || info.scene->planes.length > 0)
    {
        // sum of cast_ray's bounce iterations (renderer.cpp:260, 467-470) over the contexts' scanlines
        for (int32 g = 0; g < atray_multi_count(info.scene->gpus); ++g)
        {
            uint64_t casts = 0;
            if (atray_last_ray_casts(atray_multi_ctx(info.scene->gpus, g), &casts) == ATRAY_OK) info.total_ray_casts += (int64)casts;
        }
    }
    else if (info.hit_prim && info.camera->render_settings.samples_per_pixel == 1)
    {
        // ray_casts = bounce iterations that hit geometry (renderer.cpp:222-231,260), per band
        const int32 W = info.camera_tex->bmb.width;
        for (int32 b = 0; b < info.twq.jobs.size; ++b)
        {
            int64 c = 0;
            for (int32 y = info.twq.jobs[b].tile.left_bottom.y; y <= info.twq.jobs[b].tile.right_top.y; ++y)
                for (int32 x = 0; x < W; ++x) c += info.hit_prim[(size_t)y * W + x] != ATRAY_MISS;
            info.twq.jobs[b].ray_casts = c;
            info.total_ray_casts += c;
        }
    }
    return FALSE;
}
