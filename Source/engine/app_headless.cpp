// Headless replay of the reference's render_app call sequence (engine/app.cpp:59-195) against the drop-in host layer:
// load → get_AABB → translate_to → set_camera → materials/scene → prep_scene → start_render_from_camera →
// poll wait_for_render_from_camera_to_finish(33 ms) [→ info.twq.clear() + wait_for_pool on --cancel-after-ms] → stats
// (+ optional BMP and raw plane dumps). What app.cpp hard-codes (app.cpp:16, 81-96, 122-126) is a flag here:
//   app_headless --obj model.obj [--res WxH] [--spp N] [--aa 0|1] [--bounces N] [--leaf-cap N] [--method sah|center]
//                [--gpus N | --devices a,b,..] [--mode tree|notree] [--fp strict|fast] [--rescale S] [--loader-threads N]
//                [--gather bgrx|all] [--frames N] [--cancel-after-ms T] [--probe-ray ox,oy,oz,dx,dy,dz] [--bmp base] [--planes out.bin]
//   (old positional form, still accepted: app_headless model.obj [W H] [rescale] [bmp_base] [planes.bin])
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "engine/renderer/renderer.h"
#include "engine/tools/OBJ_loader.h"

struct Args
{
    std::string obj, bmp, planes, devices;
    uint32 W = 1280, H = 720, spp = 1, leaf_cap = 300, frames = 1;
    int32 aa = 0, bounces = 1, gpus = 1, loader_threads = 8, cancel_after_ms = -1;
    bool sah = true, notree = false, fast = false, probe_ray = false;
    bool all_planes = true;   // --gather all: bitmap + hit distance + primitive id reach the host; bgrx: the bitmap only (the reference's output)
    f32 rescale = 0.f, probe[6] = {0, 0, 0, 0, 0, -1};
};

static int usage(const char* exe)
{
    std::fprintf(stderr,
                 "usage: %s --obj model.obj [--res WxH] [--spp N] [--aa 0|1] [--bounces N] [--leaf-cap N] [--method sah|center]\n"
                 "          [--gpus N | --devices a,b,..] [--mode tree|notree] [--fp strict|fast] [--rescale S] [--loader-threads N]\n"
                 "          [--gather bgrx|all] [--frames N] [--cancel-after-ms T] [--probe-ray ox,oy,oz,dx,dy,dz] [--bmp base] [--planes out.bin]\n", exe);
    return 2;
}

static bool parse(int argc, char** argv, Args& a)
{
    if (argc >= 2 && std::strncmp(argv[1], "--", 2) != 0)   // positional form
    {
        a.obj = argv[1];
        if (argc > 3) { a.W = (uint32)std::atoi(argv[2]); a.H = (uint32)std::atoi(argv[3]); }
        if (argc > 4) a.rescale = (f32)std::atof(argv[4]);
        if (argc > 5 && argv[5][0] && argv[5][0] != '-') a.bmp = argv[5];
        if (argc > 6) a.planes = argv[6];
        return true;
    }
    for (int i = 1; i < argc; ++i)
    {
        const std::string k = argv[i];
        if (i + 1 >= argc) return false;
        const char* v = argv[++i];
        if (k == "--obj") a.obj = v;
        else if (k == "--probe-ray")
        {
            a.probe_ray = std::sscanf(v, "%f,%f,%f,%f,%f,%f", &a.probe[0], &a.probe[1], &a.probe[2], &a.probe[3], &a.probe[4], &a.probe[5]) == 6;
            if (!a.probe_ray) return false;
        }
        else if (k == "--res") { if (std::sscanf(v, "%ux%u", &a.W, &a.H) != 2) return false; }
        else if (k == "--spp") a.spp = (uint32)std::atoi(v);
        else if (k == "--aa") a.aa = std::atoi(v);
        else if (k == "--bounces") a.bounces = std::atoi(v);
        else if (k == "--leaf-cap") a.leaf_cap = (uint32)std::atoi(v);
        else if (k == "--method") a.sah = std::strcmp(v, "center") != 0;
        else if (k == "--gpus") a.gpus = std::atoi(v);
        else if (k == "--devices") a.devices = v;
        else if (k == "--mode") a.notree = std::strcmp(v, "notree") == 0;
        else if (k == "--fp") a.fast = std::strcmp(v, "fast") == 0;
        else if (k == "--rescale") a.rescale = (f32)std::atof(v);
        else if (k == "--loader-threads") a.loader_threads = std::atoi(v);
        else if (k == "--gather") a.all_planes = std::strcmp(v, "bgrx") != 0;
        else if (k == "--frames") a.frames = (uint32)std::atoi(v);
        else if (k == "--cancel-after-ms") a.cancel_after_ms = std::atoi(v);
        else if (k == "--bmp") a.bmp = v;
        else if (k == "--planes") a.planes = v;
        else return false;
    }
    return !a.obj.empty() && a.W > 0 && a.H > 0 && a.spp > 0 && a.gpus >= 1 && a.gpus <= 8 && a.frames >= 1;
}

int run(const Args& a);

int main(int argc, char** argv)
{
    Args a;
    if (!parse(argc, argv, a)) return usage(argv[0]);
    return run(a);
}

int run(const Args& a)
{
    const uint32 W = a.W, H = a.H;
    Texture texture;
    if (!Setup_Texture(texture, TextureFileType::BMP, W, H)) { std::fprintf(stderr, "Setup_Texture: %s\n", atray_last_error()); return 1; }
    ThreadPool thread_pool;
    thread_pool.threads.allocate(a.loader_threads > 0 ? a.loader_threads : 1);  // app.cpp:16

    Model model = {};
    load_model_data(model.data, a.obj.c_str(), thread_pool);
    if (!model.data.handle) { std::fprintf(stderr, "load_model_data: %s\n", atray_last_error()); return 1; }
    model.surrounding_aabb = get_AABB(model.data);
    if (a.rescale > 0) resize_scale(model, a.rescale);
    translate_to(model, {0.f, -15.f, -38.f});
    model.kd_tree.division_method = a.sah ? KD_Division_Method::SAH : KD_Division_Method::CENTER;
    model.kd_tree.max_no_faces_per_node = a.leaf_cap;

    Camera cm;
    RenderSettings rs;
    rs.anti_aliasing = a.aa ? TRUE : FALSE;
    rs.resolution.x = texture.bmb.width;
    rs.resolution.y = texture.bmb.height;
    rs.samples_per_pixel = a.spp;   // the reference's default is 5 spp x 5 bounces (app.cpp:81-85): --spp 5 --bounces 5 --aa 1
    rs.bounce_limit = a.bounces;
    set_camera(cm, {0.1f, 2.f, 0.f}, {-.1f, -0.5f, -1.f}, rs, 1.0f);

    Scene scene;
    Material skybox = {{0.3f, 0.4f, 0.5f}, {0.2f, 0.3f, 0.4f}, 0.3f};
    Material model_mat = {{0.4f, 0.2f, 0.2f}, {0.92f, 0.5f, 0.0f}, 0.3f};
    scene.materials.add(skybox);
    scene.materials.add(model_mat);
    model.data.material = &scene.materials[1];
    scene.models.add_nocpy(model);
    if (!a.devices.empty())
        for (const char* p = a.devices.c_str(); *p;)
        {
            char* end = nullptr;
            const long d = std::strtol(p, &end, 10);
            if (end == p) break;
            scene.gpu_devices.add((int32)d);
            p = *end == ',' ? end + 1 : end;
        }
    else if (a.gpus > 1)
    {
        int visible = 0;
        atray_device_count(&visible);
        for (int32 i = 0; i < a.gpus; ++i) scene.gpu_devices.add(visible >= 2 * a.gpus ? i * (visible / a.gpus) : i);
    }

    uint32 kd_tree_max_nodes;
    auto s0 = std::chrono::steady_clock::now();
    prep_scene(scene, kd_tree_max_nodes);
    if (!scene.gpu) { std::fprintf(stderr, "prep_scene: %s\n", atray_last_error()); return 1; }
    if (a.notree) atray_multi_scene_use_kd_tree(scene.gpus, 0);   // run-time form of building without USE_KD_TREE (renderer.h:8)
    const double prep_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - s0).count();

    std::vector<f32> t((size_t)W * H);
    std::vector<uint32> prim((size_t)W * H);
    RenderInfo info;
    info.camera_tex = &texture;
    info.camera = &cm;
    info.scene = &scene;
    info.hit_stack_capacity = kd_tree_max_nodes;
    info.leaf_stack_capacity = kd_tree_max_nodes;
    info.hit_t = a.all_planes ? t.data() : nullptr;
    info.hit_prim = a.all_planes ? prim.data() : nullptr;
    info.fp_mode = a.fast ? ATRAY_FP_FAST : ATRAY_FP_STRICT;

    double ms = 0.0, best = 1e30;
    int polls = 0;
    bool cancelled = false;
    for (uint32 frame = 0; frame < a.frames && !cancelled; ++frame)
    {
        info.total_ray_casts = 0;
        polls = 0;
        auto t0 = std::chrono::steady_clock::now();
        start_render_from_camera(info, thread_pool);
        if (a.cancel_after_ms >= 0)
        {
            // the cancel of app.cpp:165-172: the window was closed while the frame was rendering
            std::this_thread::sleep_for(std::chrono::milliseconds(a.cancel_after_ms));
            const bool was_running = wait_for_render_from_camera_to_finish(info, thread_pool, 0) == TRUE;
            auto c0 = std::chrono::steady_clock::now();
            info.twq.clear();
            wait_for_pool(thread_pool, 0xffffffffu);
            const double cms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - c0).count();
            std::printf("cancel after %d ms: frame was %s, twq.clear() + wait_for_pool returned in %.3f ms, still running: %d\n", a.cancel_after_ms,
                        was_running ? "running" : "finished", cms, (int)wait_for_render_from_camera_to_finish(info, thread_pool, 0));
            cancelled = true;
            break;
        }
        while (wait_for_render_from_camera_to_finish(info, thread_pool, 33)) ++polls;
        ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (ms < best) best = ms;
        if (info.last_error != ATRAY_OK) { std::fprintf(stderr, "render: %s\n", atray_last_error()); return 1; }
    }
    if (!cancelled)
        std::printf("nodes %u | %ux%u | gpus %d | tiles %d/%d | total_ray_casts %lld | %.3f ms wall (%d polls) | best of %u: %.3f ms | prep %.1f ms\n",
                    kd_tree_max_nodes, W, H, atray_multi_count(scene.gpus), (int)info.twq.jobs_done, (int)info.twq.jobs.size,
                    info.total_ray_casts, ms, polls, a.frames, best, prep_ms);
    if (a.probe_ray)
    {
        Optimized_Ray r;
        r.ray.origin = {a.probe[0], a.probe[1], a.probe[2]};
        r.ray.direction = {a.probe[3], a.probe[4], a.probe[5]};
        r.inv_ray_d = {1 / r.ray.direction.x, 1 / r.ray.direction.y, 1 / r.ray.direction.z};   // renderer.cpp:41-44
        r.inv_signs = {r.inv_ray_d.x < 0, r.inv_ray_d.y < 0, r.inv_ray_d.z < 0};
        TriangleIntersectionData td = {ATRAY_MISS, -1.f, -1.f};
        const f32 dist = get_ray_kd_tree_intersection(r, scene.models[0].kd_tree, td, nullptr, nullptr);
        uint32 tb, ub, vb;
        std::memcpy(&tb, &dist, 4); std::memcpy(&ub, &td.u, 4); std::memcpy(&vb, &td.v, 4);
        std::printf("probe t_bits %08x face %u u_bits %08x v_bits %08x\n", tb, td.face_index, ub, vb);
    }
    if (!a.bmp.empty()) Write_To_File(texture, a.bmp.c_str());
    if (!a.planes.empty())
    {
        FILE* f = std::fopen(a.planes.c_str(), "wb");
        std::fwrite(texture.bmb.buffer_memory, 4, (size_t)W * H, f);
        std::fwrite(t.data(), 4, (size_t)W * H, f);
        std::fwrite(prim.data(), 4, (size_t)W * H, f);
        std::fclose(f);
    }
    info.twq.on_clear = nullptr;
    info.twq.jobs.clear();
    free_scene_memory(scene);
    Free_Texture(texture);
    thread_pool.threads.clear();
    return 0;
}
