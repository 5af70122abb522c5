// Headless replay of the reference's render_app call sequence (engine/app.cpp:59-195) against the drop-in host layer:
// load → get_AABB → translate_to → set_camera → materials/scene → prep_scene → start_render_from_camera →
// poll wait_for_render_from_camera_to_finish(33 ms) → stats (+ optional BMP and raw plane dumps).
//   app_headless <obj> [W H] [rescale] [bmp_base] [out_planes.bin]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "engine/renderer/renderer.h"
#include "engine/tools/OBJ_loader.h"

int main(int argc, char** argv)
{
    if (argc < 2) { std::fprintf(stderr, "usage: %s model.obj [W H] [rescale] [bmp_base] [planes.bin]\n", argv[0]); return 2; }
    const uint32 W = argc > 3 ? (uint32)std::atoi(argv[2]) : 1280, H = argc > 3 ? (uint32)std::atoi(argv[3]) : 720;
    const f32 rescale = argc > 4 ? (f32)std::atof(argv[4]) : 0.f;
    Texture texture;
    if (!Setup_Texture(texture, TextureFileType::BMP, W, H)) { std::fprintf(stderr, "Setup_Texture: %s\n", atray_last_error()); return 1; }
    ThreadPool thread_pool;
    thread_pool.threads.allocate(8);  // app.cpp:16

    Model model = {};
    load_model_data(model.data, argv[1], thread_pool);
    if (!model.data.handle) { std::fprintf(stderr, "load_model_data: %s\n", atray_last_error()); return 1; }
    model.surrounding_aabb = get_AABB(model.data);
    if (rescale > 0) resize_scale(model, rescale);
    translate_to(model, {0.f, -15.f, -38.f});
    model.kd_tree.division_method = KD_Division_Method::SAH;
    model.kd_tree.max_no_faces_per_node = 300;

    Camera cm;
    RenderSettings rs;
    rs.anti_aliasing = FALSE;
    rs.resolution.x = texture.bmb.width;
    rs.resolution.y = texture.bmb.height;
    rs.samples_per_pixel = 1;  // primary-ray path (the reference default is 5 spp × 5 bounces: SURVEY §8f "next")
    rs.bounce_limit = 1;
    set_camera(cm, {0.1f, 2.f, 0.f}, {-.1f, -0.5f, -1.f}, rs, 1.0f);

    Scene scene;
    Material skybox = {{0.3f, 0.4f, 0.5f}, {0.2f, 0.3f, 0.4f}, 0.3f};
    Material model_mat = {{0.4f, 0.2f, 0.2f}, {0.92f, 0.5f, 0.0f}, 0.3f};
    scene.materials.add(skybox);
    scene.materials.add(model_mat);
    model.data.material = &scene.materials[1];
    scene.models.add_nocpy(model);

    uint32 kd_tree_max_nodes;
    prep_scene(scene, kd_tree_max_nodes);
    if (!scene.gpu) { std::fprintf(stderr, "prep_scene: %s\n", atray_last_error()); return 1; }

    std::vector<f32> t((size_t)W * H);
    std::vector<uint32> prim((size_t)W * H);
    RenderInfo info;
    info.camera_tex = &texture;
    info.camera = &cm;
    info.scene = &scene;
    info.hit_stack_capacity = kd_tree_max_nodes;
    info.leaf_stack_capacity = kd_tree_max_nodes;
    info.hit_t = t.data();
    info.hit_prim = prim.data();

    auto t0 = std::chrono::steady_clock::now();
    start_render_from_camera(info, thread_pool);
    int polls = 0;
    while (wait_for_render_from_camera_to_finish(info, thread_pool, 33)) ++polls;
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (info.last_error != ATRAY_OK) { std::fprintf(stderr, "render: %s\n", atray_last_error()); return 1; }
    std::printf("nodes %u | %ux%u | tiles %d/%d | total_ray_casts %lld | %.3f ms wall (%d polls)\n", kd_tree_max_nodes, W, H,
                (int)info.twq.jobs_done, (int)info.twq.jobs.size, info.total_ray_casts, ms, polls);
    if (argc > 5 && argv[5][0] && argv[5][0] != '-') Write_To_File(texture, argv[5]);
    if (argc > 6)
    {
        FILE* f = std::fopen(argv[6], "wb");
        std::fwrite(texture.bmb.buffer_memory, 4, (size_t)W * H, f);
        std::fwrite(t.data(), 4, (size_t)W * H, f);
        std::fwrite(prim.data(), 4, (size_t)W * H, f);
        std::fclose(f);
    }
    info.twq.jobs.clear();
    free_scene_memory(scene);
    Free_Texture(texture);
    thread_pool.threads.clear();
    return 0;
}
