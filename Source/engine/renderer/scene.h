// Scene (engine/renderer/scene.h:13-24) restricted to what the path uses (materials, models); spheres and planes are
// SURVEY §8(f) "next" rows. `gpu` is the device context prep_scene creates and free_scene_memory destroys.
#pragma once
#include "model.h"
struct Scene
{
    DBuffer<Material> materials;  // materials[0] is the sky (renderer.cpp:152)
    DBuffer<Model> models;
    atray_ctx* gpu = nullptr;
    int gpu_device = 0;
};
inline void free_scene_memory(Scene& scene)  // scene.cpp:3-30
{
    for (int32 i = 0; i < scene.models.length; ++i) { atray_model_free(scene.models[i].data.handle); scene.models[i].data.handle = nullptr; }
    scene.models.clear_buffer();
    scene.materials.clear_buffer();
    if (scene.gpu) atray_ctx_destroy(scene.gpu);
    scene.gpu = nullptr;
}
