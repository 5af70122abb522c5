// Scene (engine/renderer/scene.h:13-24): materials, models, spheres, planes. `gpus` is the device side prep_scene creates and
// free_scene_memory destroys: one context per entry of `gpu_devices` (empty: the single device `gpu_device`), driven together as the
// reference drives its pool threads (atray_multi, include/atray_b200.h). `gpu` stays the first context for per-context calls.
#pragma once
#include "model.h"
#include "plane.h"
#include "sphere.h"
struct Scene
{
    DBuffer<Material> materials;  // materials[0] is the sky (renderer.cpp:152)
    DBuffer<Model> models;
    DBuffer<Sphere> spheres;
    DBuffer<Plane> planes;
    atray_multi* gpus = nullptr;
    atray_ctx* gpu = nullptr;
    int gpu_device = 0;
    DBuffer<int32> gpu_devices;   // e.g. {0, 4}: render every frame on these GPUs together
};
inline void free_scene_memory(Scene& scene)  // scene.cpp:3-30
{
    for (int32 i = 0; i < scene.models.length; ++i) { atray_model_free(scene.models[i].data.handle); scene.models[i].data.handle = nullptr; }
    scene.models.clear_buffer();
    scene.materials.clear_buffer();
    scene.spheres.clear_buffer();
    scene.planes.clear_buffer();
    scene.gpu_devices.clear_buffer();
    if (scene.gpus) atray_multi_destroy(scene.gpus);
    scene.gpus = nullptr;
    scene.gpu = nullptr;
}
