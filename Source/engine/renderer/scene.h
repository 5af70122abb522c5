// Scene (engine/renderer/scene.h:13-24): materials, models, spheres, planes. `gpu` is the device context prep_scene creates
// and free_scene_memory destroys.
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
    atray_ctx* gpu = nullptr;
    int gpu_device = 0;
};
inline void free_scene_memory(Scene& scene)  // scene.cpp:3-30
{
    for (int32 i = 0; i < scene.models.length; ++i) { atray_model_free(scene.models[i].data.handle); scene.models[i].data.handle = nullptr; }
    scene.models.clear_buffer();
    scene.materials.clear_buffer();
    scene.spheres.clear_buffer();
    scene.planes.clear_buffer();
    if (scene.gpu) atray_ctx_destroy(scene.gpu);
    scene.gpu = nullptr;
}
