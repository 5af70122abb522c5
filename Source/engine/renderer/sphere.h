// Sphere (engine/renderer/sphere.h:5-10). The intersection itself (sphere.h:12-39) runs in the CUDA kernel.
#pragma once
#include "material.h"
struct Sphere { vec3f center = {0.f, 0.f, 0.f}; f32 radius = 0; Material* material = nullptr; };
