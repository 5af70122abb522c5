// Ray / Optimized_Ray (engine/renderer/ray.h:7-22) as plain data: the single-ray query of kd_tree.h takes them. The slab test's
// inverse direction and sign bits are recomputed on the device from ray.direction (renderer.cpp:41-44 does the same per query).
#pragma once
#include "PL/PL_math.h"
static constexpr f32 tolerance = 0.0001f;   // ray.h:5
typedef Vec3<int32> vec3i;
struct Ray { vec3f origin, direction; };
struct Optimized_Ray
{
    Ray ray;
    vec3f inv_ray_d;
    vec3i inv_signs;
};
