// Plain-data records shared by the host layer and the C ABI. Field order and sizes are part of the drop-in contract:
// applications written against the reference fill these structs member by member (app.cpp:81-96), and Material is
// memcpy-compatible with AtrayMaterial (28 bytes: two rgb triples and the scatter weight).
#pragma once
#include "PL/PL_math.h"
#include "atray_b200.h"

// what app.cpp:91,96 aggregate-initialises as { emission, reflection, scatter }
struct Material
{
    vec3f emission_color, reflection_color;
    f32 scatter;
};
static_assert(sizeof(Material) == sizeof(AtrayMaterial), "Material must stay memcpy-compatible with AtrayMaterial");

// what app.cpp:81-85 fills before set_camera; atray_camera_set receives the four values as arguments
struct RenderSettings
{
    vec2i resolution;            // pixels
    b32 anti_aliasing;           // jitter each sample inside its pixel (renderer.cpp:336-347)
    uint32 samples_per_pixel;
    int32 bounce_limit;          // 1 = primary rays only
};
