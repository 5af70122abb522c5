// Plane (engine/renderer/plane.h:5-10): `distance` along the normal from the origin. Intersection (plane.h:12-22) in CUDA.
#pragma once
#include "material.h"
struct Plane { vec3f normal = {0.f, 0.f, 0.f}; f32 distance = 0; Material* material = nullptr; };
