// Material (engine/renderer/material.h:4-9), same 28-byte layout as AtrayMaterial.
#pragma once
#include "PL/PL_math.h"
struct Material { vec3f emission_color; vec3f reflection_color; f32 scatter; };
