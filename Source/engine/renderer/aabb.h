// AABB (engine/renderer/aabb.h:5-17): min then max, 24 bytes. The ray/box tests themselves run on the GPU.
#pragma once
#include "PL/PL_math.h"
struct AABB { vec3f min; vec3f max; };
