// Include-path compatibility: the reference's applications include "engine/renderer/material.h" for Material.
#pragma once
#include "render_types.h"
