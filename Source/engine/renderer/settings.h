// Include-path compatibility: the reference's applications include "engine/renderer/settings.h" for RenderSettings.
#pragma once
#include "render_types.h"
