#pragma once
#include <cstddef>
#include "PL/PL_base_defs.h"
