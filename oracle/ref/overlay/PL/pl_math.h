// Overlay forwarder: parser.h includes "PL/pl_math.h"; on disk it is PL/PL_math.h.
#pragma once
#include "PL/PL_math.h"
