// Overlay forwarder (oracle/ref build only): the reference includes "pl_base_defs.h"
// but the file on disk is PL/PL_base_defs.h (case-insensitive FS assumed upstream).
#pragma once
#include <cstddef>
#include "PL/PL_base_defs.h"
