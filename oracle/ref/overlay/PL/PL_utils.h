// Overlay forwarder: thread_pool.h includes "PL/PL_utils.h"; on disk it is PL/pl_utils.h.
#pragma once
#include "PL/pl_utils.h"
