// ThreadPool (engine/tools/thread_pool.h:9-26 of the reference): kept in the signatures for drop-in compatibility.
// The GPU renderer ignores it; load_model_data uses threads.size as the loader's chunk count.
#pragma once
#include "PL/PL_math.h"
struct ThreadHandle { void* thread_handle = nullptr; };
struct Thread { ThreadHandle handle; };
struct ThreadPool { FDBuffer<Thread> threads; };
