// WorkQueue (engine/tools/work_queue.h:6-16): jobs + jobs_done progress counter read by the caller (app.cpp:174-181).
// clear() is the reference's cancel (app.cpp:165-172: the pool threads find the queue empty after the tile they are on). The
// jobs here are progress records, the work itself runs in a kernel, so clear() also has to reach the device: start_render_from_camera
// registers the context's cancel entry (atray_render_cancel through `on_clear`), which raises the flag the kernel's persistent loops
// poll and returns once the streams have drained; wait_for_pool(tpool, UINT32MAX) after it returns at once, as in the reference.
#pragma once
#include "thread_pool.h"
template <typename t, typename size_type = int32> struct WorkQueue
{
    FDBuffer<t, volatile int32> jobs;
    volatile size_type jobs_done = 0;
    void (*on_clear)(void*) = nullptr;   // set by start_render_from_camera
    void* on_clear_arg = nullptr;
    void clear()
    {
        jobs.size = 0;
        if (on_clear) on_clear(on_clear_arg);
        jobs.clear();
    }
};
