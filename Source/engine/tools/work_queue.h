// WorkQueue (engine/tools/work_queue.h:6-16): jobs + jobs_done progress counter read by the caller (app.cpp:174-181).
#pragma once
#include "thread_pool.h"
template <typename t, typename size_type = int32> struct WorkQueue
{
    FDBuffer<t, volatile int32> jobs;
    volatile size_type jobs_done = 0;
    void clear() { jobs.size = 0; jobs.clear(); }  // cancel path, app.cpp:165-172
};
