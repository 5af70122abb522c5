// ThreadPool (engine/tools/thread_pool.h:4-26 of the reference). The GPU renderer does not use it (start_render_from_camera keeps
// the parameter for drop-in compatibility; load_model_data uses threads.size as the loader's chunk count), but applications written
// against the reference call activate_pool / wait_for_pool themselves (app.cpp:169 after a cancel, OBJ_loader.cpp:332-333), so both
// exist with the reference's semantics on std::thread: activate_pool starts one thread per slot on `proc(data)`; wait_for_pool
// returns TRUE when the wait timed out and FALSE when every thread has finished (pl_wait_for_all_threads, pl_utils.h:170). A pool
// that was never activated has nothing to wait for and reports FALSE at once, which is what app.cpp:169 sees after info.twq.clear().
#pragma once
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>
#include "PL/PL_math.h"
typedef void (*ThreadProc)(void*);   // pl_utils.h:158
struct ThreadState
{
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
};
struct ThreadHandle { void* thread_handle = nullptr; };   // pl_utils.h:152-156 (here: ThreadState*)
struct Thread { ThreadHandle handle; };
struct ThreadPool { FDBuffer<Thread> threads; };

inline void release_thread(ThreadHandle& h)   // pl_close_thread, pl_utils.h:164: joins a finished thread and frees its state
{
    ThreadState* st = (ThreadState*)h.thread_handle;
    if (!st) return;
    if (st->th.joinable()) st->th.join();
    delete st;
    h.thread_handle = nullptr;
}

inline void activate_pool(ThreadPool& pool, ThreadProc proc, void* data)   // thread_pool.h:14-20
{
    for (int32 i = 0; i < pool.threads.size; i++)
    {
        release_thread(pool.threads[i].handle);   // a slot that ran before: its thread has been waited for
        ThreadState* st = new ThreadState();
        st->th = std::thread([st, proc, data]() {
            proc(data);
            std::lock_guard<std::mutex> lock(st->m);
            st->done = true;
            st->cv.notify_all();
        });
        pool.threads[i].handle.thread_handle = st;
    }
}

// thread_pool.h:22-26: TRUE if the wait timed out, FALSE if all threads are finished
inline b32 wait_for_pool(ThreadPool& pool, uint32 time_out_in_ms)
{
    const bool forever = time_out_in_ms == 0xffffffffu;
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(forever ? 0u : time_out_in_ms);
    for (int32 i = 0; i < pool.threads.size; i++)
    {
        ThreadState* st = (ThreadState*)pool.threads[i].handle.thread_handle;
        if (!st) continue;
        std::unique_lock<std::mutex> lock(st->m);
        if (forever) st->cv.wait(lock, [st] { return st->done; });
        else if (!st->cv.wait_until(lock, deadline, [st] { return st->done; })) return TRUE;
    }
    for (int32 i = 0; i < pool.threads.size; i++) release_thread(pool.threads[i].handle);
    return FALSE;
}
