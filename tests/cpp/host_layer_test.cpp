// CPU test of the host layer's pool and queue (Source/engine/tools): activate_pool / wait_for_pool with the reference's return
// convention (thread_pool.h:22-26: TRUE = timed out, FALSE = all finished) and WorkQueue::clear() reaching the registered hook.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <thread>
#include "engine/tools/work_queue.h"

static std::atomic<int> g_ran{0};
static void worker(void* data)
{
    std::this_thread::sleep_for(std::chrono::milliseconds(*(int*)data));
    g_ran += 1;
}
static int g_cleared = 0;
static void on_clear(void* p) { g_cleared += *(int*)p; }

#define CHECK(c) do { if (!(c)) { std::printf("FAILED line %d: %s\n", __LINE__, #c); return 1; } } while (0)

int main()
{
    ThreadPool pool;
    pool.threads.allocate(4);
    CHECK(wait_for_pool(pool, 0xffffffffu) == FALSE);   // never activated: nothing to wait for (app.cpp:169 after a cancel)
    int sleep_ms = 150;
    activate_pool(pool, worker, &sleep_ms);
    CHECK(wait_for_pool(pool, 10) == TRUE);              // still running: timed out
    CHECK(wait_for_pool(pool, 0xffffffffu) == FALSE);    // all finished
    CHECK(g_ran == 4);
    CHECK(wait_for_pool(pool, 0) == FALSE);
    sleep_ms = 1;
    activate_pool(pool, worker, &sleep_ms);              // slots are reusable
    CHECK(wait_for_pool(pool, 5000) == FALSE);
    CHECK(g_ran == 8);
    pool.threads.clear();

    struct Job { int v; };
    WorkQueue<Job> q;
    q.jobs.allocate(7);
    q.jobs_done = 3;
    int inc = 5;
    q.on_clear = on_clear;
    q.on_clear_arg = &inc;
    q.clear();
    CHECK(g_cleared == 5 && q.jobs.size == 0 && q.jobs.front == nullptr);
    q.on_clear = nullptr;
    q.clear();                                            // no hook: plain clear
    CHECK(g_cleared == 5);
    std::printf("ok\n");
    return 0;
}
