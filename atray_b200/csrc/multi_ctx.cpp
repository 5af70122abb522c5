// Several GPUs of one node driven from ONE process (include/atray_b200.h: atray_ctx_create_multi, atray_multi_*): the reference's
// tile thread pool shares one address space and one bitmap (renderer.cpp:371-455); this is the same shape with GPUs in place of pool
// threads. Built on the library's own public C ABI only: a context per device, the scene replicated on each, band b of band_rows
// scanlines rendered by context b % n, every context copying exactly its scanlines into the caller's host frame (each over its own
// PCIe link) or storing them straight into planes on the first device (peer stores over NVLink), and the per-frame view pass shared
// between the contexts (atray_view_share_*, blocks connected by pointer after atray_ctx_enable_peer).
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/atray_b200.h"

namespace atray { int set_error(int code, const char* msg); }   // capi.cpp: fills atray_last_error() of the calling thread
using atray::set_error;

// One issuing thread per context after the first: queueing a frame on a GPU (three kernel launches, up to a dozen waits and copies
// of the slab pipeline) takes tens of microseconds of host time, and eight of them in a row from one thread were as long as the frame
// (0.50 ms per 4K frame on 8 GPUs against 0.41 ms on 4). The workers sleep on a condition variable between frames.
struct IssueWorker
{
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<int()> job;      // set by the caller, run once by the worker
    bool has_job = false, done = false, quit = false;
    int rc = ATRAY_OK;
    std::string err;

    IssueWorker()
    {
        th = std::thread([this] {
            std::unique_lock<std::mutex> lock(m);
            for (;;)
            {
                cv.wait(lock, [this] { return has_job || quit; });
                if (quit) return;
                has_job = false;
                lock.unlock();
                const int r = job();
                std::string e = r != ATRAY_OK ? atray_last_error() : "";
                lock.lock();
                rc = r;
                err.swap(e);
                done = true;
                cv.notify_all();
            }
        });
    }
    ~IssueWorker()
    {
        {
            std::lock_guard<std::mutex> lock(m);
            quit = true;
        }
        cv.notify_all();
        if (th.joinable()) th.join();
    }
    void start(std::function<int()> f)
    {
        std::lock_guard<std::mutex> lock(m);
        job = std::move(f);
        has_job = true;
        done = false;
        cv.notify_all();
    }
    int wait()
    {
        std::unique_lock<std::mutex> lock(m);
        cv.wait(lock, [this] { return done; });
        return rc;
    }
};

struct atray_multi
{
    std::vector<std::unique_ptr<IssueWorker>> workers;   // workers[i - 1] issues for context i (ATRAY_MULTI_THREADS=0: none)
    std::vector<atray_ctx*> ctx;
    std::vector<int> device;
    bool has_scene = false;
    bool peers_on = false;        // atray_ctx_enable_peer done for every ordered pair
    int share_w = 0, share_h = 0; // frame size the shared view pass is connected for (0: not connected)
    bool share_wanted = false;    // ATRAY_MULTI_SHARE=1: share the per-frame view pass between the contexts (measured slower than every
                                  // context computing the whole pass itself at 2, 4 and 8 GPUs: profiles/r02i, r02j — opt-in)
    int in_flight = 0;            // contexts whose frame has not been collected by atray_multi_render_poll
    std::vector<char> running;    // per context
    int32_t rows_total = 0;
    int32_t band_rows = 8;
};

namespace {

int enable_peers(atray_multi* m)
{
    if (m->peers_on) return ATRAY_OK;
    for (size_t i = 0; i < m->ctx.size(); ++i)
        for (size_t j = 0; j < m->ctx.size(); ++j)
            if (i != j)
            {
                int rc = atray_ctx_enable_peer(m->ctx[i], m->ctx[j]);
                if (rc != ATRAY_OK) return rc;
            }
    m->peers_on = true;
    return ATRAY_OK;
}

// (re)connect the shared view pass for frames up to W x H; a failure (no peer access) leaves every context on its own pass
void connect_share(atray_multi* m, int W, int H)
{
    const int n = int(m->ctx.size());
    if (n < 2 || !m->share_wanted || (W <= m->share_w && H <= m->share_h)) return;
    // two contexts on one device would have kernels of the same GPU waiting for each other's flags: no shared pass then
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j)
            if (m->device[size_t(i)] == m->device[size_t(j)]) { m->share_wanted = false; return; }
    for (atray_ctx* c : m->ctx) atray_view_share_off(c);
    m->share_w = m->share_h = 0;
    if (enable_peers(m) != ATRAY_OK) { m->share_wanted = false; return; }
    std::vector<void*> blocks(n, nullptr);
    for (int i = 0; i < n; ++i)
        if (atray_view_share_create(m->ctx[i], W, H, &blocks[i], nullptr) != ATRAY_OK) { m->share_wanted = false; return; }
    for (int i = 0; i < n; ++i)
        if (atray_view_share_connect(m->ctx[i], i, n, blocks.data()) != ATRAY_OK)
        {
            for (atray_ctx* c : m->ctx) atray_view_share_off(c);
            m->share_wanted = false;
            return;
        }
    m->share_w = W;
    m->share_h = H;
}

}  // namespace

extern "C" {

int atray_ctx_create_multi(const int32_t* devices, int32_t n, atray_multi** out)
{
    if (!out || n < 1 || n > 8) return set_error(ATRAY_EINVAL, "atray_ctx_create_multi: 1 to 8 devices");
    int visible = 0;
    int rc = atray_device_count(&visible);
    if (rc != ATRAY_OK) return rc;
    atray_multi* m = new atray_multi();
    for (int i = 0; i < n; ++i)
    {
        // default placement: spread over the visible devices when there are at least twice as many (on the 8-GPU B200 boxes GPUs
        // 0-3 and 4-7 reach host memory through different paths: 2 or 4 contexts should sit on both halves), else 0..n-1
        const int dev = devices ? devices[i] : ((n > 1 && visible >= 2 * n) ? i * (visible / n) : i);
        atray_ctx* c = nullptr;
        rc = atray_ctx_create(dev, &c);
        if (rc != ATRAY_OK) { atray_multi_destroy(m); return rc; }
        m->ctx.push_back(c);
        m->device.push_back(dev);
    }
    m->running.assign(size_t(n), 0);
    bool threads = true;
    if (const char* env = std::getenv("ATRAY_MULTI_THREADS")) threads = std::atoi(env) != 0;
    if (threads)
        for (int i = 1; i < n; ++i) m->workers.emplace_back(new IssueWorker());
    if (const char* env = std::getenv("ATRAY_MULTI_SHARE")) m->share_wanted = std::atoi(env) != 0;
    *out = m;
    return ATRAY_OK;
}

void atray_multi_destroy(atray_multi* m)
{
    if (!m) return;
    m->workers.clear();   // joins the issuing threads
    for (atray_ctx* c : m->ctx) atray_ctx_destroy(c);
    delete m;
}

int atray_multi_count(const atray_multi* m) { return m ? int(m->ctx.size()) : 0; }

atray_ctx* atray_multi_ctx(atray_multi* m, int32_t i) { return (m && i >= 0 && size_t(i) < m->ctx.size()) ? m->ctx[size_t(i)] : nullptr; }

int atray_multi_device(const atray_multi* m, int32_t i) { return (m && i >= 0 && size_t(i) < m->device.size()) ? m->device[size_t(i)] : -1; }

int atray_multi_scene_upload(atray_multi* m, const atray_model* model, const AtrayMaterial* sky, const AtrayMaterial* model_mat)
{
    if (!m || !model || !sky || !model_mat) return set_error(ATRAY_EINVAL, "null argument");
    if (m->in_flight) return set_error(ATRAY_ESTATE, "render in flight");
    // the uploads are independent (one flatten + a few copies per device): one host thread per context
    std::vector<int> rc(m->ctx.size(), ATRAY_OK);
    std::vector<std::string> err(m->ctx.size());
    std::vector<std::thread> th;
    for (size_t i = 1; i < m->ctx.size(); ++i)
        th.emplace_back([&, i] { rc[i] = atray_scene_upload(m->ctx[i], model, sky, model_mat); if (rc[i] != ATRAY_OK) err[i] = atray_last_error(); });
    rc[0] = atray_scene_upload(m->ctx[0], model, sky, model_mat);
    for (std::thread& t : th) t.join();
    m->share_w = m->share_h = 0;   // the blocks are sized by the scene's slot count
    for (size_t i = 0; i < rc.size(); ++i)
        if (rc[i] != ATRAY_OK) return i ? set_error(rc[i], err[i].c_str()) : rc[i];
    m->has_scene = true;
    return ATRAY_OK;
}

int atray_multi_scene_set_primitives(atray_multi* m, const AtraySphere* spheres, uint32_t n_spheres, const AtrayPlane* planes, uint32_t n_planes)
{
    if (!m) return set_error(ATRAY_EINVAL, "null argument");
    for (atray_ctx* c : m->ctx)
    {
        int rc = atray_scene_set_primitives(c, spheres, n_spheres, planes, n_planes);
        if (rc != ATRAY_OK) return rc;
    }
    return ATRAY_OK;
}

int atray_multi_scene_use_kd_tree(atray_multi* m, int use_kd_tree)
{
    if (!m) return set_error(ATRAY_EINVAL, "null argument");
    for (atray_ctx* c : m->ctx)
    {
        int rc = atray_scene_use_kd_tree(c, use_kd_tree);
        if (rc != ATRAY_OK) return rc;
    }
    return ATRAY_OK;
}

int atray_multi_render_begin(atray_multi* m, const AtrayCamera* cam, const AtrayRenderOpts* opts, const AtrayFrame* host_out)
{
    if (!m || !cam || !opts || !host_out) return set_error(ATRAY_EINVAL, "null argument");
    if (!m->has_scene) return set_error(ATRAY_ESTATE, "render before scene_upload");
    if (m->in_flight) return set_error(ATRAY_ESTATE, "a render is already in flight");
    const int n = int(m->ctx.size());
    AtrayRenderOpts o = *opts;
    if (n == 1)
    {
        int rc = atray_render_begin(m->ctx[0], cam, &o, host_out);
        if (rc != ATRAY_OK) return rc;
        m->running[0] = 1;
        m->in_flight = 1;
        m->rows_total = cam->height;
        return ATRAY_OK;
    }
    if (o.part.band_rows <= 0) o.part.band_rows = 8;
    m->band_rows = o.part.band_rows;
    connect_share(m, cam->width, cam->height);
    // every context queues its frame (it copies exactly rank i's scanlines into the shared host frame); contexts 1 … n-1 are
    // issued by their worker threads while this thread issues context 0
    std::vector<int> rcs((size_t)n, ATRAY_OK);
    std::vector<std::string> errs((size_t)n);
    auto issue = [m, cam, o, host_out, n](int i) -> int {
        AtrayRenderOpts oi = o;
        oi.part.rank = i;
        oi.part.world = n;
        return atray_render_begin(m->ctx[size_t(i)], cam, &oi, host_out);
    };
    const bool threaded = m->workers.size() == size_t(n - 1);
    if (threaded)
        for (int i = 1; i < n; ++i) m->workers[size_t(i - 1)]->start([issue, i] { return issue(i); });
    rcs[0] = issue(0);
    if (rcs[0] != ATRAY_OK) errs[0] = atray_last_error();
    for (int i = 1; i < n; ++i)
    {
        if (threaded) { rcs[size_t(i)] = m->workers[size_t(i - 1)]->wait(); errs[size_t(i)] = m->workers[size_t(i - 1)]->err; }
        else { rcs[size_t(i)] = issue(i); if (rcs[size_t(i)] != ATRAY_OK) errs[size_t(i)] = atray_last_error(); }
    }
    int bad = -1;
    for (int i = 0; i < n && bad < 0; ++i)
        if (rcs[size_t(i)] != ATRAY_OK) bad = i;
    if (bad >= 0)
    {
        for (int j = 0; j < n; ++j)
            if (rcs[size_t(j)] == ATRAY_OK) atray_render_cancel(m->ctx[size_t(j)]);
        for (int j = 0; j < n; ++j) m->running[size_t(j)] = 0;
        for (atray_ctx* c : m->ctx) atray_view_share_off(c);   // the contexts' frame counters have diverged: reconnect next time
        m->share_w = m->share_h = 0;
        m->in_flight = 0;
        return set_error(rcs[size_t(bad)], errs[size_t(bad)].c_str());
    }
    for (int i = 0; i < n; ++i) m->running[size_t(i)] = 1;
    m->in_flight = n;
    m->rows_total = cam->height;
    return ATRAY_OK;
}

int atray_multi_render_poll(atray_multi* m, uint32_t timeout_ms)
{
    if (!m) return set_error(ATRAY_EINVAL, "null argument");
    if (!m->in_flight) return 0;
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
    for (size_t i = 0; i < m->ctx.size(); ++i)
    {
        if (!m->running[i]) continue;
        uint32_t left = 0xffffffffu;
        if (timeout_ms != 0xffffffffu)
        {
            const auto now = std::chrono::steady_clock::now();
            left = now >= deadline ? 0u : uint32_t(std::chrono::duration_cast<std::chrono::milliseconds>(deadline - now).count());
        }
        int r = atray_render_poll(m->ctx[i], left);
        if (r == 1) return 1;   // still running
        m->running[i] = 0;
        m->in_flight -= 1;
        if (r < 0)
        {
            for (size_t j = 0; j < m->ctx.size(); ++j)
                if (m->running[j]) { atray_render_cancel(m->ctx[j]); m->running[j] = 0; }
            m->in_flight = 0;
            return r;
        }
    }
    m->in_flight = 0;
    return 0;
}

int atray_multi_render_progress(atray_multi* m, int32_t* rows_done, int32_t* rows_total)
{
    if (!m || !rows_done || !rows_total) return set_error(ATRAY_EINVAL, "null argument");
    *rows_total = m->rows_total;
    if (m->ctx.size() == 1) return atray_render_progress(m->ctx[0], rows_done, rows_total);
    // a context's scanlines land together when its partitioned frame completes: count the bands of the finished contexts
    const int n = int(m->ctx.size());
    const int n_bands = (m->rows_total + m->band_rows - 1) / m->band_rows;
    int64_t rows = 0;
    for (int i = 0; i < n; ++i)
    {
        bool done = !m->running[size_t(i)];
        if (!done)
        {
            int32_t rd = 0, rt = 0;
            if (atray_render_progress(m->ctx[size_t(i)], &rd, &rt) == ATRAY_OK && atray_render_poll(m->ctx[size_t(i)], 0) == 0)
            {
                m->running[size_t(i)] = 0;
                m->in_flight -= 1;
                done = true;
            }
        }
        if (done)
        {
            const int mine = i < n_bands ? (n_bands - i + n - 1) / n : 0;
            rows += int64_t(mine) * m->band_rows;
        }
    }
    *rows_done = rows > m->rows_total ? m->rows_total : int32_t(rows);
    return ATRAY_OK;
}

int atray_multi_render_cancel(atray_multi* m)
{
    if (!m) return set_error(ATRAY_EINVAL, "null argument");
    int cut = 0;
    for (size_t i = 0; i < m->ctx.size(); ++i)
    {
        int r = atray_render_cancel(m->ctx[i]);
        if (r < 0) return r;
        cut |= r;
        m->running[i] = 0;
    }
    m->in_flight = 0;
    return cut;
}

int atray_multi_render_device(atray_multi* m, const AtrayCamera* cam, const AtrayRenderOpts* opts, const AtrayFrame* device_out)
{
    if (!m || !cam || !opts || !device_out) return set_error(ATRAY_EINVAL, "null argument");
    if (!m->has_scene) return set_error(ATRAY_ESTATE, "render before scene_upload");
    if (m->in_flight) return set_error(ATRAY_ESTATE, "a render is already in flight");
    const int n = int(m->ctx.size());
    AtrayRenderOpts o = *opts;
    if (n == 1) return atray_render_device(m->ctx[0], cam, &o, device_out);
    if (o.part.band_rows <= 0) o.part.band_rows = 8;
    int rc = enable_peers(m);   // the planes live on one device: the other contexts' kernels store into them over NVLink
    if (rc != ATRAY_OK) return rc;
    connect_share(m, cam->width, cam->height);
    for (int i = 0; i < n; ++i)
    {
        o.part.rank = i;
        o.part.world = n;
        rc = atray_render_device(m->ctx[size_t(i)], cam, &o, device_out);
        if (rc != ATRAY_OK) return rc;
    }
    return ATRAY_OK;
}

int atray_multi_sync(atray_multi* m)
{
    if (!m) return set_error(ATRAY_EINVAL, "null argument");
    for (atray_ctx* c : m->ctx)
    {
        int rc = atray_ctx_sync(c);
        if (rc != ATRAY_OK) return rc;
    }
    return ATRAY_OK;
}

}  // extern "C"
