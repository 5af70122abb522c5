// Host prerequisites of the render path (fresh code; behaviour follows the reference so that the arrays handed to the
// GPU are bit-identical to what the reference's CPU renderer walks):
//   OBJ text → arrays ........ engine/tools/OBJ_loader.cpp:32-361 + utilities/parser.h:4-205 (+ tables PL_math.h:6-41)
//   model box / transforms ... engine/renderer/model.h:41-61, 108-121, 136-152
//   8-way tree build ......... engine/renderer/kd_tree.cpp:3-17 (helpers), 20-64, 67-288
//   camera basis ............. engine/renderer/camera.h:23-45, PL/PL_math.h:377-392 (normalize)
// Compile with -ffp-contract=off: every float expression below is evaluated in the reference's operation order.
#include "host_model.h"

#include <algorithm>

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

namespace atray {

static constexpr float kTolerance = 0.0001f;          // ray.h:5
static constexpr float kMaxFloat = 3.402823466e+38F;  // PL_base_defs.h:70

// ---------------------------------------------------------------------------------------------------------------
// number scanning (utilities/parser.h). The reference does not use strtod: mantissa digits are accumulated in a
// u64, converted to f64 once and scaled by one table entry; the float bits of every vertex depend on exactly that.
// ---------------------------------------------------------------------------------------------------------------
namespace {

const uint64_t kU64Pow10[20] = {1ULL, 10ULL, 100ULL, 1000ULL, 10000ULL, 100000ULL, 1000000ULL, 10000000ULL, 100000000ULL,
    1000000000ULL, 10000000000ULL, 100000000000ULL, 1000000000000ULL, 10000000000000ULL, 100000000000000ULL,
    1000000000000000ULL, 10000000000000000ULL, 100000000000000000ULL, 1000000000000000000ULL, 10000000000000000000ULL};

double pow10_entry(int e)  // F64_POWER_10[e + 28], e in [-28, 19]
{
    static const double tab[48] = {1.0e-28, 1.0e-27, 1.0e-26, 1.0e-25, 1.0e-24, 1.0e-23, 1.0e-22, 1.0e-21, 1.0e-20, 1.0e-19,
        1.0e-18, 1.0e-17, 1.0e-16, 1.0e-15, 1.0e-14, 1.0e-13, 1.0e-12, 1.0e-11, 1.0e-10, 1.0e-9, 1.0e-8, 1.0e-7, 1.0e-6,
        1.0e-5, 1.0e-4, 1.0e-3, 1.0e-2, 1.0e-1, 1.0e0, 1.0e1, 1.0e2, 1.0e3, 1.0e4, 1.0e5, 1.0e6, 1.0e7, 1.0e8, 1.0e9, 1.0e10,
        1.0e11, 1.0e12, 1.0e13, 1.0e14, 1.0e15, 1.0e16, 1.0e17, 1.0e18, 1.0e19};
    return tab[e + 28];
}

struct Scanner
{
    const char* p;
    static bool digit(char c) { return c >= '0' && c <= '9'; }
    void blanks() { while (*p == ' ' || *p == '\t' || *p == '\r') ++p; }
    int sign_char()
    {
        if (*p == '-') { ++p; return -1; }
        if (*p == '+') ++p;
        return 1;
    }
    int32_t integer()  // parse_int, parser.h:38-65
    {
        blanks();
        int s = sign_char();
        uint32_t acc = 0;
        for (; digit(*p); ++p) acc = acc * 10u + uint32_t(*p - '0');
        return int32_t(acc) * s;
    }
    double real()  // parse_f64, parser.h:113-191
    {
        blanks();
        double s = sign_char() < 0 ? -1.0 : 1.0;
        uint64_t mant = 0, frac = 0;
        int nfrac = 0;
        for (; digit(*p); ++p) mant = mant * 10 + uint64_t(*p - '0');
        if (*p == '.') ++p;
        for (; digit(*p); ++p) { frac = frac * 10 + uint64_t(*p - '0'); ++nfrac; }
        mant = mant * kU64Pow10[nfrac > 19 ? 19 : nfrac] + frac;
        int e10 = 0;
        if (*p == 'e' || *p == 'E')
        {
            ++p;
            int es = sign_char();
            for (; digit(*p); ++p) e10 = e10 * 10 + (*p - '0');
            e10 *= es;
        }
        e10 -= nfrac;
        if (e10 < -28 || e10 > 19) e10 = 0;
        double v = double(mant);
        v = v * s;
        v *= pow10_entry(e10);
        return v;
    }
    Vec3 triple()  // parse_vec3f, parser.h:194-205
    {
        Vec3 r;
        r.x = float(real());
        r.y = float(real());
        r.z = float(real());
        return r;
    }
    void corner(int32_t& v, int32_t& vt, int32_t& vn)  // one "v", "v/vt", "v//vn" or "v/vt/vn" group (OBJ_loader.cpp:88-145)
    {
        v = integer();
        if (*p != '/') return;
        ++p;
        if (*p == '/') { ++p; vn = integer(); return; }
        vt = integer();
        if (*p == '/') { ++p; vn = integer(); }
    }
};

struct ChunkOut { std::vector<Vec3> v, vn, vt; std::vector<Face> f; };

// One newline-aligned slice [b, e) of the file text (parse_chunks, OBJ_loader.cpp:32-176).
void scan_lines(const char* b, const char* e, ChunkOut& out)
{
    Scanner sc{b};
    while (sc.p < e)
    {
        char c = *sc.p;
        if (c == 'v')
        {
            ++sc.p;
            if (*sc.p == ' ') out.v.push_back(sc.triple());
            else if (*sc.p == 't') { ++sc.p; out.vt.push_back(sc.triple()); }
            else if (*sc.p == 'n') { ++sc.p; out.vn.push_back(sc.triple()); }
        }
        else if (c == 'f')
        {
            ++sc.p;
            Face fc;
            std::memset(&fc, 0, sizeof fc);
            for (int k = 0; k < 3; ++k) sc.corner(fc.v[k], fc.vt[k], fc.vn[k]);
            out.f.push_back(fc);
        }
        while (*sc.p != '\n') ++sc.p;
        ++sc.p;
    }
}

}  // namespace

int load_obj(const char* path, int threads, HostModel& m, std::string& err)
{
    FILE* fp = std::fopen(path, "rb");
    if (!fp) { err = std::string("cannot open ") + path; return -2; }
    std::fseek(fp, 0, SEEK_END);
    long sz = std::ftell(fp);
    std::fseek(fp, 0, SEEK_SET);
    std::vector<char> text(size_t(sz) + 2, 0);
    if (sz > 0 && std::fread(text.data(), 1, size_t(sz), fp) != size_t(sz)) { std::fclose(fp); err = "short read"; return -2; }
    std::fclose(fp);
    text[size_t(sz)] = '\n';  // terminating newline the reference writes past the last byte (OBJ_loader.cpp:331-332)

    // Newline-aligned slices, parsed concurrently and concatenated in file order (OBJ_loader.cpp:298-341, 190-227).
    // Unlike the reference, slice bounds never run past the buffer on tiny files (SURVEY §7 hard part 8).
    if (threads < 1) threads = 1;
    const char* base = text.data();
    const char* end = base + sz;
    std::vector<std::pair<const char*, const char*>> cuts;
    size_t step = (size_t(sz) + size_t(threads) - 1) / size_t(threads);
    const char* b = base;
    while (b < end)
    {
        const char* e = (size_t(end - b) <= step) ? end : b + step;
        while (e < end && *e != '\n') ++e;
        if (e < end) ++e;  // slice ends just after its newline
        cuts.emplace_back(b, e);
        b = e;
    }
    std::vector<ChunkOut> parts(cuts.size());
    if (cuts.size() <= 1)
    {
        if (!cuts.empty()) scan_lines(cuts[0].first, cuts[0].second, parts[0]);
    }
    else
    {
        std::atomic<size_t> next{0};
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t)
            pool.emplace_back([&]() {
                for (size_t i; (i = next.fetch_add(1)) < cuts.size();) scan_lines(cuts[i].first, cuts[i].second, parts[i]);
            });
        for (auto& th : pool) th.join();
    }
    m = HostModel();
    for (auto& c : parts)
    {
        m.vertices.insert(m.vertices.end(), c.v.begin(), c.v.end());
        m.normals.insert(m.normals.end(), c.vn.begin(), c.vn.end());
        m.tex_coords.insert(m.tex_coords.end(), c.vt.begin(), c.vt.end());
        m.faces.insert(m.faces.end(), c.f.begin(), c.f.end());
    }
    // prep_model_data (OBJ_loader.cpp:229-267): negative = relative to the END of the finished arrays; then 1-based → 0-based
    // (an absent vt/vn becomes -1).
    const int32_t nv = int32_t(m.vertices.size()), nn = int32_t(m.normals.size()), nt = int32_t(m.tex_coords.size());
    for (Face& f : m.faces)
        for (int k = 0; k < 3; ++k)
        {
            if (f.vt[k] < 0) f.vt[k] += nt + 1;
            if (f.vn[k] < 0) f.vn[k] += nn + 1;
            if (f.v[k] < 0) f.v[k] += nv + 1;
            --f.vt[k]; --f.vn[k]; --f.v[k];
        }
    for (const Face& f : m.faces)
        for (int k = 0; k < 3; ++k)
            if (f.v[k] < 0 || f.v[k] >= nv) { err = "face references a vertex outside the file"; return -1; }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// model box and transforms (model.h). min/max are the reference's `a > b ? … : …` forms (PL_math.h:330-338).
// ---------------------------------------------------------------------------------------------------------------
static inline float pick_max(float a, float b) { return a > b ? a : b; }
static inline float pick_min(float a, float b) { return a > b ? b : a; }

static Box padded_bounds(const std::vector<Vec3>& pts)  // get_AABB, model.h:41-61
{
    float hx = -kMaxFloat, lx = kMaxFloat, hy = -kMaxFloat, ly = kMaxFloat, hz = -kMaxFloat, lz = kMaxFloat;
    for (const Vec3& p : pts)
    {
        hx = pick_max(hx, p.x); lx = pick_min(lx, p.x);
        hy = pick_max(hy, p.y); ly = pick_min(ly, p.y);
        hz = pick_max(hz, p.z); lz = pick_min(lz, p.z);
    }
    Box b;
    b.hi = {hx + kTolerance, hy + kTolerance, hz + kTolerance};
    b.lo = {lx - kTolerance, ly - kTolerance, lz - kTolerance};
    return b;
}

void compute_aabb(HostModel& m)
{
    m.aabb = padded_bounds(m.vertices);
    m.has_aabb = true;
}

void resize_scale(HostModel& m, float new_max_scale)  // model.h:108-121
{
    float ex = m.aabb.hi.x - m.aabb.lo.x, ey = m.aabb.hi.y - m.aabb.lo.y, ez = m.aabb.hi.z - m.aabb.lo.z;
    float k = new_max_scale / pick_max(pick_max(ex, ey), ez);
    for (Vec3& p : m.vertices) { p.x *= k; p.y *= k; p.z *= k; }
    m.aabb.hi = {m.aabb.hi.x * k, m.aabb.hi.y * k, m.aabb.hi.z * k};
    m.aabb.lo = {m.aabb.lo.x * k, m.aabb.lo.y * k, m.aabb.lo.z * k};
}

void translate_to(HostModel& m, Vec3 c)  // model.h:136-152
{
    Vec3 mid = {m.aabb.lo.x + (m.aabb.hi.x - m.aabb.lo.x) / 2.0f, m.aabb.lo.y + (m.aabb.hi.y - m.aabb.lo.y) / 2.0f,
                m.aabb.lo.z + (m.aabb.hi.z - m.aabb.lo.z) / 2.0f};
    Vec3 d = {c.x - mid.x, c.y - mid.y, c.z - mid.z};
    for (Vec3& p : m.vertices) { p.x += d.x; p.y += d.y; p.z += d.z; }
    m.aabb.hi = {m.aabb.hi.x + d.x, m.aabb.hi.y + d.y, m.aabb.hi.z + d.z};
    m.aabb.lo = {m.aabb.lo.x + d.x, m.aabb.lo.y + d.y, m.aabb.lo.z + d.z};
}

// ---------------------------------------------------------------------------------------------------------------
// tree build (kd_tree.cpp:20-288). Semantics that must be reproduced for the node array to match the reference:
//  * a node splits iff it holds MORE than `cap` primitives; split point = box midpoint (CENTER) or the
//    area-weighted mean of triangle centres (SAH; f32 vector sum, f64 area sum); a SAH point outside the node box
//    aborts the split and the node stays an over-full leaf (kd_tree.cpp:107-112);
//  * child c covers the low/high half per axis with bit2 = x, bit1 = y, bit0 = z (kd_tree.cpp:120-148);
//  * a primitive goes to every child whose box contains at least one of its VERTICES, bounds inclusive
//    (kd_tree.cpp:10-17, aabb.h:19-27) — straddling triangles without a vertex inside are dropped there;
//  * the 8 children are appended contiguously when the parent is split, and nodes are processed LIFO, so the last
//    child is split first (kd_tree.cpp:259-280).
// ---------------------------------------------------------------------------------------------------------------
static inline bool inside(const Vec3& p, const Box& b)
{
    return p.x >= b.lo.x && p.x <= b.hi.x && p.y >= b.lo.y && p.y <= b.hi.y && p.z >= b.lo.z && p.z <= b.hi.z;
}

static inline float half_cross_length(const PrimRef& t)  // area_of_triangle, kd_tree.cpp:3-8: |(a-c) x (a-b)| / 2
{
    Vec3 ab = {t.a.x - t.b.x, t.a.y - t.b.y, t.a.z - t.b.z};
    Vec3 ac = {t.a.x - t.c.x, t.a.y - t.c.y, t.a.z - t.c.z};
    Vec3 n = {ac.y * ab.z - ac.z * ab.y, ac.z * ab.x - ac.x * ab.z, ac.x * ab.y - ac.y * ab.x};
    return std::sqrt((n.x * n.x) + (n.y * n.y) + (n.z * n.z)) / 2;
}

static bool split_point(const TreeNode& n, int method, Vec3& s)
{
    if (method == 0)
    {
        s = {(n.box.hi.x - n.box.lo.x) / 2.0f, (n.box.hi.y - n.box.lo.y) / 2.0f, (n.box.hi.z - n.box.lo.z) / 2.0f};
        s = {s.x + n.box.lo.x, s.y + n.box.lo.y, s.z + n.box.lo.z};
        return true;
    }
    Vec3 acc = {0.f, 0.f, 0.f};
    double area_total = 0.0;
    for (const PrimRef& t : n.prims)
    {
        Vec3 ctr = {((t.a.x + t.b.x) + t.c.x) / 3.0f, ((t.a.y + t.b.y) + t.c.y) / 3.0f, ((t.a.z + t.b.z) + t.c.z) / 3.0f};
        float area = half_cross_length(t);
        acc = {acc.x + ctr.x * area, acc.y + ctr.y * area, acc.z + ctr.z * area};
        area_total += area;
    }
    float w = float(area_total);
    s = {acc.x / w, acc.y / w, acc.z / w};
    return inside(s, n.box);
}

// One step of the builder (kd_tree.cpp:67-288): splits L[id] into 8 children appended at the end of L when it holds more than
// `cap` primitives and its split point lies inside its box. Returns false when the node stays a leaf.
static bool split_node(std::vector<TreeNode>& L, uint32_t id, int method, uint32_t cap, size_t& live_refs)
{
    if (L[id].prims.size() <= cap) return false;
    Vec3 s;
    if (!split_point(L[id], method, s)) return false;
    const Box pb = L[id].box;
    std::vector<PrimRef> parent;
    parent.swap(L[id].prims);
    const uint32_t first = uint32_t(L.size());
    L[id].children_start = int32_t(first);
    L.resize(L.size() + 8);
    for (int c = 0; c < 8; ++c)
    {
        Box& cb = L[first + c].box;
        cb.lo.x = (c & 4) ? s.x : pb.lo.x; cb.hi.x = (c & 4) ? pb.hi.x : s.x;
        cb.lo.y = (c & 2) ? s.y : pb.lo.y; cb.hi.y = (c & 2) ? pb.hi.y : s.y;
        cb.lo.z = (c & 1) ? s.z : pb.lo.z; cb.hi.z = (c & 1) ? pb.hi.z : s.z;
    }
    live_refs -= parent.size();
    for (const PrimRef& t : parent)
        for (int c = 0; c < 8; ++c)
        {
            const Box& cb = L[first + c].box;
            if (inside(t.a, cb) || inside(t.b, cb) || inside(t.c, cb)) { L[first + c].prims.push_back(t); ++live_refs; }
        }
    return true;
}

static const char* kGuardMessage = "tree build exceeded the node/primitive guard (degenerate split; the reference would not terminate)";

// The reference's order: a LIFO work list, the 8 children of a node appended when the node is taken — a pre-order walk that visits
// the children 7, 6, … 0, every subtree complete before the next. L[0] is the subtree's root (with its primitives).
static int build_sequential(std::vector<TreeNode>& L, int method, uint32_t cap, std::string& err)
{
    std::vector<uint32_t> todo{0u};
    size_t live_refs = L[0].prims.size();
    while (!todo.empty())
    {
        const uint32_t id = todo.back();
        todo.pop_back();
        if (!split_node(L, id, method, cap, live_refs)) continue;
        const uint32_t first = uint32_t(L[id].children_start);
        for (uint32_t c = 0; c < 8; ++c) todo.push_back(first + c);
        if (live_refs > 400000000u || L.size() > 40000000u) { err = kGuardMessage; return -5; }
    }
    return 0;
}

// SURVEY §8 f4 (first half): the same tree, built by several threads. Because every subtree is complete before the next one starts,
// the nodes below child c of a split node are one contiguous run of the reference's array, and the runs follow each other in the
// order 7, 6, … 0 behind the node's 8 children. So the 8 subtrees are built independently (each in its own array with local
// numbering, `depth` levels of fan-out) and spliced behind one another with their child indices rebased: bit-identical node array
// (tests/test_host_model.py compares it with the reference build's).
static int build_parallel(std::vector<TreeNode>& L, int method, uint32_t cap, int depth, std::string& err)
{
    if (depth <= 0) return build_sequential(L, method, cap, err);
    size_t live_refs = L[0].prims.size();
    if (!split_node(L, 0, method, cap, live_refs)) return 0;   // L = {root, 8 children}
    std::vector<TreeNode> sub[8];
    int rc[8];
    std::string errs[8];
    std::vector<std::thread> th;
    for (int c = 0; c < 8; ++c)
    {
        sub[c].resize(1);
        sub[c][0].box = L[1 + c].box;
        sub[c][0].prims.swap(L[1 + c].prims);
        rc[c] = 0;
    }
    for (int c = 1; c < 8; ++c)
        th.emplace_back([&, c] { rc[c] = build_parallel(sub[c], method, cap, depth - 1, errs[c]); });
    rc[0] = build_parallel(sub[0], method, cap, depth - 1, errs[0]);
    for (std::thread& t : th) t.join();
    for (int c = 0; c < 8; ++c)
        if (rc[c] != 0) { err = errs[c]; return rc[c]; }
    size_t total = L.size();
    for (int c = 0; c < 8; ++c) total += sub[c].size() - 1;
    if (total > 40000000u) { err = kGuardMessage; return -5; }
    L.reserve(total);
    for (int c = 7; c >= 0; --c)
    {
        const int32_t base = int32_t(L.size());   // local index i >= 1 of this subtree lands at base + i - 1
        TreeNode& root = L[1 + c];
        root.prims.swap(sub[c][0].prims);
        root.children_start = sub[c][0].children_start > 0 ? base + sub[c][0].children_start - 1 : 0;
        for (size_t i = 1; i < sub[c].size(); ++i)
        {
            L.emplace_back();
            TreeNode& n = L.back();
            n.box = sub[c][i].box;
            n.children_start = sub[c][i].children_start > 0 ? base + sub[c][i].children_start - 1 : 0;
            n.prims.swap(sub[c][i].prims);
        }
        std::vector<TreeNode>().swap(sub[c]);
    }
    return 0;
}

int build_tree(HostModel& m, int method, uint32_t cap, std::string& err, int threads)
{
    m.tree.clear();
    m.max_faces_per_node = cap;
    m.division_method = method;
    m.tree.emplace_back();
    m.tree[0].box = padded_bounds(m.vertices);  // build_KD_tree recomputes get_AABB from the current vertices (kd_tree.cpp:23)
    m.tree[0].prims.resize(m.faces.size());
    for (size_t i = 0; i < m.faces.size(); ++i)
    {
        const Face& f = m.faces[i];
        m.tree[0].prims[i] = {m.vertices[f.v[0]], m.vertices[f.v[1]], m.vertices[f.v[2]], uint32_t(i)};
    }
    // one level of fan-out (8 subtrees) from a few thousand faces on, two (64) for large meshes; small ones are not worth a thread
    int depth = 0;
    if (threads > 1 && m.faces.size() >= 4096) depth = (threads > 8 && m.faces.size() >= 500000) ? 2 : 1;
    return build_parallel(m.tree, method, cap, depth, err);
}

// ---------------------------------------------------------------------------------------------------------------
// flat GPU form. Edges ab = b - a, ac = c - a are the first two operations of the reference's triangle
// test (model.h:78-79); computing them here in f32 yields the same bits the CPU path computes per ray.
// Within a leaf the records are stored in Morton order of their centroids (spatially compact packets of 8 for the
// film-space culling of the render kernel); each record carries its position in the reference's leaf order, which is
// what the sequential strict-< scan of kd_tree.cpp:437-462 breaks ties on.
// ---------------------------------------------------------------------------------------------------------------
static inline float bits_to_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

static inline uint32_t spread10(uint32_t v)  // 10 bits → every third bit
{
    v &= 1023u;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

static void morton_order(const std::vector<PrimRef>& prims, std::vector<uint32_t>& order)
{
    const size_t n = prims.size();
    order.resize(n);
    for (size_t i = 0; i < n; ++i) order[i] = uint32_t(i);
    if (n <= 8) return;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    std::vector<double> ctr(n * 3);
    for (size_t i = 0; i < n; ++i)
    {
        const PrimRef& t = prims[i];
        const double c[3] = {(double(t.a.x) + t.b.x + t.c.x) / 3.0, (double(t.a.y) + t.b.y + t.c.y) / 3.0,
                             (double(t.a.z) + t.b.z + t.c.z) / 3.0};
        for (int k = 0; k < 3; ++k)
        {
            ctr[3 * i + k] = c[k];
            if (c[k] < lo[k]) lo[k] = c[k];
            if (c[k] > hi[k]) hi[k] = c[k];
        }
    }
    std::vector<uint32_t> code(n, 0u);
    for (size_t i = 0; i < n; ++i)
    {
        uint32_t q[3];
        for (int k = 0; k < 3; ++k)
        {
            const double e = hi[k] - lo[k];
            double f = e > 0.0 ? (ctr[3 * i + k] - lo[k]) / e : 0.0;
            if (!(f >= 0.0)) f = 0.0;   // NaN-safe
            if (f > 1.0) f = 1.0;
            q[k] = uint32_t(f * 1023.0);
        }
        code[i] = (spread10(q[0]) << 2) | (spread10(q[1]) << 1) | spread10(q[2]);
    }
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return code[x] < code[y]; });
}

void flatten(const HostModel& m, FlatScene& out)
{
    out = FlatScene();
    out.n_nodes = uint32_t(m.tree.size());
    out.nodes.assign(size_t(out.n_nodes) * 8, 0.f);
    uint32_t slots = 0;
    for (const TreeNode& n : m.tree)
        if (n.children_start == 0) slots += (uint32_t(n.prims.size()) + 7u) & ~7u;
    out.n_slots = slots;
    out.tris.assign(size_t(slots / 8) * 24 * 4, 0.f);
    out.packet_leaf.assign(slots / 8, 0u);
    out.pos_face.assign(slots, 0xffffffffu);
    // Leaf slots are laid out in depth-first order of the tree (children in index order), not in node order: the packets
    // of every subtree are then one contiguous range node_pk[node] = [begin, end), and the per-frame pixel range of an
    // inner node is a union over that range (render.cu: k_view_nodes) instead of a bottom-up pass over the levels.
    std::vector<uint32_t> first_slot(out.n_nodes, 0u);
    out.node_pk.assign(size_t(out.n_nodes) * 2, 0u);
    {
        uint32_t at = 0;
        std::vector<std::pair<uint32_t, int>> st;   // (node, next child to descend into; -1 = entering)
        if (out.n_nodes) st.push_back({0u, -1});
        while (!st.empty())
        {
            auto& top = st.back();
            const uint32_t nd = top.first;
            const TreeNode& n = m.tree[nd];
            if (top.second < 0)
            {
                out.node_pk[2 * size_t(nd)] = at / 8;
                if (n.children_start == 0)
                {
                    first_slot[nd] = at;
                    at += (uint32_t(n.prims.size()) + 7u) & ~7u;
                    out.node_pk[2 * size_t(nd) + 1] = at / 8;
                    st.pop_back();
                    continue;
                }
                top.second = 0;
            }
            if (top.second < 8)
            {
                const uint32_t ch = uint32_t(n.children_start) + uint32_t(top.second++);
                st.push_back({ch, -1});
            }
            else
            {
                out.node_pk[2 * size_t(nd) + 1] = at / 8;
                st.pop_back();
            }
        }
    }
    std::vector<uint32_t> order;
    for (uint32_t i = 0; i < out.n_nodes; ++i)
    {
        const TreeNode& n = m.tree[i];
        const uint32_t cursor = first_slot[i];
        float* q = &out.nodes[size_t(i) * 8];
        q[0] = n.box.lo.x; q[1] = n.box.lo.y; q[2] = n.box.lo.z;
        q[4] = n.box.hi.x; q[5] = n.box.hi.y; q[6] = n.box.hi.z;
        if (n.children_start != 0)
        {
            q[3] = bits_to_float(uint32_t(n.children_start));
            q[7] = bits_to_float(0u);
            continue;
        }
        const uint32_t cnt = uint32_t(n.prims.size());
        q[3] = bits_to_float(~cursor);
        q[7] = bits_to_float(cnt);
        ++out.n_leaves;
        if (cnt > out.max_leaf_prims) out.max_leaf_prims = cnt;
        const uint32_t padded = (cnt + 7u) & ~7u;
        morton_order(n.prims, order);
        for (uint32_t k = 0; k < cnt; ++k) out.pos_face[cursor + k] = n.prims[k].face;
        for (uint32_t k = 0; k < padded; ++k)
        {
            const uint32_t slot = cursor + k;
            float* pkt = &out.tris[size_t(slot / 8) * 24 * 4];
            float* A = pkt + size_t(slot % 8) * 4;
            float* AB = A + 8 * 4;
            float* AC = A + 16 * 4;
            out.packet_leaf[slot / 8] = i;
            if (k < cnt)
            {
                const PrimRef& t = n.prims[order[k]];
                A[0] = t.a.x; A[1] = t.a.y; A[2] = t.a.z; A[3] = bits_to_float(t.face);
                AB[0] = t.b.x - t.a.x; AB[1] = t.b.y - t.a.y; AB[2] = t.b.z - t.a.z; AB[3] = bits_to_float(order[k]);
                AC[0] = t.c.x - t.a.x; AC[1] = t.c.y - t.a.y; AC[2] = t.c.z - t.a.z;
            }
            else
            {
                A[3] = bits_to_float(0xffffffffu);  // padding slot: zero edges → det = 0 < tolerance → never a hit
                AB[3] = bits_to_float(0xffffffffu);
            }
        }
    }
    // depth of every node (children always follow their parent in the array), inner nodes bucketed deepest first
    std::vector<uint32_t> depth(out.n_nodes, 0u);
    uint32_t max_depth = 0;
    for (uint32_t i = 0; i < out.n_nodes; ++i)
        if (m.tree[i].children_start != 0)
        {
            for (uint32_t c = 0; c < 8; ++c) depth[uint32_t(m.tree[i].children_start) + c] = depth[i] + 1;
            if (depth[i] > max_depth) max_depth = depth[i];
        }
    // static insertion order of the leaves and the parent links (rasteriser tables)
    out.leaf_seq.assign(out.n_nodes, 0u);
    out.parent.assign(out.n_nodes, 0xffffffffu);
    if (out.n_nodes && m.tree[0].children_start != 0)
    {
        std::vector<uint32_t> stack(1, 0u);
        while (!stack.empty())
        {
            const uint32_t cur = stack.back();
            stack.pop_back();
            const uint32_t c0 = uint32_t(m.tree[cur].children_start);
            for (uint32_t c = 0; c < 8; ++c)
            {
                const uint32_t ch = c0 + c;
                out.parent[ch] = cur;
                if (m.tree[ch].children_start != 0) stack.push_back(ch);
                else
                {
                    out.leaf_seq[ch] = uint32_t(out.seq_leaf.size());
                    out.seq_leaf.push_back(ch);
                }
            }
        }
    }
    out.level_start.assign(1, 0u);
    for (int d = int(max_depth); d >= 0; --d)
    {
        for (uint32_t i = 0; i < out.n_nodes; ++i)
            if (m.tree[i].children_start != 0 && depth[i] == uint32_t(d)) out.inner_order.push_back(i);
        out.level_start.push_back(uint32_t(out.inner_order.size()));
    }
}

// camera basis (camera.h:23-45); normalize = v * (1/sqrt(|v|²)) with |v|² = (x² + y²) + z² (PL_math.h:377-392)
static inline Vec3 unit(Vec3 v)
{
    float k = 1.0f / std::sqrt((v.x * v.x) + (v.y * v.y) + (v.z * v.z));
    return {v.x * k, v.y * k, v.z * k};
}
static inline Vec3 cross3(Vec3 a, Vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

CameraBasis make_camera(int w, int h, Vec3 eye, Vec3 facing, float h_fov)
{
    CameraBasis c;
    c.h_fov = h_fov;
    c.aspect = w / float(h);
    c.eye = eye;
    Vec3 f = unit(facing);
    c.fc = {eye.x + f.x, eye.y + f.y, eye.z + f.z};
    c.cz = {-f.x, -f.y, -f.z};
    c.cx = unit(cross3({0.f, 1.f, 0.f}, c.cz));
    c.cy = unit(cross3(c.cz, c.cx));
    c.hpw = (0.5f * h_fov) / float(w);
    c.hph = 0.5f / float(h);
    return c;
}

}  // namespace atray
