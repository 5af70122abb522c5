// sm_100a kernels of the ATRay render path: camera ray generation → 8-way tree traversal → culled Möller–Trumbore →
// flat shade → {t, primitive id, BGRX} store. Replaces the reference's tile thread pool and everything below it:
//   render_tile_from_camera  engine/renderer/renderer.cpp:294-369   (pixel loop, film coordinates, SetRay, pack)
//   cast_ray @ bounce 1      renderer.cpp:213-262                    (sky / model emission, 3 RNG draws per hit)
//   get_intersection_data    renderer.cpp:34-57                      (Optimized_Ray, accept t in (1e-4, MAX_FLOAT))
//   traverse_oct_tree_new    engine/renderer/kd_tree.cpp:337-465     (two-phase traversal)
//   ray/box slab tests       engine/renderer/aabb.h:29-93
//   triangle test            engine/renderer/model.h:75-103
//
// Execution model: persistent CTAs; each warp pulls 8×4-pixel tiles (or 32-ray groups) from an atomic
// counter. Phase 1 runs one ray per lane: LIFO walk over inner nodes with the ≤5-accepted-children cut, leaves
// collected into a per-lane list kept sorted by entry distance (stable insertion). Phase 2 is warp-cooperative: the
// warp takes its rays one at a time, the 32 lanes test 32 consecutive triangle slots of the current leaf with
// coalesced float4 loads of precomputed-edge records, a lexicographic (t, slot) warp arg-min reproduces the
// reference's sequential strict-< scan, and the ray stops after the first leaf that yields a hit.
//
// Arithmetic: in STRICT mode every decision-critical float op is an explicit round-to-nearest intrinsic
// (__fmul_rn/__fadd_rn/__fsub_rn never contract into FMA; __frcp_rn / __fdiv_rn / __fsqrt_rn are IEEE) in the
// reference's operation order → bit-identical t and ids. FAST mode lets the triangle test contract into FMA.
#include "device_types.h"

#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace atray {

#define FULL_MASK 0xffffffffu
static constexpr float kTol = 0.0001f;               // ray.h:5
static constexpr float kMaxFloat = 3.402823466e+38F;  // PL_base_defs.h:70
#ifndef ATRAY_LEAF_CAP_LOCAL
#define ATRAY_LEAF_CAP_LOCAL 48
#endif
static constexpr int kLeafCap = ATRAY_LEAF_CAP_LOCAL;                   // per-lane leaf list (local memory); overflow → per-warp spill area
static constexpr int kStackCap = ATRAY_LEAF_CAP_LOCAL;                  // per-lane inner-node stack
static constexpr int kBlock = 256;                    // 8 warps per CTA
#ifndef ATRAY_RENDER_MIN_BLOCKS
#define ATRAY_RENDER_MIN_BLOCKS 3                     // resident CTAs per SM the culled render kernel is compiled for (register cap)
#endif

struct RayL { float ox, oy, oz, dx, dy, dz; };

// Cancel word of the context (DevScene::cancel, device memory): the host raises it from a second stream while a kernel runs
// (atray_render_cancel — the reference's WorkQueue::clear() + "threads finish the tile they are on", app.cpp:165-172). Every
// persistent loop reads it next to its work cursor and stops taking work once it is set.
__device__ __forceinline__ bool cancel_requested(const DevScene& sc)
{
    return *reinterpret_cast<const volatile uint32_t*>(sc.cancel) != 0u;
}

__device__ __forceinline__ float mulr(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float addr(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float subr(float a, float b) { return __fsub_rn(a, b); }

template <bool STRICT> struct Fp
{
    static __device__ __forceinline__ float mul(float a, float b) { return STRICT ? __fmul_rn(a, b) : a * b; }
    static __device__ __forceinline__ float add(float a, float b) { return STRICT ? __fadd_rn(a, b) : a + b; }
    static __device__ __forceinline__ float sub(float a, float b) { return STRICT ? __fsub_rn(a, b) : a - b; }
    // dot = (ax*bx + ay*by) + az*bz ; cross component = p*q - r*s  (PL_math.h:416, 422)
    static __device__ __forceinline__ float dot(float ax, float ay, float az, float bx, float by, float bz)
    {
        return add(add(mul(ax, bx), mul(ay, by)), mul(az, bz));
    }
    static __device__ __forceinline__ float crs(float p, float q, float r, float s) { return sub(mul(p, q), mul(r, s)); }
};

struct WorkCount
{
    unsigned int box, tri, pdet, pu, pv, listed, scanned, ovf, rect, tsky, tbusy;
    __device__ void clear() { box = tri = pdet = pu = pv = listed = scanned = ovf = rect = tsky = tbusy = 0; }
};

// ---- ray/box slab tests (aabb.h). `near`/`far` selection by the sign of 1/d is the reference's bounds[sign] indexing.
struct BoxT { float tmin, tmax; bool ok; };

__device__ __forceinline__ BoxT slab_xy_z(const RayL& r, float ix, float iy, float iz, float4 lo, float4 hi, bool full)
{
    BoxT o;
    const bool sx = ix < 0.0f, sy = iy < 0.0f, sz = iz < 0.0f;
    float tmin = mulr(subr(sx ? hi.x : lo.x, r.ox), ix);
    float tmax = mulr(subr(sx ? lo.x : hi.x, r.ox), ix);
    const float tymin = mulr(subr(sy ? hi.y : lo.y, r.oy), iy);
    const float tymax = mulr(subr(sy ? lo.y : hi.y, r.oy), iy);
    o.ok = !((tmin > tymax) || (tymin > tmax));
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    const float tzmin = mulr(subr(sz ? hi.z : lo.z, r.oz), iz);
    const float tzmax = mulr(subr(sz ? lo.z : hi.z, r.oz), iz);
    o.ok = o.ok && !((tmin > tzmax) || (tzmin > tmax));
    if (full)
    {
        if (tzmin > tmin) tmin = tzmin;
        if (tzmax < tmax) tmax = tzmax;
    }
    o.tmin = tmin;
    o.tmax = tmax;
    return o;
}
// check_ray_AABB_intersection, aabb.h:65-93 (no sign test: boxes behind the ray pass)
__device__ __forceinline__ bool box_check(const RayL& r, float ix, float iy, float iz, float4 lo, float4 hi)
{
    return slab_xy_z(r, ix, iy, iz, lo, hi, false).ok;
}
// get_ray_AABB_intersection, aabb.h:29-63: tmin if > 0, else tmax if > 0, else 0
__device__ __forceinline__ float box_entry(const RayL& r, float ix, float iy, float iz, float4 lo, float4 hi)
{
    BoxT b = slab_xy_z(r, ix, iy, iz, lo, hi, true);
    if (!b.ok) return 0.0f;
    if (b.tmin > 0.0f) return b.tmin;
    if (b.tmax > 0.0f) return b.tmax;
    return 0.0f;
}

// ---- phase 1: collect the sorted leaf list of one ray (kd_tree.cpp:363-435). Returns false on capacity overflow.
// The reference's list is ascending in the entry distance with ties in insertion order; the insertion order of a ray is
// the static full-tree order restricted to the leaves it lists (FlatScene::leaf_seq), so the list is "ascending
// (distance, leaf_seq)" and does not depend on the order in which subtrees are visited. The children of a node are still
// examined in index order (the "at most 5 accepted" cut counts in that order), but the accepted inner children are
// pushed far-to-near for this ray's direction signs (child bit 4/2/1 = upper x/y/z half, kd_tree.cpp:120-148): subtrees
// are then visited front to back, entry distances arrive almost ascending and the insertion rarely has to shift.
template <bool COUNT>
__device__ bool collect_leaves(const float4* __restrict__ nodes, const uint32_t* __restrict__ leaf_seq, const RayL& r, float ix, float iy,
                               float iz, uint32_t* stack, int stack_cap, uint2* leaves, int leaf_cap, int& nleaf, WorkCount& wc)
{
    nleaf = 0;
    int sp = 1;
    stack[0] = 0u;
    const unsigned near_mask = (ix < 0.0f ? 4u : 0u) | (iy < 0.0f ? 2u : 0u) | (iz < 0.0f ? 1u : 0u);
    while (sp > 0)
    {
        const uint32_t cur = stack[--sp];
        const uint32_t c0 = (uint32_t)__float_as_int(__ldg(&nodes[2 * cur]).w);
        int accepted = 0;
        unsigned inner_acc = 0u;
#pragma unroll 1
        for (int i = 0; i < 8 && accepted <= 4; ++i)  // at most 5 accepted children (kd_tree.cpp:374)
        {
            const uint32_t ch = c0 + i;
            const float4 lo = __ldg(&nodes[2 * ch]);
            const float4 hi = __ldg(&nodes[2 * ch + 1]);
            if (COUNT) wc.box++;
            if (__float_as_int(lo.w) > 0)  // inner child
            {
                if (box_check(r, ix, iy, iz, lo, hi))
                {
                    ++accepted;
                    inner_acc |= 1u << i;
                }
            }
            else  // leaf child (possibly empty)
            {
                const float dis = box_entry(r, ix, iy, iz, lo, hi);
                if (dis > 0.0f)
                {
                    ++accepted;
                    if (nleaf >= leaf_cap) return false;
                    int k = nleaf;
                    while (k > 0)
                    {
                        const float pd = __uint_as_float(leaves[k - 1].y);
                        if (!(dis < pd || (dis == pd && __ldg(&leaf_seq[ch]) < __ldg(&leaf_seq[leaves[k - 1].x])))) break;
                        leaves[k] = leaves[k - 1];
                        --k;
                    }
                    leaves[k] = make_uint2(ch, __float_as_uint(dis));
                    ++nleaf;
                    if (COUNT) wc.listed++;
                }
            }
        }
        if (inner_acc)
        {
            if (sp + __popc(inner_acc) > stack_cap) return false;
#pragma unroll
            for (int j = 7; j >= 0; --j)  // farthest octant first → the nearest is popped next
            {
                const unsigned i = (unsigned)j ^ near_mask;
                if ((inner_acc >> i) & 1u) stack[sp++] = c0 + i;
            }
        }
    }
    return true;
}

// ---- phase 2: warp-cooperative scan of ONE ray's leaf list (kd_tree.cpp:437-462). All 32 lanes call this with the
// same `src` lane; the ray and the list entries are broadcast from that lane. Lane `src` receives the result.
template <bool STRICT, bool COUNT, bool UV = false>
__device__ __forceinline__ void scan_ray_leaves(const DevScene& sc, int lane, int src, const RayL& mine, const uint2* my_list,
                                                int my_nleaf, float& out_t, uint32_t& out_face, WorkCount& wc,
                                                float* out_u = nullptr, float* out_v = nullptr)
{
    typedef Fp<STRICT> F;
    const float ox = __shfl_sync(FULL_MASK, mine.ox, src), oy = __shfl_sync(FULL_MASK, mine.oy, src),
                oz = __shfl_sync(FULL_MASK, mine.oz, src);
    const float dx = __shfl_sync(FULL_MASK, mine.dx, src), dy = __shfl_sync(FULL_MASK, mine.dy, src),
                dz = __shfl_sync(FULL_MASK, mine.dz, src);
    const int n = __shfl_sync(FULL_MASK, my_nleaf, src);
    for (int k = 0; k < n; ++k)
    {
        uint32_t node = 0;
        if (lane == src) node = my_list[k].x;
        node = __shfl_sync(FULL_MASK, node, src);
        const uint32_t first = ~(uint32_t)__float_as_int(__ldg(&sc.nodes[2 * node]).w);
        const uint32_t cnt = (uint32_t)__float_as_int(__ldg(&sc.nodes[2 * node + 1]).w);
        if (COUNT && lane == src) wc.scanned++;
        float bt = kMaxFloat;
        float bu = 0.0f, bv = 0.0f;  // barycentrics of the best hit (TriangleIntersectionData.u/v), UV only
        uint32_t bpos = 0xffffffffu, bface = 0xffffffffu;
        const uint32_t padded = (cnt + 7u) & ~7u;  // slots of this leaf (packets of 8)
        for (uint32_t base = 0; base < cnt; base += 32)
        {
            // 32 consecutive slots = 4 packets; lane → packet (lane >> 3), slot in packet (lane & 7)
            const uint32_t slot = first + base + lane;
            const bool in_leaf = base + lane < padded;
            const float4* rec = sc.tris + (size_t)(slot >> 3) * 24 + (slot & 7);
            float4 A = make_float4(0.f, 0.f, 0.f, __uint_as_float(0xffffffffu)), AB = make_float4(0.f, 0.f, 0.f, 0.f), AC = AB;
            if (in_leaf) { A = __ldg(rec); AB = __ldg(rec + 8); AC = __ldg(rec + 16); }
            const bool real = __float_as_uint(A.w) != 0xffffffffu;
            // get_triangle_ray_intersection_culled, model.h:75-103 (ab, ac precomputed with the same f32 subtraction)
            const float px = F::crs(dy, AC.z, dz, AC.y), py = F::crs(dz, AC.x, dx, AC.z), pz = F::crs(dx, AC.y, dy, AC.x);
            const float det = F::dot(AB.x, AB.y, AB.z, px, py, pz);
            bool alive = !(det < kTol);
            if (COUNT) { wc.tri += real ? 1u : 0u; wc.pdet += (real && alive) ? 1u : 0u; }
            if (!__any_sync(FULL_MASK, alive)) continue;
            const float det_inv = __frcp_rn(det);  // 1 / det, IEEE round-to-nearest
            const float tx = F::sub(ox, A.x), ty = F::sub(oy, A.y), tz = F::sub(oz, A.z);
            const float u = F::mul(F::dot(tx, ty, tz, px, py, pz), det_inv);
            alive = alive && !(u < 0.0f || u > 1.0f);
            if (COUNT) wc.pu += (real && alive) ? 1u : 0u;
            if (!__any_sync(FULL_MASK, alive)) continue;
            const float qx = F::crs(ty, AB.z, tz, AB.y), qy = F::crs(tz, AB.x, tx, AB.z), qz = F::crs(tx, AB.y, ty, AB.x);
            const float v = F::mul(F::dot(dx, dy, dz, qx, qy, qz), det_inv);
            alive = alive && !(v < 0.0f || F::add(u, v) > 1.0f);
            if (COUNT) wc.pv += (real && alive) ? 1u : 0u;
            if (!__any_sync(FULL_MASK, alive)) continue;
            const float t = F::mul(F::dot(qx, qy, qz, AC.x, AC.y, AC.z), det_inv);
            // caller's accept: distance < closest && distance > tolerance, scanning in the reference's leaf order: among equal
            // distances the lower position wins (the records are Morton-sorted, so a lane does not meet them in that order)
            if (alive && t > kTol && (t < bt || (t == bt && __float_as_uint(AB.w) < bpos)))
            {
                bt = t;
                bpos = __float_as_uint(AB.w);  // position in the reference's leaf order
                bface = __float_as_uint(A.w);
                if (UV) { bu = u; bv = v; }
            }
        }
        if (!__any_sync(FULL_MASK, bt < kMaxFloat)) continue;  // no triangle of this leaf accepted the ray: next leaf
        // sequential strict-< scan ≡ lexicographic arg-min over (t, slot)
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
        {
            const float ot = __shfl_xor_sync(FULL_MASK, bt, off);
            const uint32_t op = __shfl_xor_sync(FULL_MASK, bpos, off);
            const uint32_t of = __shfl_xor_sync(FULL_MASK, bface, off);
            float ou = 0.0f, ov = 0.0f;
            if (UV) { ou = __shfl_xor_sync(FULL_MASK, bu, off); ov = __shfl_xor_sync(FULL_MASK, bv, off); }
            if (ot < bt || (ot == bt && op < bpos)) { bt = ot; bpos = op; bface = of; if (UV) { bu = ou; bv = ov; } }
        }
        if (bt < kMaxFloat)  // this leaf produced a hit → stop (kd_tree.cpp:457-460)
        {
            if (lane == src) { out_t = bt; out_face = bface; if (UV) { *out_u = bu; *out_v = bv; } }
            break;
        }
    }
}

// ---- the no-tree model query, renderer.cpp:58-82 (#ifndef USE_KD_TREE): if get_ray_AABB_intersection(model box) is
// non-zero, every face in file order through the triangle test, accepting t > tolerance && t < best (strict: the first
// face with the smallest t wins). One ray per lane; the 32 lanes walk the face table together (uniform loads).
template <bool STRICT, bool UV>
__device__ __forceinline__ void brute_warp(const DevScene& sc, bool active, const RayL& ray, float& t_out, uint32_t& face_out,
                                           float* u_out, float* v_out)
{
    typedef Fp<STRICT> F;
    t_out = kMaxFloat;
    face_out = 0xffffffffu;
    const float ix = __frcp_rn(ray.dx), iy = __frcp_rn(ray.dy), iz = __frcp_rn(ray.dz);
    const float4 lo = make_float4(sc.aabb_lo[0], sc.aabb_lo[1], sc.aabb_lo[2], 0.f), hi = make_float4(sc.aabb_hi[0], sc.aabb_hi[1], sc.aabb_hi[2], 0.f);
    const bool go = active && box_entry(ray, ix, iy, iz, lo, hi) != 0.0f;
    if (!__any_sync(FULL_MASK, go)) return;
    float best = kMaxFloat;
    for (uint32_t j = 0; j < sc.n_faces; ++j)
    {
        const float4 A = __ldg(&sc.faces[3 * (size_t)j]), AB = __ldg(&sc.faces[3 * (size_t)j + 1]), AC = __ldg(&sc.faces[3 * (size_t)j + 2]);
        const float px = F::crs(ray.dy, AC.z, ray.dz, AC.y), py = F::crs(ray.dz, AC.x, ray.dx, AC.z), pz = F::crs(ray.dx, AC.y, ray.dy, AC.x);
        const float det = F::dot(AB.x, AB.y, AB.z, px, py, pz);
        if (!go || det < kTol) continue;
        const float det_inv = __frcp_rn(det);
        const float tx = F::sub(ray.ox, A.x), ty = F::sub(ray.oy, A.y), tz = F::sub(ray.oz, A.z);
        const float u = F::mul(F::dot(tx, ty, tz, px, py, pz), det_inv);
        if (u < 0.0f || u > 1.0f) continue;
        const float qx = F::crs(ty, AB.z, tz, AB.y), qy = F::crs(tz, AB.x, tx, AB.z), qz = F::crs(tx, AB.y, ty, AB.x);
        const float v = F::mul(F::dot(ray.dx, ray.dy, ray.dz, qx, qy, qz), det_inv);
        if (v < 0.0f || F::add(u, v) > 1.0f) continue;
        const float t = F::mul(F::dot(qx, qy, qz, AC.x, AC.y, AC.z), det_inv);
        if (t > kTol && t < best)
        {
            best = t;
            face_out = j;
            if (UV) { *u_out = u; *v_out = v; }
        }
    }
    t_out = best;
}

// ---- one warp traces up to 32 rays (one per lane). Result: t (MAX_FLOAT on miss) and face (MISS) per lane.
template <bool STRICT, bool COUNT, bool UV = false>
__device__ __forceinline__ void trace_warp(const DevScene& sc, const DevScratch& scratch, int warp_slot, int lane, bool active,
                                           const RayL& ray, uint2* leaves, uint32_t* stack, float& t_out, uint32_t& face_out,
                                           WorkCount& wc, float* u_out = nullptr, float* v_out = nullptr)
{
    if (sc.brute)  // warp-uniform: the no-tree branch of get_intersection_data
    {
        brute_warp<STRICT, UV>(sc, active, ray, t_out, face_out, u_out, v_out);
        return;
    }
    t_out = kMaxFloat;
    face_out = 0xffffffffu;
    int nleaf = 0;
    bool overflow = false;
    // Optimized_Ray, renderer.cpp:41-44
    const float ix = __frcp_rn(ray.dx), iy = __frcp_rn(ray.dy), iz = __frcp_rn(ray.dz);
    if (active)
    {
        const float4 lo = __ldg(&sc.nodes[0]);
        const float4 hi = __ldg(&sc.nodes[1]);
        if (COUNT) wc.box++;
        if (box_check(ray, ix, iy, iz, lo, hi))  // kd_tree.cpp:339
        {
            if (sc.root_is_leaf)  // kd_tree.cpp:344-361: scan the root's primitives, no entry-distance filter
            {
                leaves[0] = make_uint2(0u, 0u);
                nleaf = 1;
            }
            else
                overflow = !collect_leaves<COUNT>(sc.nodes, sc.leaf_seq, ray, ix, iy, iz, stack, kStackCap, leaves, kLeafCap, nleaf, wc);
            if (overflow) nleaf = 0;
        }
    }
    unsigned todo = __ballot_sync(FULL_MASK, nleaf > 0);
    while (todo)
    {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        scan_ray_leaves<STRICT, COUNT, UV>(sc, lane, src, ray, leaves, nleaf, t_out, face_out, wc, u_out, v_out);
    }
    // rare: list or stack outgrew the per-lane arrays → redo that ray with the warp's spill area in global memory
    unsigned spill = __ballot_sync(FULL_MASK, overflow);
    while (spill)
    {
        const int src = __ffs(spill) - 1;
        spill &= spill - 1;
        uint2* gl = scratch.leaves + (size_t)warp_slot * scratch.leaf_cap;
        uint32_t* gs = scratch.stack + (size_t)warp_slot * scratch.stack_cap;
        int gn = 0;
        if (lane == src)
        {
            if (COUNT) wc.ovf++;
            WorkCount dummy;
            dummy.clear();
            // the aborted first attempt already counted a partial walk; the redo is not counted again
            collect_leaves<false>(sc.nodes, sc.leaf_seq, ray, ix, iy, iz, gs, (int)scratch.stack_cap, gl, (int)scratch.leaf_cap, gn, dummy);
        }
        __syncwarp();
        scan_ray_leaves<STRICT, COUNT, UV>(sc, lane, src, ray, gl, gn, t_out, face_out, wc, u_out, v_out);
        __syncwarp();
    }
}

// ---- camera ray, renderer.cpp:317,329,350-351 + SetRay ray.h:24-29 + normalize PL_math.h:387-392 (always strict)
__device__ __forceinline__ RayL camera_ray(const DevCamera& cm, float fx, float fy)
{
    RayL r;
    r.ox = cm.eye[0]; r.oy = cm.eye[1]; r.oz = cm.eye[2];
    const float px = addr(addr(cm.fc[0], mulr(cm.cx[0], fx)), mulr(cm.cy[0], fy));
    const float py = addr(addr(cm.fc[1], mulr(cm.cx[1], fx)), mulr(cm.cy[1], fy));
    const float pz = addr(addr(cm.fc[2], mulr(cm.cx[2], fx)), mulr(cm.cy[2], fy));
    const float dx = subr(px, r.ox), dy = subr(py, r.oy), dz = subr(pz, r.oz);
    const float len2 = addr(addr(mulr(dx, dx), mulr(dy, dy)), mulr(dz, dz));
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(len2));
    r.dx = mulr(dx, inv); r.dy = mulr(dy, inv); r.dz = mulr(dz, inv);
    return r;
}
__device__ __forceinline__ float film_coord(int i, int n) { return addr(-1.0f, mulr(2.0f, __fdiv_rn((float)i, (float)n))); }

// clamp (PL_math.h:430-436), rgb_float_to_byte (:489, truncation), Set_Pixel packing (texture.h:27-38)
__device__ __forceinline__ uint32_t pack_bgrx(float r, float g, float b)
{
    r = 0.0f > (r > 1.0f ? 1.0f : r) ? 0.0f : (r > 1.0f ? 1.0f : r);
    g = 0.0f > (g > 1.0f ? 1.0f : g) ? 0.0f : (g > 1.0f ? 1.0f : g);
    b = 0.0f > (b > 1.0f ? 1.0f : b) ? 0.0f : (b > 1.0f ? 1.0f : b);
    const uint32_t R = (uint32_t)(unsigned char)(int)mulr(r, 255.0f), G = (uint32_t)(unsigned char)(int)mulr(g, 255.0f),
                   B = (uint32_t)(unsigned char)(int)mulr(b, 255.0f);
    return B | (G << 8) | (R << 16);
}

// PCG32 (PL_math.h:506-541) with the per-pixel stream of SURVEY §8(d) C3
struct Pcg { unsigned long long state, stream; };
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x)
{
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}
__device__ __forceinline__ uint32_t pcg_u32(Pcg& s)
{
    const unsigned long long old = s.state;
    s.state = old * 6364136223846793005ULL + (s.stream | 1ULL);
    const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    const uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((0u - rot) & 31u));
}
__device__ __forceinline__ float pcg_bi(Pcg& s) { return addr(-1.0f, mulr(2.0f, mulr((float)pcg_u32(s), 2.328306437e-10F))); }

__device__ __forceinline__ void flush_counters(DevCounters* c, const WorkCount& wc, unsigned int rays)
{
    auto red = [](unsigned int v) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL_MASK, v, off);
        return (unsigned long long)v;
    };
    const unsigned long long r = red(rays), b = red(wc.box), t = red(wc.tri), d = red(wc.pdet), u = red(wc.pu), v = red(wc.pv),
                             l = red(wc.listed), s = red(wc.scanned), o = red(wc.ovf), rc = red(wc.rect), ts = red(wc.tsky),
                             tb = red(wc.tbusy);
    if ((threadIdx.x & 31) == 0)
    {
        atomicAdd(&c->rays, r); atomicAdd(&c->box_tests, b); atomicAdd(&c->tri_tests, t); atomicAdd(&c->pass_det, d);
        atomicAdd(&c->pass_u, u); atomicAdd(&c->pass_v, v); atomicAdd(&c->leaves_listed, l); atomicAdd(&c->leaves_scanned, s);
        atomicAdd(&c->overflow_rays, o); atomicAdd(&c->rect_tests, rc); atomicAdd(&c->tiles_sky, ts); atomicAdd(&c->tiles_busy, tb);
    }
}

// ---- K1: fused primary-ray render (replaces render_tile_from_camera and the tile thread pool) ----------------
template <bool STRICT, bool COUNT>
__global__ void __launch_bounds__(kBlock) k_render(DevScene sc, DevCamera cm, DevPartition part, DevOut out, DevScratch scratch,
                                                   unsigned int* tile_counter, DevCounters* counters,
                                                   unsigned long long jitter_seed)
{
    const int lane = threadIdx.x & 31;
    const int warp_slot = (blockIdx.x * kBlock + threadIdx.x) >> 5;
    const int tiles_x = (cm.W + 7) >> 3;
    const int rows_per_band = part.band_rows >> 2;  // tile rows per band (band_rows is a multiple of 4)
    const int n_bands = (cm.H + part.band_rows - 1) / part.band_rows;
    const int my_bands = part.rank < n_bands ? (n_bands - part.rank + part.world - 1) / part.world : 0;
    const unsigned int tiles_per_band = (unsigned int)(tiles_x * rows_per_band);
    const unsigned int n_tiles = (unsigned int)my_bands * tiles_per_band;
    WorkCount wc;
    wc.clear();
    uint2 leaves[kLeafCap];    // per-lane sorted leaf list (local memory)
    uint32_t stack[kStackCap]; // per-lane inner-node stack of the unculled traversal
    unsigned int rays_done = 0;
    for (;;)
    {
        unsigned int tile = 0;
        if (lane == 0) tile = cancel_requested(sc) ? 0xffffffffu : atomicAdd(tile_counter, 1u);
        tile = __shfl_sync(FULL_MASK, tile, 0);
        if (tile >= n_tiles) break;
        const unsigned int lb = tile / tiles_per_band, rem = tile - lb * tiles_per_band;
        const int trow = (int)(rem / (unsigned int)tiles_x), tcol = (int)(rem - (unsigned int)trow * (unsigned int)tiles_x);
        const int x = tcol * 8 + (lane & 7);
        const int y = ((int)lb * part.world + part.rank) * part.band_rows + trow * 4 + (lane >> 3);
        const bool active = x < cm.W && y < cm.H;
        const float film_y = film_coord(y, cm.H);                                                    // renderer.cpp:317
        const float film_x = mulr(mulr(film_coord(x, cm.W), cm.h_fov), cm.aspect);                  // renderer.cpp:329
        float t = kMaxFloat;
        uint32_t face = 0xffffffffu;
        float cr = 0.0f, cg = 0.0f, cb = 0.0f;
        if (!cm.aa)
        {
            const RayL ray = camera_ray(cm, film_x, film_y);  // renderer.cpp:350-351
            trace_warp<STRICT, COUNT>(sc, scratch, warp_slot, lane, active, ray, leaves, stack, t, face, wc);
            const bool hit = face != 0xffffffffu;
            const float er = hit ? sc.mdl[0] : sc.sky[0], eg = hit ? sc.mdl[1] : sc.sky[1], eb = hit ? sc.mdl[2] : sc.sky[2];
            for (uint32_t s = 0; s < cm.spp; ++s) { cr = addr(cr, er); cg = addr(cg, eg); cb = addr(cb, eb); }  // :353-356
            if (active) rays_done += 1;
        }
        else
        {
            const unsigned long long p = (unsigned long long)y * (unsigned long long)cm.W + (unsigned long long)x;
            Pcg rng;
            rng.state = splitmix64(jitter_seed ^ p);
            rng.stream = 2ULL * p + 1ULL;
            for (uint32_t s = 0; s < cm.spp; ++s)  // renderer.cpp:336-347
            {
                const float x_off = addr(mulr(pcg_bi(rng), cm.hpw), film_x);
                const float y_off = addr(mulr(pcg_bi(rng), cm.hph), film_y);
                const RayL ray = camera_ray(cm, x_off, y_off);
                trace_warp<STRICT, COUNT>(sc, scratch, warp_slot, lane, active, ray, leaves, stack, t, face, wc);
                const bool hit = face != 0xffffffffu;
                if (hit) { pcg_bi(rng); pcg_bi(rng); pcg_bi(rng); }  // cast_ray's random bounce draws (renderer.cpp:246)
                cr = addr(cr, hit ? sc.mdl[0] : sc.sky[0]);
                cg = addr(cg, hit ? sc.mdl[1] : sc.sky[1]);
                cb = addr(cb, hit ? sc.mdl[2] : sc.sky[2]);
                if (active) rays_done += 1;
            }
        }
        if (active)
        {
            const float n = (float)cm.spp;
            const size_t idx = (size_t)y * (size_t)cm.W + (size_t)x;
            if (out.bgrx) out.bgrx[idx] = pack_bgrx(__fdiv_rn(cr, n), __fdiv_rn(cg, n), __fdiv_rn(cb, n));  // :358-365
            if (out.t) out.t[idx] = t;
            if (out.prim) out.prim[idx] = face;
        }
    }
    if (COUNT) flush_counters(counters, wc, rays_done);
}

// ---- per-frame film-space bounds (view_bounds.h) as packed pixel ranges ---------------------------------------------------
// uint2 {x0 | y0 << 16, x1 | y1 << 16}, inclusive; coordinates < 65535 (checked on the host), empty = {0xffffffff, 0}.
__device__ __forceinline__ uint2 px_empty() { return make_uint2(0xffffffffu, 0u); }
__device__ __forceinline__ uint2 px_pack(const PxRect& p)
{
    return make_uint2((uint32_t)p.x0 | ((uint32_t)p.y0 << 16), (uint32_t)p.x1 | ((uint32_t)p.y1 << 16));
}
__device__ __forceinline__ uint2 px_union(uint2 a, uint2 b) { return make_uint2(__vminu2(a.x, b.x), __vmaxu2(a.y, b.y)); }
// both halfwords of a.lo <= b.hi and of a.hi >= b.lo (an empty range overlaps nothing: its lo halves are 0xffff)
__device__ __forceinline__ bool px_overlap(uint2 a, uint2 b)
{
    return (__vcmpleu2(a.x, b.y) & __vcmpgeu2(a.y, b.x)) == 0xffffffffu;
}
__device__ __forceinline__ bool px_contains(uint2 a, uint32_t p) { return (__vcmpleu2(a.x, p) & __vcmpgeu2(a.y, p)) == 0xffffffffu; }
__device__ __forceinline__ bool px_is_empty(uint2 a) { return (a.x & 0xffffu) > (a.y & 0xffffu) || (a.x >> 16) > (a.y >> 16); }
__device__ __forceinline__ uint2 px_union_xor(uint2 r, int off)
{
    return px_union(r, make_uint2(__shfl_xor_sync(FULL_MASK, r.x, off), __shfl_xor_sync(FULL_MASK, r.y, off)));
}
// tile row ty (4 scanlines) belongs to this rank: band_rows is a multiple of 4, band b -> rank b % world
__device__ __forceinline__ bool owns_tile_row(const DevPartition& part, int ty)
{
    return part.world <= 1 || ((ty * 4) / part.band_rows) % part.world == part.rank;
}
// some scanline of bin row `by` (kBinH scanlines) belongs to this rank
__device__ __forceinline__ bool owns_bin_row(const DevPartition& part, int by, int H)
{
    if (part.world <= 1) return true;
    const int y0 = by << kBinShiftY;
    int y1 = y0 + (1 << kBinShiftY) - 1;
    if (y1 > H - 1) y1 = H - 1;
    const int br = part.band_rows;
    const bool p2 = (br & (br - 1)) == 0 && (part.world & (part.world - 1)) == 0;   // the usual case: shifts and masks
    const int b0 = p2 ? y0 >> (__ffs(br) - 1) : y0 / br, b1 = p2 ? y1 >> (__ffs(br) - 1) : y1 / br;
    if (b1 - b0 + 1 >= part.world) return true;
    const int d = p2 ? ((part.rank - b0) & (part.world - 1)) : ((part.rank - b0) % part.world + part.world) % part.world;   // bands from b0 to this rank's next
    return d <= b1 - b0;
}

// One thread per triangle slot: film rectangle (FP64, view_bounds.h) -> pixel range, unions per packet (8 lanes) and per
// chunk of 32 packets (the block), tile stamps and tile costs. One GPU (pe.n == 1): results go to this GPU's tables and only
// the tiles of this rank's bands are stamped. Shared pass (pe.n == world): this grid covers the rank's share of the slots
// and stores every result into the tables of all ranks — plain stores through peer pointers, the fused compute + NVLink
// exchange of the pass — while the stamp and cost of a tile go to the rank that owns its band.
__global__ void __launch_bounds__(kBlock, 4) k_view_bounds(DevScene sc, ViewParams vp, DevView view, DevPartition part, ViewPeers pe)
{
    const int lane = threadIdx.x & 31;
    const uint32_t tid = blockIdx.x * kBlock + threadIdx.x;
    for (uint32_t i = tid; i < (uint32_t)kCtlTicket; i += gridDim.x * kBlock) view.busy_count[i] = 0u;  // list lengths, cursors, finished-tile counters
    for (uint32_t i = tid; i < 2u * (uint32_t)(view.bins_x * view.bins_y); i += gridDim.x * kBlock) view.bin_cnt[i] = 0u;
    const uint32_t chunk = pe.block_lo + blockIdx.x;
    const uint32_t slot = chunk * kBlock + threadIdx.x;
    const bool shared = pe.n > 1;
    const bool in_range = slot < sc.n_slots && chunk < pe.block_hi;   // (an empty share still launches one block for the resets above)
    const int n_tiles = vp.tiles_x * vp.tiles_y;
    // owner of tile row ty in a shared pass (band b -> rank b % world); -1: not stamped (another rank's row, one-GPU pass)
    const int rpb = part.band_rows >> 2;   // tile rows per band; shifts when both are powers of two (the usual 8-row bands on 2/4/8 GPUs)
    const bool pow2 = (rpb & (rpb - 1)) == 0 && (part.world & (part.world - 1)) == 0;
    const int rsh = __ffs(rpb) - 1, wmask = part.world - 1;
    auto owner = [&](int ty) -> int {
        if (pe.ranges_only) return -1;   // shared pass, ranges only: every rank stamps its own tiles afterwards (k_view_stamp)
        if (pe.debug_skip & 32u) return -1;
        if (part.world <= 1) return 0;
        const int o = pow2 ? ((ty >> rsh) & wmask) : ((ty / rpb) % part.world);
        return shared ? o : (o == part.rank ? 0 : -1);
    };
    const bool skip_cost = pe.debug_skip & 1u, skip_stamps = pe.debug_skip & 2u, skip_ranges = pe.debug_skip & 4u;
    // ---- 1: which slots need the double-precision path. About half the slots of a closed mesh face away from the eye; a
    // single-precision test finds them (tri_backfacing_f32), the others are compacted so that the FP64 code runs on full warps.
    __shared__ uint2 s_slot_px[kBlock];
    __shared__ unsigned short s_list[kBlock];
    __shared__ int s_warp_n[kBlock / 32];
    bool need = false;
    if (in_range)
    {
        const float4* rec = sc.tris + (size_t)(slot >> 3) * 24 + (slot & 7);
        const float4 A = __ldg(rec), AB = __ldg(rec + 8), AC = __ldg(rec + 16);
        need = __float_as_uint(A.w) != 0xffffffffu && !tri_backfacing_f32(vp, &A.x, &AB.x, &AC.x);
        if (view.tri_rect && !need) view.tri_rect[slot] = make_float4(INFINITY, INFINITY, -INFINITY, -INFINITY);
    }
    s_slot_px[threadIdx.x] = px_empty();
    const unsigned nm = __ballot_sync(FULL_MASK, need);
    if (lane == 0) s_warp_n[threadIdx.x >> 5] = __popc(nm);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kBlock / 32; ++w)
    {
        const int c = s_warp_n[w];
        if (w < (int)(threadIdx.x >> 5)) before += c;
        total += c;
    }
    if (need) s_list[before + __popc(nm & ((1u << lane) - 1u))] = (unsigned short)threadIdx.x;
    __syncthreads();
    // ---- 2: the compacted slots, full warps: film rectangle -> pixel range -> tile stamps
    int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
    bool big = false;
    if ((int)threadIdx.x < total)
    {
        const uint32_t local = s_list[threadIdx.x], sl = chunk * kBlock + local;
        const float4* rec = sc.tris + (size_t)(sl >> 3) * 24 + (sl & 7);
        const float4 A = __ldg(rec), AB = __ldg(rec + 8), AC = __ldg(rec + 16);
        const FilmRect fr = tri_film_rect(vp, &A.x, &AB.x, &AC.x);
        if (view.tri_rect) view.tri_rect[sl] = make_float4(fr.xlo, fr.ylo, fr.xhi, fr.yhi);
        PxRect p;
        if (film_rect_px(vp, fr, p))
        {
            s_slot_px[local] = px_pack(p);
            tx0 = p.x0 >> 3; tx1 = p.x1 >> 3; ty0 = p.y0 >> 2; ty1 = p.y1 >> 2;
            const int nt = (tx1 - tx0 + 1) * (ty1 - ty0 + 1);
            if (2 * nt >= n_tiles)
            {
                for (int q = 0; q < pe.n; ++q) *pe.all_flag[q] = vp.frame_id;
                tx1 = tx0 - 1;
            }
            else if (nt > 48) big = true;
            else
            {
                for (int ty = ty0; ty <= ty1; ++ty)
                {
                    const int o = owner(ty);
                    if (o < 0 || (skip_stamps && o != pe.self)) continue;
                    uint32_t* st = pe.stamps[o] + ty * vp.tiles_x;
                    for (int tx = tx0; tx <= tx1; ++tx) st[tx] = vp.frame_id;
                }
            }
        }
    }
    // large ranges: queued for the node blocks of k_view_nodes (a warp per range; a chunk of slots close to the eye can hold two hundred of them, which
    // made the warps of that one block the critical path of the pass) — unless the stamps go to other GPUs (first shared version)
    // or the queue is full: then the whole warp stamps the tiles of one triangle at a time
    unsigned bm = __ballot_sync(FULL_MASK, big);
    if (bm && !shared && !pe.ranges_only && !pe.inline_big)
    {
        uint32_t at = 0;
        if (lane == 0) at = atomicAdd(&view.busy_count[kCtlBig], (uint32_t)__popc(bm));
        at = __shfl_sync(FULL_MASK, at, 0);
        if (at + (uint32_t)__popc(bm) <= kBigCap)
        {
            if (big) view.big_list[at + __popc(bm & ((1u << lane) - 1u))] = s_slot_px[s_list[threadIdx.x]];
            bm = 0u;
        }
    }
    if (pe.ranges_only) bm = 0u;
    while (bm)
    {
        const int src = __ffs(bm) - 1;
        bm &= bm - 1;
        const int ax0 = __shfl_sync(FULL_MASK, tx0, src), ax1 = __shfl_sync(FULL_MASK, tx1, src);
        const int ay0 = __shfl_sync(FULL_MASK, ty0, src), ay1 = __shfl_sync(FULL_MASK, ty1, src);
        const int w = ax1 - ax0 + 1, n = w * (ay1 - ay0 + 1);
        const float rw = __frcp_rn((float)w);
        for (int i = lane; i < n; i += 32)
        {
            int r = (int)(((float)i + 0.5f) * rw);   // i / w through the float reciprocal, corrected by one where it rounds off
            if (r * w > i) --r;
            if ((r + 1) * w <= i) ++r;
            const int ty = ay0 + r, o = owner(ty);
            if (o >= 0 && !(skip_stamps && o != pe.self)) pe.stamps[o][ty * vp.tiles_x + ax0 + (i - r * w)] = vp.frame_id;
        }
    }
    __syncthreads();
    // ---- 3: ranges per slot, packet and chunk
    const uint2 px = s_slot_px[threadIdx.x];
    if (in_range)
        for (int q = 0; q < pe.n; ++q)
            if (!skip_ranges || q == pe.self) pe.tri_px[q][slot] = px;
    uint2 p = px;
    p = px_union_xor(p, 4); p = px_union_xor(p, 2); p = px_union_xor(p, 1);
    if (in_range && (lane & 7) == 0)
        for (int q = 0; q < pe.n; ++q)
            if (!skip_ranges || q == pe.self) pe.pkt_px[q][slot >> 3] = p;
    // tile cost = packet ranges touching the tile (what the leaf scan of the render kernel tests against the tile); the
    // packet's 8 lanes share the tiles of its range (every lane holds the same p after the butterfly)
    // Sampled: one packet in 2^cost_sample adds to the costs — the classes are octaves of the count, so a quarter of the packets
    // orders the tiles as well. From a quarter on (at most one sampled packet per warp: its first) the whole warp shares that
    // packet's tiles, lanes flat over the range, instead of 8 lanes walking its rows while 24 wait at the block barrier.
    if (view.cost_sample >= 2)
    {
        const uint2 p0 = make_uint2(__shfl_sync(FULL_MASK, p.x, 0), __shfl_sync(FULL_MASK, p.y, 0));
        const bool sampled = ((((slot - (uint32_t)lane) >> 3) & ((1u << view.cost_sample) - 1u)) == 0u) && !(pe.debug_skip & 16u);
        if (sampled && !px_is_empty(p0))
        {
            const int qx0 = (int)(p0.x & 0xffffu) >> 3, qy0 = (int)(p0.x >> 16) >> 2, qx1 = (int)(p0.y & 0xffffu) >> 3, qy1 = (int)(p0.y >> 16) >> 2;
            const int w = qx1 - qx0 + 1, n = w * (qy1 - qy0 + 1);
            if (n <= 64)
            {
                const float rw = __frcp_rn((float)w);
                for (int i = lane; i < n; i += 32)
                {
                    int r = (int)(((float)i + 0.5f) * rw);
                    if (r * w > i) --r;
                    if ((r + 1) * w <= i) ++r;
                    const int ty = qy0 + r, o = owner(ty);
                    if (o >= 0 && !(skip_cost && o != pe.self)) atomicAdd(pe.tile_cost[o] + ty * vp.tiles_x + qx0 + (i - r * w), 1u);
                }
            }
        }
    }
    else if (!px_is_empty(p) && !(pe.debug_skip & 16u) && ((slot >> 3) & ((1u << view.cost_sample) - 1u)) == 0u)
    {
        const int qx0 = (int)(p.x & 0xffffu) >> 3, qy0 = (int)(p.x >> 16) >> 2, qx1 = (int)(p.y & 0xffffu) >> 3, qy1 = (int)(p.y >> 16) >> 2;
        if ((qx1 - qx0 + 1) * (qy1 - qy0 + 1) <= 64)
            for (int ty = qy0; ty <= qy1; ++ty)   // rows in turn, the 8 lanes stride over the columns (no division per tile)
            {
                const int o = owner(ty);
                if (o < 0 || (skip_cost && o != pe.self)) continue;
                uint32_t* row = pe.tile_cost[o] + ty * vp.tiles_x;
                for (int tx = qx0 + (lane & 7); tx <= qx1; tx += 8) atomicAdd(row + tx, 1u);
            }
    }
    // chunk = the 32 packets of this block
    p = px_union_xor(p, 8); p = px_union_xor(p, 16);
    __shared__ uint2 s_chunk[kBlock / 32];
    if (lane == 0) s_chunk[threadIdx.x >> 5] = p;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        uint2 c = s_chunk[0];
#pragma unroll
        for (int w = 1; w < kBlock / 32; ++w) c = px_union(c, s_chunk[w]);
        if (chunk < pe.block_hi || pe.block_hi == 0u)
            for (int q = 0; q < pe.n; ++q)
                if (!skip_ranges || q == pe.self) pe.chunk_px[q][chunk] = c;
    }
    if (pe.ticket)
    {
        // every store of this block is visible system-wide before the block counts itself done; the last block of the grid
        // tells every rank that this rank's share of the frame's tables is complete
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0)
        {
            const uint32_t done = atomicAdd(pe.ticket, 1u);
            if (done == gridDim.x - 1u)
            {
                __threadfence_system();
                *pe.ticket = 0u;
                for (int q = 0; q < pe.n; ++q) *((volatile uint32_t*)(pe.flags[q] + pe.self)) = pe.seq;
            }
        }
    }
}

// Shared pass: the tables are complete once every other rank's flag in this GPU's flag row has reached the frame's sequence
// number (raised by the last block of that rank's k_view_bounds, on another GPU, after its stores). The calling kernel has
// not read anything of the tables yet, so no stale line can sit in L1. Ends with a block barrier.
__device__ __forceinline__ void wait_for_peer_tables(const ViewWait& vwait)
{
    if (threadIdx.x < (unsigned)vwait.n && (int)threadIdx.x != vwait.rank)
    {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        const volatile uint32_t* f = vwait.my_flags + threadIdx.x;
        while ((int32_t)(*f - vwait.seq) < 0)
        {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > vwait.timeout_ns) { *vwait.timed_out = vwait.seq; break; }
            __nanosleep(64);
        }
        __threadfence_system();
    }
    __syncthreads();
}

// Shared pass, second kernel: the slot ranges of the whole scene are in this GPU's tables (1/N computed here, the rest stored by
// the peers over NVLink); every rank now stamps the tiles of ITS bands and adds up their costs from the ranges — local stores and
// local atomics only (the first version of the shared pass sent each stamp and each cost increment to the owning GPU: scattered
// 4-byte remote writes, slower than computing the whole pass alone; profiles/r02i). One thread per slot, a block per chunk of 256
// slots; chunks whose union range misses this rank's rows (or is empty: back-facing / off-screen) leave after one load.
__global__ void __launch_bounds__(kBlock) k_view_stamp(DevScene sc, ViewParams vp, DevView view, DevPartition part, ViewWait vwait, uint32_t dbg)
{
    const int lane = threadIdx.x & 31;
    if (vwait.my_flags) wait_for_peer_tables(vwait);
    const uint32_t chunk = blockIdx.x;
    const uint32_t slot = chunk * kBlock + threadIdx.x;
    const int rpb = part.band_rows >> 2;
    const bool pow2 = (rpb & (rpb - 1)) == 0 && (part.world & (part.world - 1)) == 0;
    const int rsh = __ffs(rpb) - 1, wmask = part.world - 1;
    auto mine = [&](int ty) -> bool {
        if (part.world <= 1) return true;
        return (pow2 ? ((ty >> rsh) & wmask) : ((ty / rpb) % part.world)) == part.rank;
    };
    {
        const uint2 c = view.chunk_px[chunk];
        if (px_is_empty(c)) return;
        const int cy0 = (int)(c.x >> 16) >> 2, cy1 = (int)(c.y >> 16) >> 2;
        bool any = cy1 - cy0 + 1 >= rpb * part.world;
        for (int ty = cy0; ty <= cy1 && !any; ++ty) any = mine(ty);
        if (!any) return;
    }
    const int n_tiles = vp.tiles_x * vp.tiles_y;
    const uint2 px = slot < sc.n_slots ? view.tri_px[slot] : px_empty();
    int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
    bool big = false;
    if (dbg & 8u) return;
    if (!px_is_empty(px) && !(dbg & 2u))
    {
        tx0 = (int)(px.x & 0xffffu) >> 3; tx1 = (int)(px.y & 0xffffu) >> 3; ty0 = (int)(px.x >> 16) >> 2; ty1 = (int)(px.y >> 16) >> 2;
        const int nt = (tx1 - tx0 + 1) * (ty1 - ty0 + 1);
        if (2 * nt >= n_tiles) tx1 = tx0 - 1;   // (k_view_bounds raised the all-tiles flag on every rank)
        else if (nt > 48) big = true;
        else
            for (int ty = ty0; ty <= ty1; ++ty)
            {
                if (!mine(ty)) continue;
                uint32_t* st = view.stamps + ty * vp.tiles_x;
                for (int tx = tx0; tx <= tx1; ++tx) st[tx] = vp.frame_id;
            }
    }
    unsigned bm = __ballot_sync(FULL_MASK, big);
    if (bm)   // queued for k_view_nodes, as in k_view_bounds
    {
        uint32_t at = 0;
        if (lane == 0) at = atomicAdd(&view.busy_count[kCtlBig], (uint32_t)__popc(bm));
        at = __shfl_sync(FULL_MASK, at, 0);
        if (at + (uint32_t)__popc(bm) <= kBigCap)
        {
            if (big) view.big_list[at + __popc(bm & ((1u << lane) - 1u))] = px;
            bm = 0u;
        }
    }
    while (bm)
    {
        const int src = __ffs(bm) - 1;
        bm &= bm - 1;
        const int ax0 = __shfl_sync(FULL_MASK, tx0, src), ax1 = __shfl_sync(FULL_MASK, tx1, src);
        const int ay0 = __shfl_sync(FULL_MASK, ty0, src), ay1 = __shfl_sync(FULL_MASK, ty1, src);
        const int w = ax1 - ax0 + 1, n = w * (ay1 - ay0 + 1);
        const float rw = __frcp_rn((float)w);
        for (int i = lane; i < n; i += 32)   // flat over the range's tiles (a row-by-row loop leaves most lanes idle on narrow ranges)
        {
            int r = (int)(((float)i + 0.5f) * rw);   // i / w through the float reciprocal, corrected by one where it rounds off
            if (r * w > i) --r;
            if ((r + 1) * w <= i) ++r;
            const int ty = ay0 + r;
            if (mine(ty)) view.stamps[ty * vp.tiles_x + ax0 + (i - r * w)] = vp.frame_id;
        }
    }
    // tile cost = packet ranges touching the tile, as in k_view_bounds
    uint2 p = px;
    p = px_union_xor(p, 4); p = px_union_xor(p, 2); p = px_union_xor(p, 1);
    if (!px_is_empty(p) && !(dbg & 1u) && ((slot >> 3) & ((1u << view.cost_sample) - 1u)) == 0u)
    {
        const int qx0 = (int)(p.x & 0xffffu) >> 3, qy0 = (int)(p.x >> 16) >> 2, qx1 = (int)(p.y & 0xffffu) >> 3, qy1 = (int)(p.y >> 16) >> 2;
        if ((qx1 - qx0 + 1) * (qy1 - qy0 + 1) <= 64)
            for (int ty = qy0; ty <= qy1; ++ty)
            {
                if (!mine(ty)) continue;
                uint32_t* row = view.tile_cost + ty * vp.tiles_x;
                for (int tx = qx0 + (lane & 7); tx <= qx1; tx += 8) atomicAdd(row + tx, 1u);
            }
    }
}

// Second pass, two kinds of blocks running side by side:
//  * blocks [0, node_blocks): one warp per tree node (leaf or inner) unions the packet ranges of the node's subtree — one
//    contiguous range of packets, whole chunks of 32 read from chunk_px — and appends the node to the bins its range
//    touches (flat per-bin lists the render kernel filters per tile instead of walking the tree);
//  * the other blocks compact the stamps of this rank's tiles into the busy list, kTilesPerThread tiles per thread and ONE
//    global atomic per block and slab (the warps reserve their places in shared memory first; per-warp global atomics
//    would serialise on the handful of slab counters).
static constexpr int kTilesPerThread = 4;
// floor(log2(cost)) - shift, clamped to the classes. Measured on the 4K frame (tools/cost_sweep.sh): 8 classes
// (shift 0: 1, 2-3, 4-7, … ≥ 128) take the render kernel from 275 to 257 us; 16 half-octave classes do no better.
__device__ __forceinline__ int tile_cost_class(uint32_t cost, int shift)
{
    const int c = 31 - __clz((int)(cost | 1u)) - shift;
    return c < 0 ? 0 : (c > kCostBuckets - 1 ? kCostBuckets - 1 : c);
}
__global__ void __launch_bounds__(kBlock) k_view_nodes(DevScene sc, DevView view, DevCamera cm, DevPartition part, DevSlabs slabs,
                                                       uint32_t node_blocks, ViewWait vwait)
{
    const int lane = threadIdx.x & 31;
    if (vwait.my_flags) wait_for_peer_tables(vwait);
    if (blockIdx.x < node_blocks)
    {
        const uint32_t warp = (blockIdx.x * kBlock + threadIdx.x) >> 5;
        const uint32_t n_warps = (node_blocks * kBlock) >> 5;
        // First the queued large ranges (DevView::big_list, filled by k_view_bounds / k_view_stamp): one warp per range stamps the
        // tiles of this rank's bands, lanes flat over the range's tiles. The compaction blocks below read the stamps, so every
        // node block counts itself done once its share is stamped and they wait for that count (the grid — at most 592 blocks of
        // 32 registers — is resident at once, and the node blocks have the lowest indices).
        {
            const uint32_t n_big = min(view.busy_count[kCtlBig], kBigCap);
            const int rpb = part.band_rows >> 2;
            const bool pow2 = (rpb & (rpb - 1)) == 0 && (part.world & (part.world - 1)) == 0;
            const int rsh = __ffs(rpb) - 1, wmask = part.world - 1;
            for (uint32_t k = warp; k < n_big; k += n_warps)
            {
                const uint2 px = view.big_list[k];
                const int ax0 = (int)(px.x & 0xffffu) >> 3, ax1 = (int)(px.y & 0xffffu) >> 3, ay0 = (int)(px.x >> 16) >> 2, ay1 = (int)(px.y >> 16) >> 2;
                const int w = ax1 - ax0 + 1, n = w * (ay1 - ay0 + 1);
                const float rw = __frcp_rn((float)w);
                for (int i = lane; i < n; i += 32)
                {
                    int r = (int)(((float)i + 0.5f) * rw);   // i / w through the float reciprocal, corrected by one where it rounds off
                    if (r * w > i) --r;
                    if ((r + 1) * w <= i) ++r;
                    const int ty = ay0 + r;
                    const bool mine = part.world <= 1 || (pow2 ? ((ty >> rsh) & wmask) : ((ty / rpb) % part.world)) == part.rank;
                    if (mine) view.stamps[ty * view.tiles_x + ax0 + (i - r * w)] = view.frame_id;
                }
            }
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0)
            {
                const uint32_t d = atomicAdd(&view.busy_count[kCtlBigDone], 1u);
                if (d == node_blocks - 1u) view.busy_count[kCtlBig] = 0u;   // every node block has read the queue's length
            }
        }
        for (uint32_t node = warp; node < sc.n_nodes; node += n_warps)
        {
            const uint2 pk = __ldg(&sc.node_pk[node]);
            const uint32_t pb = pk.x, pe = pk.y;
            uint32_t hb = (pb + 31u) & ~31u;   // packets before the first whole chunk
            if (hb > pe) hb = pe;
            uint2 acc = px_empty();
            for (uint32_t k = pb + lane; k < hb; k += 32) acc = px_union(acc, view.pkt_px[k]);
            if (hb < pe)
            {
                const uint32_t ce = pe >> 5;
#pragma unroll 4
                for (uint32_t c = (hb >> 5) + lane; c < ce; c += 32) acc = px_union(acc, view.chunk_px[c]);   // (root: ~30 rounds, 4 in flight)
                uint32_t tb = ce << 5;
                if (tb < hb) tb = hb;
                for (uint32_t k = tb + lane; k < pe; k += 32) acc = px_union(acc, view.pkt_px[k]);
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) acc = px_union_xor(acc, off);
            if (lane == 0) view.node_px[node] = acc;
            if (px_is_empty(acc) || node == 0u) continue;   // (the root is tested by every tile directly)
            const float4 n0 = __ldg(&sc.nodes[2 * node]), n1 = __ldg(&sc.nodes[2 * node + 1]);
            const bool inner = __float_as_int(n0.w) > 0;
            float4 sp4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (inner) sp4 = __ldg(&sc.nodes[2 * (uint32_t)__float_as_int(n0.w) + 1]);   // split point = hi corner of child 0
            const int bx0 = (int)(acc.x & 0xffffu) >> kBinShiftX, by0 = (int)(acc.x >> 16) >> kBinShiftY;
            const int bx1 = (int)(acc.y & 0xffffu) >> kBinShiftX, by1 = (int)(acc.y >> 16) >> kBinShiftY;
            const int w = bx1 - bx0 + 1, n = w * (by1 - by0 + 1);
            if (inner && n > kWideBins)
            {
                // near the root: one entry in the wide list (every tile checks it) instead of hundreds of bin entries
                uint32_t at = 0;
                if (lane == 0) at = atomicAdd(&view.busy_count[kCtlWide], 1u);
                at = __shfl_sync(FULL_MASK, at, 0);
                if (at < (uint32_t)kWideCap)
                {
                    if (lane == 0)
                    {
                        view.wide_inner[2 * at] = make_uint4(acc.x, acc.y, node, 0u);
                        view.wide_inner[2 * at + 1] = make_uint4(__float_as_uint(sp4.x), __float_as_uint(sp4.y), __float_as_uint(sp4.z), 0u);
                    }
                    continue;
                }
            }
            const float rw = __frcp_rn((float)w);   // i / w through the float reciprocal (corrected by one below)
            for (int i = lane; i < n; i += 32)
            {
                int r = (int)(((float)i + 0.5f) * rw);
                if (r * w > i) --r;
                if ((r + 1) * w <= i) ++r;
                const int by = by0 + r, bx = bx0 + (i - r * w);
                if (!owns_bin_row(part, by, cm.H)) continue;
                const uint32_t b = (uint32_t)(by * view.bins_x + bx);
                if (inner)
                {
                    const uint32_t at = atomicAdd(&view.bin_cnt[2 * b + 1], 1u);
                    if (at < view.bin_inner_cap)
                    {
                        uint4* e = view.bin_inner + ((size_t)b * view.bin_inner_cap + at) * 2;
                        e[0] = make_uint4(acc.x, acc.y, node, 0u);
                        e[1] = make_uint4(__float_as_uint(sp4.x), __float_as_uint(sp4.y), __float_as_uint(sp4.z), 0u);
                    }
                }
                else
                {
                    const uint32_t at = atomicAdd(&view.bin_cnt[2 * b], 1u);
                    if (at < view.bin_leaf_cap)
                    {
                        uint4* e = view.bin_leaf + ((size_t)b * view.bin_leaf_cap + at) * 3;
                        e[0] = make_uint4(acc.x, acc.y, __ldg(&sc.leaf_seq[node]), ~(uint32_t)__float_as_int(n0.w));
                        e[1] = make_uint4(__float_as_uint(n0.x), __float_as_uint(n0.y), __float_as_uint(n0.z), __float_as_uint(n1.w));
                        e[2] = make_uint4(__float_as_uint(n1.x), __float_as_uint(n1.y), __float_as_uint(n1.z), node);
                    }
                }
            }
        }
        return;
    }
    // (the large ranges are stamped by the node blocks: wait for them; bounded, as every cross-block wait here)
    if (threadIdx.x == 0)
    {
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (*((volatile uint32_t*)&view.busy_count[kCtlBigDone]) < node_blocks)
        {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 1000000000ull) break;
            __nanosleep(32);
        }
        __threadfence();
    }
    __syncthreads();
    // busy tiles of this rank, in the render kernel's local numbering (band-major, then tile row, then tile column)
    const int tiles_x = (cm.W + 7) >> 3;
    const int rows_per_band = part.band_rows >> 2;
    const int n_bands = (cm.H + part.band_rows - 1) / part.band_rows;
    const int my_bands = part.rank < n_bands ? (n_bands - part.rank + part.world - 1) / part.world : 0;
    const unsigned int tiles_per_band = (unsigned int)(tiles_x * rows_per_band);
    const unsigned int n_tiles = (unsigned int)my_bands * tiles_per_band;
    const bool all_busy = *((volatile uint32_t*)view.all_flag) == view.frame_id;
    const unsigned int blk = blockIdx.x - node_blocks, n_blk = gridDim.x - node_blocks;
    constexpr unsigned int kStep = kBlock * kTilesPerThread;
    __shared__ unsigned int s_count[kMaxSlabs * kCostBuckets], s_at[kMaxSlabs * kCostBuckets];
    for (unsigned int base = blk * kStep; base < n_tiles; base += n_blk * kStep)
    {
        if (threadIdx.x < kMaxSlabs * kCostBuckets) s_count[threadIdx.x] = 0u;
        __syncthreads();
        bool busy[kTilesPerThread];
        int slab[kTilesPerThread];
#pragma unroll
        for (int k = 0; k < kTilesPerThread; ++k)
        {
            const unsigned int tile = base + k * kBlock + threadIdx.x;
            busy[k] = false;
            int cls = 0;
            if (tile < n_tiles)
            {
                const unsigned int lb = tile / tiles_per_band, rem = tile - lb * tiles_per_band;
                const int trow = (int)(rem / (unsigned int)tiles_x), tcol = (int)(rem - (unsigned int)trow * (unsigned int)tiles_x);
                const int y0 = ((int)lb * part.world + part.rank) * part.band_rows + trow * 4;
                if (y0 < cm.H)
                {
                    const int gt = (y0 >> 2) * view.tiles_x + tcol;
                    busy[k] = all_busy || view.stamps[gt] == view.frame_id;
                    // every tile this rank examines starts the next frame at cost zero (k_view_bounds adds cost to this
                    // rank's tiles only, so no tile keeps a stale count when the camera or the partition changes)
                    const uint32_t cost = view.tile_cost[gt];
                    if (cost) view.tile_cost[gt] = 0u;
                    if (busy[k])
                    {
                        cls = tile_cost_class(cost, view.cost_shift);
                        if (view.tile_log) { view.tile_log[tile].z = cost; view.tile_log[tile].w = (uint32_t)cls; }
                    }
                }
            }
            int sl = 0;
            while (sl + 1 < slabs.n && tile >= slabs.tile_begin[sl + 1]) ++sl;
            slab[k] = sl * kCostBuckets + cls;   // list key: (slab, cost class)
        }
        // pass 1: how many entries does this block add to each slab's segment
#pragma unroll
        for (int k = 0; k < kTilesPerThread; ++k)
        {
            unsigned todo = __ballot_sync(FULL_MASK, busy[k]);
            while (todo)
            {
                const int leader = __ffs(todo) - 1;
                const int ls = __shfl_sync(FULL_MASK, slab[k], leader);
                const unsigned m = __ballot_sync(FULL_MASK, busy[k] && slab[k] == ls);
                todo &= ~m;
                if (lane == leader) atomicAdd(&s_count[ls], (unsigned int)__popc(m));
            }
        }
        __syncthreads();
        if (threadIdx.x < (unsigned)(slabs.n * kCostBuckets) && s_count[threadIdx.x])
            s_at[threadIdx.x] = atomicAdd(&view.busy_count[kCtlCount + threadIdx.x], s_count[threadIdx.x]);
        __syncthreads();
        // pass 2: places inside the block's reservation (segment s starts at busy_list[tile_begin[s]]: room for every tile)
#pragma unroll
        for (int k = 0; k < kTilesPerThread; ++k)
        {
            unsigned todo = __ballot_sync(FULL_MASK, busy[k]);
            while (todo)
            {
                const int leader = __ffs(todo) - 1;
                const int ls = __shfl_sync(FULL_MASK, slab[k], leader);
                const unsigned m = __ballot_sync(FULL_MASK, busy[k] && slab[k] == ls);
                todo &= ~m;
                unsigned int at = 0;
                if (lane == leader) at = atomicAdd(&s_at[ls], (unsigned int)__popc(m));
                at = __shfl_sync(FULL_MASK, at, leader);
                if (busy[k] && slab[k] == ls)
                    view.busy_list[(unsigned int)(ls % kCostBuckets) * view.bucket_stride + slabs.tile_begin[ls / kCostBuckets] + at +
                                   __popc(m & ((1u << lane) - 1u))] = base + k * kBlock + threadIdx.x;
            }
        }
        __syncthreads();
    }
}

// ---- culled busy tile, flat: ONE warp, one 8x4 tile, no tree walk ----------------------------------------------------------
// For one camera ray the reference returns (kd_tree.cpp:337-465): among the leaves the ray lists, in ascending (entry
// distance, insertion order) order, the first one in which some triangle passes the float test, and inside it the minimum
// (t, position). A leaf L is listed iff every inner node above it passes check_ray_AABB_intersection, its own entry distance
// is > 0 and no "at most 5 accepted children" cut (kd_tree.cpp:374) removes it. With the per-frame pixel ranges:
//  * a ray whose pixel lies outside node_px[L] is accepted by no triangle of L (view_bounds.h), so L can be dropped from its
//    list without changing the result — the candidates of a tile are the leaves of its bin whose range touches the tile;
//  * for a ray with finite origin, direction and reciprocals every slab value is a monotone function of the plane it comes
//    from and the child boxes nest (kd_tree.cpp:120-148): a leaf that passes its own slab test passes the test of every
//    ancestor, the leaf's entry distance alone decides (tiles with another ray take the unculled core);
//  * the cut at an inner node n can only trigger for a ray whose three split-plane parameters (s - o)*inv of n's split
//    point s are not pairwise different: the children are the 8 octants around s, both octants that share a split plane
//    use the same float t(s), and with pairwise different parameters at most 4 octants contain a ray parameter in their
//    closed slab intervals. The inner nodes above any candidate leaf of the tile are in the bin's inner list (their range
//    contains the leaf's); a ray with a tie at one of them ("risky", a ray through a split edge up to rounding) is traced
//    by the unculled core instead;
//  * the order in which the leaves are examined is free: per ray the best (entry distance, leaf_seq) key with a hit is
//    kept, a leaf is skipped for the rays that already hold a better key, and (t, position) is minimised inside a leaf —
//    the lexicographic minimum the reference's list order + early exit (kd_tree.cpp:437-462) computes. The candidates of
//    a chunk are taken nearest first (entry distance of the tile's first ray) so that later leaves are mostly skipped.
#ifndef ATRAY_PAIR_CAP
#define ATRAY_PAIR_CAP 256   // (128: render kernel 254 -> 247 us on the stand-in with 256, same box)
#endif
static constexpr int kPairCap = ATRAY_PAIR_CAP;
struct WarpPairs
{
    unsigned long long key[32];        // per ray: min (t bits, position in the reference's leaf order) in the current leaf
    uint32_t face[32];
    uint32_t pair_slot[kPairCap];      // queue of (ray lane, slot) pairs whose pixel lies inside the triangle's range
    unsigned char pair_lane[kPairCap];
};
struct WarpScratch { WarpPairs pairs; };

// The reference triangle test (model.h:75-103) for one ray and one record; true iff the caller would accept the
// distance (t > tolerance && t < MAX_FLOAT, renderer.cpp:52 / kd_tree.cpp:446).
template <bool STRICT, bool COUNT>
__device__ __forceinline__ bool tri_accept(const float4* __restrict__ rec, float ox, float oy, float oz, float dx, float dy, float dz,
                                           float& t_out, uint32_t& pos_out, uint32_t& face_out, WorkCount& wc)
{
    typedef Fp<STRICT> F;
    const float4 AB = __ldg(rec + 8), AC = __ldg(rec + 16);
    const float px = F::crs(dy, AC.z, dz, AC.y), py = F::crs(dz, AC.x, dx, AC.z), pz = F::crs(dx, AC.y, dy, AC.x);
    const float det = F::dot(AB.x, AB.y, AB.z, px, py, pz);
    if (COUNT) wc.tri++;
    if (det < kTol) return false;
    if (COUNT) wc.pdet++;
    const float4 A = __ldg(rec);
    const float det_inv = __frcp_rn(det);
    const float tx = F::sub(ox, A.x), ty = F::sub(oy, A.y), tz = F::sub(oz, A.z);
    const float u = F::mul(F::dot(tx, ty, tz, px, py, pz), det_inv);
    if (u < 0.0f || u > 1.0f) return false;
    if (COUNT) wc.pu++;
    const float qx = F::crs(ty, AB.z, tz, AB.y), qy = F::crs(tz, AB.x, tx, AB.z), qz = F::crs(tx, AB.y, ty, AB.x);
    const float v = F::mul(F::dot(dx, dy, dz, qx, qy, qz), det_inv);
    if (v < 0.0f || F::add(u, v) > 1.0f) return false;
    if (COUNT) wc.pv++;
    const float t = F::mul(F::dot(qx, qy, qz, AC.x, AC.y, AC.z), det_inv);
    t_out = t;
    pos_out = __float_as_uint(AB.w);
    face_out = __float_as_uint(A.w);
    return t > kTol && t < kMaxFloat;
}

// Runs the triangle test on every queued pair, 32 pairs per step: each lane takes one pair, fetches the ray direction
// from its owner lane by shuffle (every camera ray starts at the eye) and gathers the record. Accepted hits lower
// key[ray] = (t bits, position) — t > 0, so the float bits order like the values, and the lexicographic minimum equals
// the reference's sequential strict-< scan.
template <bool STRICT, bool COUNT>
__device__ __forceinline__ void drain_pairs(const DevScene& sc, WarpPairs& wp, int lane, int& qn, const RayL& ray, WorkCount& wc)
{
    __syncwarp();
    for (int base = 0; base < qn; base += 32)
    {
        const bool have = base + lane < qn;
        const int src = have ? (int)wp.pair_lane[base + lane] : lane;
        const uint32_t s = have ? wp.pair_slot[base + lane] : 0u;
        const float dx = __shfl_sync(FULL_MASK, ray.dx, src), dy = __shfl_sync(FULL_MASK, ray.dy, src), dz = __shfl_sync(FULL_MASK, ray.dz, src);
        float t = 0.f;
        uint32_t pos = 0u, face = 0u;
        bool acc = false;
        if (have) acc = tri_accept<STRICT, COUNT>(sc.tris + (size_t)(s >> 3) * 24 + (s & 7), ray.ox, ray.oy, ray.oz, dx, dy, dz, t, pos, face, wc);
        const unsigned long long mk = ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)pos;
        if (acc) atomicMin(&wp.key[src], mk);
        __syncwarp();
        if (acc && wp.key[src] == mk) wp.face[src] = face;  // (t, pos) is unique per ray and leaf: exactly one writer
        __syncwarp();
    }
    qn = 0;
}

// One leaf against the rays `ing` of the tile: packets whose range touches the tile, then their triangles, then the rays
// inside each triangle's range (a bit mask computed from the range and the tile origin), queued as pairs and tested.
// On return key[lane] / face[lane] hold the leaf's best hit of ray `lane` (all-ones when it has none).
template <bool STRICT, bool COUNT>
__device__ __forceinline__ void scan_leaf_flat(const DevScene& sc, const DevView& vw, WarpPairs& wp, int lane, const RayL& ray,
                                               uint32_t first, uint32_t cnt, bool ing, uint2 tr, int tile_x0, int tile_y0, WorkCount& wc)
{
    const unsigned gmask = __ballot_sync(FULL_MASK, ing);
    const uint32_t npk = (cnt + 7u) >> 3, pk0 = first >> 3;
    int qn = 0;
    for (uint32_t pb = 0; pb < npk; pb += 32)
    {
        bool pov = false;
        if (pb + lane < npk)
        {
            pov = px_overlap(__ldg(&vw.pkt_px[pk0 + pb + lane]), tr);
            if (COUNT) wc.rect++;
        }
        unsigned pmask = __ballot_sync(FULL_MASK, pov);
        while (pmask)
        {
            // the next (up to) four touching packets, one per group of 8 lanes
            int sel = -1;
            {
                unsigned m = pmask;
                const int j = lane >> 3;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    const int b = m ? __ffs(m) - 1 : -1;
                    if (q == j) sel = b;
                    m &= m - 1;  // m == 0 stays 0
                }
                pmask = m;
            }
            uint32_t myslot = 0, m32 = 0;
            if (sel >= 0)
            {
                myslot = ((pk0 + pb + (uint32_t)sel) << 3) + (lane & 7);
                const uint2 r = __ldg(&vw.tri_px[myslot]);
                if (COUNT) wc.rect++;
                if (px_overlap(r, tr))
                {
                    // rays of the tile inside the range: columns cx0..cx1 and rows ry0..ry1 relative to the tile origin
                    const int cx0 = max((int)(r.x & 0xffffu) - tile_x0, 0), cx1 = min((int)(r.y & 0xffffu) - tile_x0, 7);
                    const int ry0 = max((int)(r.x >> 16) - tile_y0, 0), ry1 = min((int)(r.y >> 16) - tile_y0, 3);
                    const uint32_t cols = (0xffu >> (7 - cx1)) & (0xffu << cx0) & 0xffu;
                    const uint32_t rows = (0xfu >> (3 - ry1)) & (0xfu << ry0) & 0xfu;
                    const uint32_t spread = (rows & 1u) | ((rows & 2u) << 7) | ((rows & 4u) << 14) | ((rows & 8u) << 21);
                    m32 = (cols * spread) & gmask;
                }
            }
            // pairs of this step: exclusive prefix sum of the per-lane pair counts
            const int n = __popc(m32);
            int incl = n;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1)
            {
                const int v = __shfl_up_sync(FULL_MASK, incl, off);
                if (lane >= off) incl += v;
            }
            const int total = __shfl_sync(FULL_MASK, incl, 31);
            if (total == 0) continue;
            if (qn + total > kPairCap) drain_pairs<STRICT, COUNT>(sc, wp, lane, qn, ray, wc);
            if (total <= kPairCap)
            {
                int at = qn + incl - n;
                while (m32)
                {
                    const int b = __ffs(m32) - 1;
                    m32 &= m32 - 1;
                    wp.pair_slot[at] = myslot;
                    wp.pair_lane[at] = (unsigned char)b;
                    ++at;
                }
                qn += total;
            }
            else
            {
                // more pairs than the queue holds (large triangles over the whole tile): one triangle at a time
                unsigned cm = __ballot_sync(FULL_MASK, m32 != 0u);
                while (cm)
                {
                    const int b = __ffs(cm) - 1;
                    cm &= cm - 1;
                    const uint32_t m = __shfl_sync(FULL_MASK, m32, b), s = __shfl_sync(FULL_MASK, myslot, b);
                    if ((m >> lane) & 1u)
                    {
                        const int at = qn + __popc(m & ((1u << lane) - 1u));
                        wp.pair_slot[at] = s;
                        wp.pair_lane[at] = (unsigned char)lane;
                    }
                    qn += __popc(m);
                    if (qn > kPairCap - 32) drain_pairs<STRICT, COUNT>(sc, wp, lane, qn, ray, wc);
                }
            }
        }
    }
    drain_pairs<STRICT, COUNT>(sc, wp, lane, qn, ray, wc);
}

// ---- the exact per-ray search (tree walk + sorted leaf list, as trace_warp) for CAMERA rays of the current frame, with the
// leaf scan restricted to the triangles whose pixel range contains the ray's pixel: a triangle outside cannot be accepted
// (view_bounds.h), so leaving it out changes nothing. Dense views (eye inside the model, close-ups) put hundreds of leaves
// into a bin — more than the flat lists hold — and their tiles land here; a leaf of 300 triangles then costs one range load
// per slot and one triangle-test round for the handful that cover the pixel instead of ten rounds.
template <bool STRICT, bool COUNT>
__device__ __forceinline__ void scan_ray_leaves_culled(const DevScene& sc, const DevView& vw, WarpPairs& wp, int lane, int src, const RayL& mine,
                                                       uint32_t mypx, const uint2* my_list, int my_nleaf, float& out_t, uint32_t& out_face,
                                                       WorkCount& wc)
{
    typedef Fp<STRICT> F;
    const float ox = __shfl_sync(FULL_MASK, mine.ox, src), oy = __shfl_sync(FULL_MASK, mine.oy, src), oz = __shfl_sync(FULL_MASK, mine.oz, src);
    const float dx = __shfl_sync(FULL_MASK, mine.dx, src), dy = __shfl_sync(FULL_MASK, mine.dy, src), dz = __shfl_sync(FULL_MASK, mine.dz, src);
    const uint32_t p = __shfl_sync(FULL_MASK, mypx, src);
    const int n = __shfl_sync(FULL_MASK, my_nleaf, src);
    for (int k = 0; k < n; ++k)
    {
        uint32_t node = 0;
        if (lane == src) node = my_list[k].x;
        node = __shfl_sync(FULL_MASK, node, src);
        const uint32_t first = ~(uint32_t)__float_as_int(__ldg(&sc.nodes[2 * node]).w);
        const uint32_t cnt = (uint32_t)__float_as_int(__ldg(&sc.nodes[2 * node + 1]).w);
        if (COUNT && lane == src) wc.scanned++;
        float bt = kMaxFloat;
        uint32_t bpos = 0xffffffffu, bface = 0xffffffffu;
        int qn = 0;
        const uint32_t padded = (cnt + 7u) & ~7u;
        for (uint32_t base = 0; base < padded || qn > 0; base += 32)
        {
            // queue the slots of the next 32 whose range contains the pixel; test when 32 are waiting or the leaf ends
            if (base < padded)
            {
                const uint32_t slot = first + base + lane;
                const bool cand = base + lane < padded && px_contains(vw.tri_px[slot], p);
                if (COUNT && base + lane < padded) wc.rect++;
                const unsigned m = __ballot_sync(FULL_MASK, cand);
                if (cand) wp.pair_slot[qn + __popc(m & ((1u << lane) - 1u))] = slot;
                qn += __popc(m);
                if (qn < 32 && base + 32 < padded) continue;
            }
            __syncwarp();
            const int take = qn < 32 ? qn : 32;
            const bool have = lane < take;
            const uint32_t slot = have ? wp.pair_slot[qn - take + lane] : 0u;   // (the last `take` entries: the order is free)
            qn -= take;
            __syncwarp();
            if (take == 0) break;
            const float4* rec = sc.tris + (size_t)(slot >> 3) * 24 + (slot & 7);
            float4 A = make_float4(0.f, 0.f, 0.f, __uint_as_float(0xffffffffu)), AB = make_float4(0.f, 0.f, 0.f, 0.f), AC = AB;
            if (have) { A = __ldg(rec); AB = __ldg(rec + 8); AC = __ldg(rec + 16); }
            const bool real = __float_as_uint(A.w) != 0xffffffffu;
            const float px = F::crs(dy, AC.z, dz, AC.y), py = F::crs(dz, AC.x, dx, AC.z), pz = F::crs(dx, AC.y, dy, AC.x);
            const float det = F::dot(AB.x, AB.y, AB.z, px, py, pz);
            bool alive = have && !(det < kTol);
            if (COUNT) { wc.tri += real ? 1u : 0u; wc.pdet += (real && alive) ? 1u : 0u; }
            if (!__any_sync(FULL_MASK, alive)) continue;
            const float det_inv = __frcp_rn(det);
            const float tx = F::sub(ox, A.x), ty = F::sub(oy, A.y), tz = F::sub(oz, A.z);
            const float u = F::mul(F::dot(tx, ty, tz, px, py, pz), det_inv);
            alive = alive && !(u < 0.0f || u > 1.0f);
            if (COUNT) wc.pu += (real && alive) ? 1u : 0u;
            if (!__any_sync(FULL_MASK, alive)) continue;
            const float qx = F::crs(ty, AB.z, tz, AB.y), qy = F::crs(tz, AB.x, tx, AB.z), qz = F::crs(tx, AB.y, ty, AB.x);
            const float v = F::mul(F::dot(dx, dy, dz, qx, qy, qz), det_inv);
            alive = alive && !(v < 0.0f || F::add(u, v) > 1.0f);
            if (COUNT) wc.pv += (real && alive) ? 1u : 0u;
            if (!__any_sync(FULL_MASK, alive)) continue;
            const float t = F::mul(F::dot(qx, qy, qz, AC.x, AC.y, AC.z), det_inv);
            const uint32_t pos = __float_as_uint(AB.w);
            if (alive && t > kTol && (t < bt || (t == bt && pos < bpos)))   // per lane over several rounds: keep the (t, position) minimum
            {
                bt = t;
                bpos = pos;
                bface = __float_as_uint(A.w);
            }
        }
        if (!__any_sync(FULL_MASK, bt < kMaxFloat)) continue;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
        {
            const float ot = __shfl_xor_sync(FULL_MASK, bt, off);
            const uint32_t op = __shfl_xor_sync(FULL_MASK, bpos, off);
            const uint32_t of = __shfl_xor_sync(FULL_MASK, bface, off);
            if (ot < bt || (ot == bt && op < bpos)) { bt = ot; bpos = op; bface = of; }
        }
        if (bt < kMaxFloat)
        {
            if (lane == src) { out_t = bt; out_face = bface; }
            break;
        }
    }
}

// collect_leaves for a CAMERA ray of the current frame: children whose pixel range (node_px, the union over the subtree's
// triangles) does not contain the ray's pixel are not examined — no triangle below them can be accepted for this ray, and
// dropping leaves that cannot produce a hit from the sorted list does not change which leaf produces the first one. The "at most
// 5 accepted children" cut (kd_tree.cpp:374) can only trigger at a node where two of the ray's three split-plane parameters
// coincide (at most 4 octants contain the ray otherwise, see trace_tile_flat); such a node sends the ray to the exact walk.
// Finite rays only (the caller has checked).
template <bool COUNT>
__device__ bool collect_leaves_culled(const float4* __restrict__ nodes, const uint32_t* __restrict__ leaf_seq, const uint2* node_px,
                                      uint32_t mypx, const RayL& r, float ix, float iy, float iz, uint32_t* stack, int stack_cap, uint2* leaves,
                                      int leaf_cap, int& nleaf, WorkCount& wc)
{
    nleaf = 0;
    int sp = 1;
    stack[0] = 0u;
    const unsigned near_mask = (ix < 0.0f ? 4u : 0u) | (iy < 0.0f ? 2u : 0u) | (iz < 0.0f ? 1u : 0u);
    while (sp > 0)
    {
        const uint32_t cur = stack[--sp];
        const uint32_t c0 = (uint32_t)__float_as_int(__ldg(&nodes[2 * cur]).w);
        {
            const float4 s = __ldg(&nodes[2 * c0 + 1]);   // split point = hi corner of child 0 (kd_tree.cpp:120-148)
            const float tsx = mulr(subr(s.x, r.ox), ix), tsy = mulr(subr(s.y, r.oy), iy), tsz = mulr(subr(s.z, r.oz), iz);
            if (COUNT) wc.box++;
            if (tsx == tsy || tsy == tsz || tsx == tsz)
                return collect_leaves<COUNT>(nodes, leaf_seq, r, ix, iy, iz, stack, stack_cap, leaves, leaf_cap, nleaf, wc);
        }
        unsigned inner_acc = 0u;
#pragma unroll 1
        for (int i = 0; i < 8; ++i)
        {
            const uint32_t ch = c0 + i;
            if (!px_contains(node_px[ch], mypx)) continue;
            const float4 lo = __ldg(&nodes[2 * ch]);
            const float4 hi = __ldg(&nodes[2 * ch + 1]);
            if (COUNT) wc.box++;
            if (__float_as_int(lo.w) > 0)
            {
                if (box_check(r, ix, iy, iz, lo, hi)) inner_acc |= 1u << i;
            }
            else
            {
                const float dis = box_entry(r, ix, iy, iz, lo, hi);
                if (dis > 0.0f)
                {
                    if (nleaf >= leaf_cap) return false;
                    int k = nleaf;
                    while (k > 0)
                    {
                        const float pd = __uint_as_float(leaves[k - 1].y);
                        if (!(dis < pd || (dis == pd && __ldg(&leaf_seq[ch]) < __ldg(&leaf_seq[leaves[k - 1].x])))) break;
                        leaves[k] = leaves[k - 1];
                        --k;
                    }
                    leaves[k] = make_uint2(ch, __float_as_uint(dis));
                    ++nleaf;
                    if (COUNT) wc.listed++;
                }
            }
        }
        if (inner_acc)
        {
            if (sp + __popc(inner_acc) > stack_cap) return false;
#pragma unroll
            for (int j = 7; j >= 0; --j)
            {
                const unsigned i = (unsigned)j ^ near_mask;
                if ((inner_acc >> i) & 1u) stack[sp++] = c0 + i;
            }
        }
    }
    return true;
}

template <bool STRICT, bool COUNT>
__device__ __forceinline__ void trace_warp_culled(const DevScene& sc, const DevScratch& scratch, const DevView& vw, WarpPairs& wp, int warp_slot,
                                                  int lane, bool active, const RayL& ray, uint32_t mypx, uint2* leaves, uint32_t* stack,
                                                  float& t_out, uint32_t& face_out, WorkCount& wc)
{
    t_out = kMaxFloat;
    face_out = 0xffffffffu;
    int nleaf = 0;
    bool overflow = false;
    const float ix = __frcp_rn(ray.dx), iy = __frcp_rn(ray.dy), iz = __frcp_rn(ray.dz);
    if (active)
    {
        const float4 lo = __ldg(&sc.nodes[0]);
        const float4 hi = __ldg(&sc.nodes[1]);
        if (COUNT) wc.box++;
        if (box_check(ray, ix, iy, iz, lo, hi))
        {
            if (sc.root_is_leaf) { leaves[0] = make_uint2(0u, 0u); nleaf = 1; }
            else overflow = !collect_leaves_culled<COUNT>(sc.nodes, sc.leaf_seq, vw.node_px, mypx, ray, ix, iy, iz, stack, kStackCap, leaves, kLeafCap, nleaf, wc);
            if (overflow) nleaf = 0;
        }
    }
    unsigned todo = __ballot_sync(FULL_MASK, nleaf > 0);
    while (todo)
    {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        scan_ray_leaves_culled<STRICT, COUNT>(sc, vw, wp, lane, src, ray, mypx, leaves, nleaf, t_out, face_out, wc);
    }
    unsigned spill = __ballot_sync(FULL_MASK, overflow);
    while (spill)
    {
        const int src = __ffs(spill) - 1;
        spill &= spill - 1;
        uint2* gl = scratch.leaves + (size_t)warp_slot * scratch.leaf_cap;
        uint32_t* gs = scratch.stack + (size_t)warp_slot * scratch.stack_cap;
        int gn = 0;
        if (lane == src)
        {
            if (COUNT) wc.ovf++;
            WorkCount dummy;
            dummy.clear();
            collect_leaves_culled<false>(sc.nodes, sc.leaf_seq, vw.node_px, mypx, ray, ix, iy, iz, gs, (int)scratch.stack_cap, gl, (int)scratch.leaf_cap, gn, dummy);
        }
        __syncwarp();
        scan_ray_leaves_culled<STRICT, COUNT>(sc, vw, wp, lane, src, ray, mypx, gl, gn, t_out, face_out, wc);
        __syncwarp();
    }
}

template <bool STRICT, bool COUNT>
__device__ __forceinline__ void trace_tile_flat(const DevScene& sc, const DevScratch& scratch, const DevView& vw, WarpScratch& ws,
                                                int warp_slot, int lane, bool active, const RayL& ray, int px_x, int px_y,
                                                int tile_x0, int tile_y0, uint2* leaves, uint32_t* stack, float& t_out,
                                                uint32_t& face_out, WorkCount& wc)
{
    WarpPairs& wp = ws.pairs;
    const float ix = __frcp_rn(ray.dx), iy = __frcp_rn(ray.dy), iz = __frcp_rn(ray.dz);
    // finite origin, direction and reciprocals (no 0*inf, inf-inf or NaN in any slab product)
    const float chk = (ray.ox - ray.ox) + (ray.oy - ray.oy) + (ray.oz - ray.oz) + (ix - ix) + (iy - iy) + (iz - iz) +
                      (ray.dx - ray.dx) + (ray.dy - ray.dy) + (ray.dz - ray.dz);
    const bool regular = chk == 0.0f;  // any NaN/Inf operand makes the sum NaN
    t_out = kMaxFloat;
    face_out = 0xffffffffu;
    if (__any_sync(FULL_MASK, active && !regular))
    {
        trace_warp<STRICT, COUNT>(sc, scratch, warp_slot, lane, active, ray, leaves, stack, t_out, face_out, wc);
        return;
    }
    const uint32_t mypx = (uint32_t)px_x | ((uint32_t)px_y << 16);
    bool part = false;   // this lane's ray takes part in the flat search
    if (active)
    {
        const float4 lo = __ldg(&sc.nodes[0]);
        const float4 hi = __ldg(&sc.nodes[1]);
        if (COUNT) wc.box++;
        part = box_check(ray, ix, iy, iz, lo, hi) && px_contains(__ldg(&vw.node_px[0]), mypx);  // kd_tree.cpp:339
    }
    if (!__any_sync(FULL_MASK, part)) return;
    // pixel range of the rays that can have work
    uint2 tr;
    {
        const uint32_t x0 = __reduce_min_sync(FULL_MASK, part ? (uint32_t)px_x : 0xffffu), y0 = __reduce_min_sync(FULL_MASK, part ? (uint32_t)px_y : 0xffffu);
        const uint32_t x1 = __reduce_max_sync(FULL_MASK, part ? (uint32_t)px_x : 0u), y1 = __reduce_max_sync(FULL_MASK, part ? (uint32_t)px_y : 0u);
        tr = make_uint2(x0 | (y0 << 16), x1 | (y1 << 16));
    }
    wp.key[lane] = ~0ull;
    __syncwarp();
    if (sc.root_is_leaf)  // kd_tree.cpp:344-361: the root's primitives are scanned without an entry-distance filter
    {
        const uint32_t first = ~(uint32_t)__float_as_int(__ldg(&sc.nodes[0]).w);
        const uint32_t cnt = (uint32_t)__float_as_int(__ldg(&sc.nodes[1]).w);
        if (COUNT && part) { wc.listed++; wc.scanned++; }
        scan_leaf_flat<STRICT, COUNT>(sc, vw, wp, lane, ray, first, cnt, part, tr, tile_x0, tile_y0, wc);
        const unsigned long long kk = wp.key[lane];
        if (part && kk != ~0ull) { t_out = __uint_as_float((uint32_t)(kk >> 32)); face_out = wp.face[lane]; }
        return;
    }
    const uint32_t bin = (uint32_t)((tile_y0 >> kBinShiftY) * vw.bins_x + (tile_x0 >> kBinShiftX));
    const uint32_t nL = __ldg(&vw.bin_cnt[2 * bin]), nI = __ldg(&vw.bin_cnt[2 * bin + 1]);
    const uint32_t nWI = min(__ldg(&vw.busy_count[kCtlWide]), (uint32_t)kWideCap);   // inner nodes near the root: one list for all bins
    if (nL > vw.bin_leaf_cap || nI > vw.bin_inner_cap)   // a list of this bin did not fit: the tile takes the unculled core
    {
        if (COUNT && active) wc.ovf++;
        // (tried instead, r02: larger lists — slower than this per-ray search from 1024 entries per bin on; candidates by a per-tile
        // tree walk + the pair scan of the flat path — 2.5x slower on the dense views: their triangles cover whole tiles, every
        // triangle becomes 32 pairs and the pair queue's bookkeeping dominates; profiles/r02_dense_views.md)
        trace_warp_culled<STRICT, COUNT>(sc, scratch, vw, wp, warp_slot, lane, part, ray, mypx, leaves, stack, t_out, face_out, wc);
        return;
    }
    // ---- rays with a tie of split-plane parameters at an inner node above a candidate leaf ("risky") ----
    bool risky = false;
    {
        const float4 s = __ldg(&sc.nodes[2 * (uint32_t)__float_as_int(__ldg(&sc.nodes[0]).w) + 1]);   // root split point
        const float tsx = mulr(subr(s.x, ray.ox), ix), tsy = mulr(subr(s.y, ray.oy), iy), tsz = mulr(subr(s.z, ray.oz), iz);
        risky = part && (tsx == tsy || tsy == tsz || tsx == tsz);
    }
    for (uint32_t base = 0; base < nWI + nI; base += 32)
    {
        const uint32_t idx = base + lane;
        const bool valid = idx < nWI + nI;
        uint4 e0 = make_uint4(0xffffffffu, 0u, 0u, 0u), e1 = make_uint4(0u, 0u, 0u, 0u);
        if (valid)
        {
            const uint4* e = idx < nWI ? vw.wide_inner + (size_t)idx * 2 : vw.bin_inner + ((size_t)bin * vw.bin_inner_cap + (idx - nWI)) * 2;
            e0 = __ldg(e); e1 = __ldg(e + 1);
        }
        unsigned m = __ballot_sync(FULL_MASK, valid && px_overlap(make_uint2(e0.x, e0.y), tr));
        while (m)
        {
            const int i = __ffs(m) - 1;
            m &= m - 1;
            const uint2 r = make_uint2(__shfl_sync(FULL_MASK, e0.x, i), __shfl_sync(FULL_MASK, e0.y, i));
            const float sx = __uint_as_float(__shfl_sync(FULL_MASK, e1.x, i)), sy = __uint_as_float(__shfl_sync(FULL_MASK, e1.y, i)),
                        sz = __uint_as_float(__shfl_sync(FULL_MASK, e1.z, i));
            const float tsx = mulr(subr(sx, ray.ox), ix), tsy = mulr(subr(sy, ray.oy), iy), tsz = mulr(subr(sz, ray.oz), iz);
            if (COUNT && part) wc.box++;
            risky = risky || (part && px_contains(r, mypx) && (tsx == tsy || tsy == tsz || tsx == tsz));
        }
    }
    const bool any_risky = __any_sync(FULL_MASK, risky);
    if (risky) part = false;
    // ---- candidate leaves of the bin ----
    uint32_t best_dis = 0xffffffffu, best_seq = 0xffffffffu;   // key of the best leaf with a hit so far (entry distance bits, leaf_seq)
    float bt = kMaxFloat;
    uint32_t bface = 0xffffffffu;
    // reference ray of the ordering heuristic: the first participating lane's
    const int ref = __ffs(__ballot_sync(FULL_MASK, part)) - 1;
    if (ref >= 0)
    {
        const float rdx = __shfl_sync(FULL_MASK, ix, ref), rdy = __shfl_sync(FULL_MASK, iy, ref), rdz = __shfl_sync(FULL_MASK, iz, ref);
        for (uint32_t base = 0; base < nL; base += 32)
        {
            const bool valid = base + lane < nL;
            uint4 e0 = make_uint4(0xffffffffu, 0u, 0u, 0u), e1 = make_uint4(0u, 0u, 0u, 0u), e2 = e1;
            if (valid)
            {
                const uint4* e = vw.bin_leaf + ((size_t)bin * vw.bin_leaf_cap + base + lane) * 3;
                e0 = __ldg(e); e1 = __ldg(e + 1); e2 = __ldg(e + 2);
            }
            const bool ov = valid && px_overlap(make_uint2(e0.x, e0.y), tr);
            // ordering key: where the reference ray enters this lane's candidate box (max of the three near-slab parameters,
            // clamped at 0); only the processing order depends on it
            uint32_t okey = 0xffffffffu;
            if (ov)
            {
                const float nx = rdx < 0.0f ? __uint_as_float(e2.x) : __uint_as_float(e1.x), ny = rdy < 0.0f ? __uint_as_float(e2.y) : __uint_as_float(e1.y),
                            nz = rdz < 0.0f ? __uint_as_float(e2.z) : __uint_as_float(e1.z);
                const float tn = fmaxf(fmaxf((nx - ray.ox) * rdx, (ny - ray.oy) * rdy), fmaxf((nz - ray.oz) * rdz, 0.0f));
                okey = (__float_as_uint(tn) & ~31u) | (uint32_t)lane;   // tn >= 0 (NaN → 0 by fmaxf): bits order like values
            }
            unsigned rem = __ballot_sync(FULL_MASK, ov);
            while (rem)
            {
                const uint32_t mn = __reduce_min_sync(FULL_MASK, ((rem >> lane) & 1u) ? okey : 0xffffffffu);
                const int i = (int)(mn & 31u);
                rem &= ~(1u << i);
                const uint2 r = make_uint2(__shfl_sync(FULL_MASK, e0.x, i), __shfl_sync(FULL_MASK, e0.y, i));
                const bool inr = part && px_contains(r, mypx);
                if (!__any_sync(FULL_MASK, inr)) continue;
                float4 lo, hi;
                lo.x = __uint_as_float(__shfl_sync(FULL_MASK, e1.x, i)); lo.y = __uint_as_float(__shfl_sync(FULL_MASK, e1.y, i));
                lo.z = __uint_as_float(__shfl_sync(FULL_MASK, e1.z, i)); hi.x = __uint_as_float(__shfl_sync(FULL_MASK, e2.x, i));
                hi.y = __uint_as_float(__shfl_sync(FULL_MASK, e2.y, i)); hi.z = __uint_as_float(__shfl_sync(FULL_MASK, e2.z, i));
                lo.w = hi.w = 0.f;
                const uint32_t seq = __shfl_sync(FULL_MASK, e0.z, i);
                const float dis = box_entry(ray, ix, iy, iz, lo, hi);   // get_ray_AABB_intersection, aabb.h:29-63
                const uint32_t db = __float_as_uint(dis);
                const bool listed = inr && dis > 0.0f;                  // kd_tree.cpp:385-388
                if (COUNT && inr) { wc.box++; wc.listed += listed ? 1u : 0u; }
                const bool ing = listed && (db < best_dis || (db == best_dis && seq < best_seq));
                if (!__any_sync(FULL_MASK, ing)) continue;
                const uint32_t first = __shfl_sync(FULL_MASK, e0.w, i), cnt = __shfl_sync(FULL_MASK, e1.w, i);
                if (COUNT && ing) wc.scanned++;
                scan_leaf_flat<STRICT, COUNT>(sc, vw, wp, lane, ray, first, cnt, ing, tr, tile_x0, tile_y0, wc);
                const unsigned long long kk = wp.key[lane];
                if (ing && kk != ~0ull)
                {
                    best_dis = db; best_seq = seq;
                    bt = __uint_as_float((uint32_t)(kk >> 32));
                    bface = wp.face[lane];
                    wp.key[lane] = ~0ull;
                }
                __syncwarp();
            }
        }
    }
    t_out = bt;
    face_out = bface;
    if (any_risky)   // exact walk for the rays that might be cut (kd_tree.cpp:374)
    {
        float t2 = kMaxFloat;
        uint32_t f2 = 0xffffffffu;
        trace_warp<STRICT, COUNT>(sc, scratch, warp_slot, lane, risky, ray, leaves, stack, t2, f2, wc);
        if (risky) { t_out = t2; face_out = f2; }
    }
}

// ---- K1c: fused primary-ray render with film-space culling ----------------------------------------------------------
// The rank's tiles are processed slab by slab (DevSlabs: contiguous ranges of whole bands, bottom of the image first).
// Phase A of a slab: every warp takes a static, interleaved share of the slab's tiles and writes the sky result into
// those no triangle rectangle can touch (≈90 % of a typical frame: a pure store stream). Phase B: the slab's busy tiles
// listed by k_view_finish are pulled one at a time from an atomic cursor (dynamic balance where the work is). A warp
// that runs out of work in a slab publishes the number of tiles it completed there (stores fenced first) and moves on:
// ctl[kCtlDone + s] == tiles of slab s ⇔ every scanline of the slab is final — the host side's copy stream waits for
// exactly that value (stream memory operation) and copies the slab out while later slabs are still being traced.
// MINB = resident CTAs per SM the kernel is compiled for: 3 (80 registers) or 4 (64 registers, more spills, more warps). With many
// busy tiles per warp (one or two ranks on a 4K frame) the extra warps win (render kernel 254 -> 240 us on the stand-in); when a
// rank's kernel is a single wave of tiles (4 or 8 ranks) the faster warps do (world 8: 74 against 79 us) — chosen per launch.
template <bool STRICT, bool COUNT, int MINB>
__global__ void __launch_bounds__(kBlock, MINB) k_render_culled(DevScene sc, DevCamera cm, DevPartition part, DevOut out, DevScratch scratch,
                                                          DevView vw, DevSlabs slabs, DevCounters* counters,
                                                          unsigned long long jitter_seed)
{
    const int lane = threadIdx.x & 31;
    const int warp_slot = (blockIdx.x * kBlock + threadIdx.x) >> 5;
    const unsigned int n_warps = (gridDim.x * kBlock) >> 5;
    const int tiles_x = (cm.W + 7) >> 3;
    const int rows_per_band = part.band_rows >> 2;
    const int n_bands = (cm.H + part.band_rows - 1) / part.band_rows;
    const int my_bands = part.rank < n_bands ? (n_bands - part.rank + part.world - 1) / part.world : 0;
    const unsigned int tiles_per_band = (unsigned int)(tiles_x * rows_per_band);
    const unsigned int n_tiles = (unsigned int)my_bands * tiles_per_band;
    const bool all_busy = __ldg(vw.all_flag) == vw.frame_id;
    __shared__ WarpScratch wscr[kBlock / 32];
    WorkCount wc;
    wc.clear();
    uint2 leaves[kLeafCap];    // per-lane sorted leaf list (local memory)
    uint32_t stack[kStackCap]; // per-lane inner-node stack of the unculled traversal
    unsigned int rays_done = 0;
    // sky pixel: cast_ray with no hit adds materials[0].emission per sample (renderer.cpp:226, 353-358)
    float sr = 0.0f, sg = 0.0f, sb = 0.0f;
    for (uint32_t s = 0; s < cm.spp; ++s) { sr = addr(sr, sc.sky[0]); sg = addr(sg, sc.sky[1]); sb = addr(sb, sc.sky[2]); }
    const float nspp = (float)cm.spp;
    const uint32_t sky_px = pack_bgrx(__fdiv_rn(sr, nspp), __fdiv_rn(sg, nspp), __fdiv_rn(sb, nspp));
    // 16-byte row segments need W % 4 == 0 and 16-byte aligned planes (caller-provided device pointers may be anything)
    const bool vec_ok = (cm.W & 3) == 0 &&
                        (((size_t)out.bgrx | (size_t)out.t | (size_t)out.prim) & 15) == 0;

  for (int slab = 0; slab < slabs.n; ++slab)
  {
    const unsigned int tile_lo = slabs.tile_begin[slab], tile_hi = min(slabs.tile_begin[slab + 1], n_tiles);
    unsigned int finished = 0;  // tiles of this slab completed by this warp
    bool cancelled = false;
    unsigned int cancel_word = 0u;
    // ---- phase A: sky tiles ------------------------------------------------------------------------------------------
    // One strip of 32 consecutive tiles (256×4 pixels) per step: one coalesced stamp load, then 512-byte row segments.
    if (!all_busy && vec_ok)
    {
        const unsigned int strips_x = (unsigned int)(tiles_x + 31) >> 5;
        const unsigned int trow_lo = tile_lo / (unsigned int)tiles_x, trow_hi = tile_hi / (unsigned int)tiles_x;
        const unsigned int n_strips = (trow_hi - trow_lo) * strips_x;
        const uint4 px4 = make_uint4(sky_px, sky_px, sky_px, sky_px);
        const uint4 t4 = make_uint4(__float_as_uint(kMaxFloat), __float_as_uint(kMaxFloat), __float_as_uint(kMaxFloat), __float_as_uint(kMaxFloat));
        const uint4 m4 = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
        for (unsigned int st = (unsigned int)warp_slot; st < n_strips; st += n_warps)
        {
            const unsigned int sr = st / strips_x, scol = st - sr * strips_x;
            const unsigned int ltr = trow_lo + sr;                                   // local tile row
            const unsigned int lb = ltr / (unsigned int)rows_per_band, trow = ltr - lb * (unsigned int)rows_per_band;
            const int y0 = ((int)lb * part.world + part.rank) * part.band_rows + (int)trow * 4;
            if (y0 >= cm.H) continue;  // tile rows past the image in the last band (not counted in live_tiles either)
            const int tcol = (int)scol * 32 + lane;
            const bool sky = tcol < tiles_x && __ldg(&vw.stamps[(y0 >> 2) * vw.tiles_x + tcol]) != vw.frame_id;
            const unsigned skym = __ballot_sync(FULL_MASK, sky);
            if (!skym) continue;
            finished += (unsigned int)__popc(skym);
            if (COUNT && lane == 0) wc.tsky += (unsigned int)__popc(skym);
#pragma unroll
            for (int h = 0; h < 2; ++h)
            {
                const int x = (int)scol * 256 + h * 128 + lane * 4;                  // 4 pixels = half a tile row
                if (!((skym >> (h * 16 + (lane >> 1))) & 1u) || x >= cm.W) continue;
#pragma unroll
                for (int r = 0; r < 4; ++r)
                {
                    const int y = y0 + r;
                    if (y >= cm.H) break;
                    const size_t idx = (size_t)y * (size_t)cm.W + (size_t)x;
                    if (out.bgrx) *reinterpret_cast<uint4*>(out.bgrx + idx) = px4;
                    if (out.t) *reinterpret_cast<uint4*>(out.t + idx) = t4;
                    if (out.prim) *reinterpret_cast<uint4*>(out.prim + idx) = m4;
                    rays_done += 4u * cm.spp;
                }
            }
        }
    }
    else if (!all_busy)
    {
        for (unsigned int tile = tile_lo + (unsigned int)warp_slot; tile < tile_hi; tile += n_warps)
        {
            const unsigned int lb = tile / tiles_per_band, rem = tile - lb * tiles_per_band;
            const int trow = (int)(rem / (unsigned int)tiles_x), tcol = (int)(rem - (unsigned int)trow * (unsigned int)tiles_x);
            const int y0 = ((int)lb * part.world + part.rank) * part.band_rows + trow * 4;
            if (y0 >= cm.H) continue;  // tile rows past the image in the last band (not counted in live_tiles either)
            if (__ldg(&vw.stamps[(y0 >> 2) * vw.tiles_x + tcol]) == vw.frame_id) continue;  // busy: phase B
            const int x = tcol * 8 + (lane & 7), y = y0 + (lane >> 3);
            if (x < cm.W && y < cm.H)
            {
                const size_t idx = (size_t)y * (size_t)cm.W + (size_t)x;
                if (out.bgrx) out.bgrx[idx] = sky_px;
                if (out.t) out.t[idx] = kMaxFloat;
                if (out.prim) out.prim[idx] = 0xffffffffu;
                rays_done += cm.spp;
            }
            finished += 1u;
            if (COUNT && lane == 0) wc.tsky++;
        }
    }
    // ---- phase B: busy tiles -----------------------------------------------------------------------------------------
    // The slab's list is walked most expensive cost class first: lane r < kCostBuckets holds the length of class
    // kCostBuckets-1-r (re-read per tile: two registers less across the trace) and where it starts in that order.
    for (;;)
    {
        unsigned int bi = 0;
        if (lane == 0) bi = atomicAdd(&vw.busy_count[kCtlCursor + slab], 1u);
        bi = __shfl_sync(FULL_MASK, bi, 0);
        // the context's cancel word, read by lane 31 one tile ahead: the (uncached) load is in flight while the tile is traced
#ifndef ATRAY_CULLED_NO_CANCEL   // (A/B builds only)
        if (__shfl_sync(FULL_MASK, cancel_word, 31)) { cancelled = true; break; }
        if (lane == 31) cancel_word = (unsigned int)cancel_requested(sc);
#endif
        const unsigned int cnt = lane < kCostBuckets ? __ldg(&vw.busy_count[kCtlCount + slab * kCostBuckets + (kCostBuckets - 1 - lane)]) : 0u;
        unsigned int end = cnt;   // inclusive prefix sum over the lanes
#pragma unroll
        for (int off = 1; off < kCostBuckets; off <<= 1)
        {
            const unsigned int v = __shfl_up_sync(FULL_MASK, end, off);
            if (lane >= off) end += v;
        }
        const unsigned hit = __ballot_sync(FULL_MASK, lane < kCostBuckets && bi < end);
        if (!hit) break;   // past the slab's last busy tile
        const int r = __ffs(hit) - 1;
        const unsigned int start = __shfl_sync(FULL_MASK, end - cnt, r);
        const unsigned int tile = __ldg(&vw.busy_list[(unsigned int)(kCostBuckets - 1 - r) * vw.bucket_stride + tile_lo + (bi - start)]);
        unsigned long long log_t0 = 0;
        if (vw.tile_log) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(log_t0));
        finished += 1u;
        const unsigned int lb = tile / tiles_per_band, rem = tile - lb * tiles_per_band;
        const int trow = (int)(rem / (unsigned int)tiles_x), tcol = (int)(rem - (unsigned int)trow * (unsigned int)tiles_x);
        const int x = tcol * 8 + (lane & 7);
        const int y = ((int)lb * part.world + part.rank) * part.band_rows + trow * 4 + (lane >> 3);
        const bool active = x < cm.W && y < cm.H;
        const int tile_y0 = y - (lane >> 3);
        const size_t idx = (size_t)y * (size_t)cm.W + (size_t)x;
        if (COUNT && lane == 0) wc.tbusy++;
        const float film_y = film_coord(y, cm.H);
        const float film_x = mulr(mulr(film_coord(x, cm.W), cm.h_fov), cm.aspect);
        float t = kMaxFloat;
        uint32_t face = 0xffffffffu;
        float cr = 0.0f, cg = 0.0f, cb = 0.0f;
        if (!cm.aa)
        {
            const RayL ray = camera_ray(cm, film_x, film_y);
            trace_tile_flat<STRICT, COUNT>(sc, scratch, vw, wscr[threadIdx.x >> 5], warp_slot, lane, active, ray, x, y, tcol * 8, tile_y0, leaves, stack, t, face, wc);
            const bool hit = face != 0xffffffffu;
            const float er = hit ? sc.mdl[0] : sc.sky[0], eg = hit ? sc.mdl[1] : sc.sky[1], eb = hit ? sc.mdl[2] : sc.sky[2];
            for (uint32_t s = 0; s < cm.spp; ++s) { cr = addr(cr, er); cg = addr(cg, eg); cb = addr(cb, eb); }
            if (active) rays_done += 1;
        }
        else
        {
            const unsigned long long p = (unsigned long long)y * (unsigned long long)cm.W + (unsigned long long)x;
            Pcg rng;
            rng.state = splitmix64(jitter_seed ^ p);
            rng.stream = 2ULL * p + 1ULL;
            for (uint32_t s = 0; s < cm.spp; ++s)
            {
                const float x_off = addr(mulr(pcg_bi(rng), cm.hpw), film_x);
                const float y_off = addr(mulr(pcg_bi(rng), cm.hph), film_y);
                const RayL ray = camera_ray(cm, x_off, y_off);
                trace_tile_flat<STRICT, COUNT>(sc, scratch, vw, wscr[threadIdx.x >> 5], warp_slot, lane, active, ray, x, y, tcol * 8, tile_y0, leaves, stack, t, face, wc);
                const bool hit = face != 0xffffffffu;
                if (hit) { pcg_bi(rng); pcg_bi(rng); pcg_bi(rng); }
                cr = addr(cr, hit ? sc.mdl[0] : sc.sky[0]);
                cg = addr(cg, hit ? sc.mdl[1] : sc.sky[1]);
                cb = addr(cb, hit ? sc.mdl[2] : sc.sky[2]);
                if (active) rays_done += 1;
            }
        }
        if (active)
        {
            if (out.bgrx) out.bgrx[idx] = pack_bgrx(__fdiv_rn(cr, nspp), __fdiv_rn(cg, nspp), __fdiv_rn(cb, nspp));
            if (out.t) out.t[idx] = t;
            if (out.prim) out.prim[idx] = face;
        }
        if (vw.tile_log && lane == 0)
        {
            unsigned long long log_t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(log_t1));
            vw.tile_log[tile].x = (uint32_t)log_t0;
            vw.tile_log[tile].y = (uint32_t)(log_t1 - log_t0);
        }
    }
    // publish this warp's share of the slab: every lane fences its own stores, then one counter update per warp
    if (__any_sync(FULL_MASK, finished != 0u))
    {
        __threadfence();
        __syncwarp();
        if (lane == 0) atomicAdd(&vw.busy_count[kCtlDone + slab], finished);
    }
    if (cancelled) break;
  }
    if (COUNT) flush_counters(counters, wc, rays_done);
}

#include "raster.cuh"

// ---- K2: ray-list trace (get_intersection_data model branch for arbitrary rays) ---------------------------------
// UV: also store the winning triangle's barycentric (u, v) (TriangleIntersectionData, model.h:18-22) at uv[2i], uv[2i+1]; a miss
// leaves them untouched, as get_ray_kd_tree_intersection leaves tri_data (kd_tree.cpp:436-462).
template <bool STRICT, bool COUNT, bool UV = false>
__global__ void __launch_bounds__(kBlock) k_trace(DevScene sc, const float* __restrict__ rays, unsigned long long n, uint32_t* prim,
                                                  float* tout, DevScratch scratch, unsigned int* tile_counter, DevCounters* counters,
                                                  float* uv = nullptr)
{
    const int lane = threadIdx.x & 31;
    const int warp_slot = (blockIdx.x * kBlock + threadIdx.x) >> 5;
    const unsigned long long n_groups = (n + 31ULL) >> 5;
    WorkCount wc;
    wc.clear();
    uint2 leaves[kLeafCap];    // per-lane sorted leaf list (local memory)
    uint32_t stack[kStackCap]; // per-lane inner-node stack of the unculled traversal
    unsigned int rays_done = 0;
    for (;;)
    {
        unsigned int g = 0;
        if (lane == 0) g = cancel_requested(sc) ? 0xffffffffu : atomicAdd(tile_counter, 1u);
        g = __shfl_sync(FULL_MASK, g, 0);
        if ((unsigned long long)g >= n_groups) break;
        const unsigned long long i = ((unsigned long long)g << 5) + (unsigned long long)lane;
        const bool active = i < n;
        RayL ray = {0.f, 0.f, 0.f, 0.f, 0.f, 1.f};
        if (active)
        {
            const float2* q = reinterpret_cast<const float2*>(rays + 6ULL * i);
            const float2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            ray.ox = a.x; ray.oy = a.y; ray.oz = b.x; ray.dx = b.y; ray.dy = c.x; ray.dz = c.y;
        }
        float t;
        uint32_t face;
        float hu = 0.0f, hv = 0.0f;
        trace_warp<STRICT, COUNT, UV>(sc, scratch, warp_slot, lane, active, ray, leaves, stack, t, face, wc, &hu, &hv);
        if (active)
        {
            prim[i] = face;
            tout[i] = t;
            if (UV && face != 0xffffffffu) { uv[2ULL * i] = hu; uv[2ULL * i + 1ULL] = hv; }
            rays_done += 1;
        }
    }
    if (COUNT) flush_counters(counters, wc, rays_done);
}

// ---- K2b: C5, incoherent random-direction rays (BASELINE.json configs[4]) -------------------------------------
// `spp` rays per pixel index p, generated on the device from the per-ray PCG32 stream {splitmix64(seed ^ i), 2i + 1}, i = p*spp + s
// (six rand_bi draws: a direction for the origin on a sphere around the model box, a target inside the box; SetRay, ray.h:24-29)
// (include/atray_b200.h: atray_trace_incoherent_device states the definition), traced through the model query of get_intersection_data
// (renderer.cpp:41-57) on the warp-cooperative core, and reduced to the pixel's flat colour: sequential f32 sum of the
// samples' emissions / spp, clamp, truncate, pack (renderer.cpp:358-365). One warp per pixel, 32 samples at a time. The
// pixels of a rank are contiguous chunks of `chunk_px` pixels dealt round-robin (chunk k -> rank k % world).
struct DevC5
{
    unsigned long long p0, p1;      // pixel indices [p0, p1) of the whole job
    uint32_t spp, chunk_px;
    unsigned long long seed;
    uint32_t* bgrx;                 // per pixel, indexed p - p0 (may be a peer pointer: GPU 0's plane)
    uint32_t* prim;                 // per ray, indexed (p - p0)*spp + s, or null
    float* t;                       // likewise
};

template <bool STRICT, bool COUNT>
__global__ void __launch_bounds__(kBlock) k_c5(DevScene sc, DevC5 job, DevPartition part, DevScratch scratch, unsigned int* counter,
                                               DevCounters* counters)
{
    const int lane = threadIdx.x & 31;
    const int warp_slot = (blockIdx.x * kBlock + threadIdx.x) >> 5;
    WorkCount wc;
    wc.clear();
    uint2 leaves[kLeafCap];
    uint32_t stack[kStackCap];
    unsigned int rays_done = 0;
    // box centre, half extent and radius of the origin sphere: c = (lo + hi)*0.5, extent = hi - lo, R = 1.5*mag(extent)
    const float cx = mulr(addr(sc.aabb_lo[0], sc.aabb_hi[0]), 0.5f), cy = mulr(addr(sc.aabb_lo[1], sc.aabb_hi[1]), 0.5f),
                cz = mulr(addr(sc.aabb_lo[2], sc.aabb_hi[2]), 0.5f);
    const float ex = subr(sc.aabb_hi[0], sc.aabb_lo[0]), ey = subr(sc.aabb_hi[1], sc.aabb_lo[1]), ez = subr(sc.aabb_hi[2], sc.aabb_lo[2]);
    const float R = mulr(1.5f, __fsqrt_rn(addr(addr(mulr(ex, ex), mulr(ey, ey)), mulr(ez, ez))));
    const float hx = mulr(ex, 0.5f), hy = mulr(ey, 0.5f), hz = mulr(ez, 0.5f);
    // pixels of this rank: local index j -> chunk j / chunk_px of this rank = global chunk (j / chunk_px)*world + rank
    const unsigned long long n_px = job.p1 - job.p0;
    const unsigned long long n_chunks = (n_px + job.chunk_px - 1) / job.chunk_px;
    const unsigned long long my_chunks = (unsigned long long)part.rank < n_chunks ? (n_chunks - part.rank + part.world - 1) / part.world : 0ull;
    for (;;)
    {
        unsigned int j = 0;
        if (lane == 0) j = cancel_requested(sc) ? 0xffffffffu : atomicAdd(counter, 1u);
        j = __shfl_sync(FULL_MASK, j, 0);
        const unsigned long long lc = j / job.chunk_px;
        if (lc >= my_chunks || j == 0xffffffffu) break;
        const unsigned long long rel = (lc * part.world + part.rank) * job.chunk_px + (j - lc * job.chunk_px);   // pixel - p0
        if (rel >= n_px) continue;
        const unsigned long long p = job.p0 + rel;
        float cr = 0.0f, cg = 0.0f, cb = 0.0f;   // lane 0: the pixel's running sum
        for (uint32_t s0 = 0; s0 < job.spp; s0 += 32)
        {
            const uint32_t s = s0 + lane;
            const bool active = s < job.spp;
            const unsigned long long i = p * job.spp + s;
            Pcg rng;
            rng.state = splitmix64(job.seed ^ i);
            rng.stream = 2ULL * i + 1ULL;
            const float b0 = pcg_bi(rng), b1 = pcg_bi(rng), b2 = pcg_bi(rng), q0 = pcg_bi(rng), q1 = pcg_bi(rng), q2 = pcg_bi(rng);
            const float binv = __fdiv_rn(1.0f, __fsqrt_rn(addr(addr(mulr(b0, b0), mulr(b1, b1)), mulr(b2, b2))));
            RayL ray;
            ray.ox = addr(cx, mulr(mulr(b0, binv), R)); ray.oy = addr(cy, mulr(mulr(b1, binv), R)); ray.oz = addr(cz, mulr(mulr(b2, binv), R));
            const float tx = addr(cx, mulr(hx, q0)), ty = addr(cy, mulr(hy, q1)), tz = addr(cz, mulr(hz, q2));
            const float dx = subr(tx, ray.ox), dy = subr(ty, ray.oy), dz = subr(tz, ray.oz);
            const float dinv = __fdiv_rn(1.0f, __fsqrt_rn(addr(addr(mulr(dx, dx), mulr(dy, dy)), mulr(dz, dz))));
            ray.dx = mulr(dx, dinv); ray.dy = mulr(dy, dinv); ray.dz = mulr(dz, dinv);
            float t;
            uint32_t face;
            trace_warp<STRICT, COUNT>(sc, scratch, warp_slot, lane, active, ray, leaves, stack, t, face, wc);
            // get_intersection_data accepts the model's distance iff t > tolerance && t < MAX_FLOAT (renderer.cpp:52)
            const bool hit = active && face != 0xffffffffu && t > kTol && t < kMaxFloat;
            if (active)
            {
                if (job.prim) job.prim[rel * job.spp + s] = face;
                if (job.t) job.t[rel * job.spp + s] = t;
                rays_done += 1;
            }
            const unsigned hm = __ballot_sync(FULL_MASK, hit);
            if (lane == 0)
            {
                const uint32_t n = min(32u, job.spp - s0);
                for (uint32_t k = 0; k < n; ++k)   // the samples in order: a sequential f32 sum (renderer.cpp:353-357)
                {
                    const bool h = (hm >> k) & 1u;
                    cr = addr(cr, h ? sc.mdl[0] : sc.sky[0]); cg = addr(cg, h ? sc.mdl[1] : sc.sky[1]); cb = addr(cb, h ? sc.mdl[2] : sc.sky[2]);
                }
            }
        }
        if (lane == 0 && job.bgrx)
        {
            const float n = (float)job.spp;
            job.bgrx[rel] = pack_bgrx(__fdiv_rn(cr, n), __fdiv_rn(cg, n), __fdiv_rn(cb, n));
        }
    }
    if (COUNT) flush_counters(counters, wc, rays_done);
}

// ---- K3: multi-bounce path ------------------------------------------------------------------------
// get_intersection_data in full (renderer.cpp:34-160: model through the tree, then spheres, then planes, hit-type
// priority plane > sphere > model > sky, smooth or flat triangle normal) and cast_ray at any bounce_limit
// (renderer.cpp:213-262). All shading arithmetic is explicit round-to-nearest in the reference's operation order, so the
// frame is bit-identical to the CPU path given the same per-pixel RNG stream. The tree part of every bounce runs on the
// warp-cooperative unculled core (secondary rays are incoherent; the film-space bounds only hold for camera rays).
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 v3add(V3 a, V3 b) { return v3(addr(a.x, b.x), addr(a.y, b.y), addr(a.z, b.z)); }
__device__ __forceinline__ V3 v3sub(V3 a, V3 b) { return v3(subr(a.x, b.x), subr(a.y, b.y), subr(a.z, b.z)); }
__device__ __forceinline__ V3 v3mul(V3 a, float k) { return v3(mulr(a.x, k), mulr(a.y, k), mulr(a.z, k)); }
__device__ __forceinline__ V3 v3had(V3 a, V3 b) { return v3(mulr(a.x, b.x), mulr(a.y, b.y), mulr(a.z, b.z)); }
__device__ __forceinline__ V3 v3neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float v3dot(V3 a, V3 b) { return addr(addr(mulr(a.x, b.x), mulr(a.y, b.y)), mulr(a.z, b.z)); }  // PL_math.h:416
__device__ __forceinline__ V3 v3cross(V3 a, V3 b)                                                                         // PL_math.h:422
{
    return v3(subr(mulr(a.y, b.z), mulr(a.z, b.y)), subr(mulr(a.z, b.x), mulr(a.x, b.z)), subr(mulr(a.x, b.y), mulr(a.y, b.x)));
}
__device__ __forceinline__ V3 v3norm(V3 a)  // normalize, PL_math.h:387-392 with invsqrt := 1/sqrtf
{
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(v3dot(a, a)));
    return v3mul(a, inv);
}
__device__ __forceinline__ V3 f4v(float4 q) { return v3(q.x, q.y, q.z); }

struct HitFull
{
    int kind;        // 0 sky, 1 triangle, 2 sphere, 3 plane (ObjectType, renderer.cpp:12-15)
    float t;         // distance_at_intersection (MAX_FLOAT for the sky)
    V3 n;            // normalised normal
    V3 em, refl;     // hit material
    float scatter;
    uint32_t face;
};

// get_sphere_ray_intersection, sphere.h:12-39
__device__ __forceinline__ float sphere_hit(V3 o, V3 d, float4 cr)
{
    const V3 pc = v3sub(o, f4v(cr));
    const float pc2 = v3dot(pc, pc);
    const float b = mulr(2.0f, v3dot(d, pc));
    const float b2 = mulr(b, b);
    const float c = subr(pc2, mulr(cr.w, cr.w));
    const float dmt = subr(b2, mulr(4.0f, c));
    if (dmt < 0.0f) return 0.0f;
    const float sq = __fsqrt_rn(dmt);
    const float ta = mulr(addr(-b, sq), 0.5f);
    const float tb = mulr(subr(-b, sq), 0.5f);
    if (ta <= 0.0f && tb <= 0.0f) return 0.0f;
    if (tb > 0.0f) return tb;
    return ta;
}
// get_plane_ray_intersection, plane.h:12-22
__device__ __forceinline__ float plane_hit(V3 o, V3 d, float4 nd)
{
    const float denom = v3dot(f4v(nd), d);
    if (denom > -kTol && denom < kTol) return 0.0f;
    return __fdiv_rn(subr(nd.w, v3dot(o, f4v(nd))), denom);
}

// Everything of get_intersection_data after the tree query: t_tri/face/u/v come from trace_warp (MISS when no triangle).
__device__ __forceinline__ void resolve_hit(const DevScene& sc, V3 o, V3 d, float t_tri, uint32_t face, float u, float v, HitFull& h)
{
    float dist = kMaxFloat;
    int near_sphere = -1, near_plane = -1;
    const bool near_model = face != 0xffffffffu && t_tri > kTol && t_tri < dist;                // renderer.cpp:52-57
    if (near_model) dist = t_tri;
    for (uint32_t i = 0; i < sc.n_spheres; ++i)                                                  // :86-96
    {
        const float t = sphere_hit(o, d, __ldg(&sc.spheres[3 * i]));
        if (t > kTol && t < dist) { dist = t; near_sphere = (int)i; }
    }
    for (uint32_t i = 0; i < sc.n_planes; ++i)                                                   // :97-107
    {
        const float t = plane_hit(o, d, __ldg(&sc.planes[3 * i]));
        if (t > kTol && t < dist) { near_plane = (int)i; dist = t; }
    }
    h.t = dist;
    h.face = 0xffffffffu;
    h.n = v3(0.f, 0.f, 0.f);
    if (near_plane >= 0)                                                                         // :111-116
    {
        const float4 a = __ldg(&sc.planes[3 * near_plane]), b = __ldg(&sc.planes[3 * near_plane + 1]), c = __ldg(&sc.planes[3 * near_plane + 2]);
        h.kind = 3; h.n = f4v(a); h.em = f4v(b); h.scatter = b.w; h.refl = f4v(c);
    }
    else if (near_sphere >= 0)                                                                   // :117-122
    {
        const float4 a = __ldg(&sc.spheres[3 * near_sphere]), b = __ldg(&sc.spheres[3 * near_sphere + 1]), c = __ldg(&sc.spheres[3 * near_sphere + 2]);
        h.kind = 2; h.n = v3sub(v3add(o, v3mul(d, dist)), f4v(a)); h.em = f4v(b); h.scatter = b.w; h.refl = f4v(c);
    }
    else if (near_model)                                                                         // :124-148
    {
        h.kind = 1; h.face = face;
        if (sc.smooth)
        {
            const V3 na = f4v(__ldg(&sc.fnorm[3 * (size_t)face])), nb = f4v(__ldg(&sc.fnorm[3 * (size_t)face + 1])),
                     nc = f4v(__ldg(&sc.fnorm[3 * (size_t)face + 2]));
            const float w = subr(subr(1.0f, u), v);
            h.n = v3add(v3add(v3mul(na, w), v3mul(nb, u)), v3mul(nc, v));
        }
        else
            h.n = f4v(__ldg(&sc.fnorm[face]));
        h.em = v3(sc.mdl[0], sc.mdl[1], sc.mdl[2]); h.refl = v3(sc.mdl_refl[0], sc.mdl_refl[1], sc.mdl_refl[2]); h.scatter = sc.mdl_scatter;
    }
    else                                                                                         // :149-155
    {
        h.kind = 0; h.em = v3(sc.sky[0], sc.sky[1], sc.sky[2]); h.refl = v3(0.f, 0.f, 0.f); h.scatter = 0.f;
    }
    h.n = v3norm(h.n);                                                                           // :157
}

template <bool STRICT>
__global__ void __launch_bounds__(kBlock) k_path(DevScene sc, DevCamera cm, DevPartition part, DevOut out, DevScratch scratch,
                                                 unsigned int* tile_counter, unsigned long long* casts_total, unsigned long long seed)
{
    const int lane = threadIdx.x & 31;
    const int warp_slot = (blockIdx.x * kBlock + threadIdx.x) >> 5;
    const int tiles_x = (cm.W + 7) >> 3;
    const int rows_per_band = part.band_rows >> 2;
    const int n_bands = (cm.H + part.band_rows - 1) / part.band_rows;
    const int my_bands = part.rank < n_bands ? (n_bands - part.rank + part.world - 1) / part.world : 0;
    const unsigned int tiles_per_band = (unsigned int)(tiles_x * rows_per_band);
    const unsigned int n_tiles = (unsigned int)my_bands * tiles_per_band;
    WorkCount wc;
    wc.clear();
    uint2 leaves[kLeafCap];
    uint32_t stack[kStackCap];
    unsigned long long casts = 0;  // ray_casts: bounce iterations that did not end in the sky (renderer.cpp:260)
    for (;;)
    {
        unsigned int tile = 0;
        if (lane == 0) tile = cancel_requested(sc) ? 0xffffffffu : atomicAdd(tile_counter, 1u);
        tile = __shfl_sync(FULL_MASK, tile, 0);
        if (tile >= n_tiles) break;
        const unsigned int lb = tile / tiles_per_band, rem = tile - lb * tiles_per_band;
        const int trow = (int)(rem / (unsigned int)tiles_x), tcol = (int)(rem - (unsigned int)trow * (unsigned int)tiles_x);
        const int x = tcol * 8 + (lane & 7);
        const int y = ((int)lb * part.world + part.rank) * part.band_rows + trow * 4 + (lane >> 3);
        const bool active = x < cm.W && y < cm.H;
        const float film_y = film_coord(y, cm.H);
        const float film_x = mulr(mulr(film_coord(x, cm.W), cm.h_fov), cm.aspect);
        const unsigned long long p = (unsigned long long)y * (unsigned long long)cm.W + (unsigned long long)x;
        Pcg rng;  // harness-defined per-pixel stream (the reference seeds per thread from rdtsc, renderer.cpp:376-378)
        rng.state = splitmix64(seed ^ p);
        rng.stream = 2ULL * p + 1ULL;
        V3 col = v3(0.f, 0.f, 0.f);
        float t_first = kMaxFloat;
        uint32_t face_first = 0xffffffffu;
        for (uint32_t s = 0; s < cm.spp; ++s)  // renderer.cpp:336-356: AA on = a jittered ray per sample, off = the same ray
        {
            float fx = film_x, fy = film_y;
            if (cm.aa)
            {
                fx = addr(mulr(pcg_bi(rng), cm.hpw), film_x);
                fy = addr(mulr(pcg_bi(rng), cm.hph), film_y);
            }
            const RayL pr = camera_ray(cm, fx, fy);
            V3 o = v3(pr.ox, pr.oy, pr.oz), d = v3(pr.dx, pr.dy, pr.dz);
            V3 color = v3(0.f, 0.f, 0.f), weight = v3(1.0f, 1.0f, 1.0f);
            bool alive = active;
            for (int32_t i = 0; i < cm.bounce; ++i)  // cast_ray, renderer.cpp:213-262
            {
                if (!__any_sync(FULL_MASK, alive)) break;
                RayL r;
                r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z;
                float t_tri, u = 0.0f, v = 0.0f;
                uint32_t face;
                trace_warp<STRICT, false, true>(sc, scratch, warp_slot, lane, alive, r, leaves, stack, t_tri, face, wc, &u, &v);
                if (!alive) continue;
                HitFull h;
                resolve_hit(sc, o, d, t_tri, face, u, v, h);
                if (i == 0) { t_first = h.t; face_first = h.face; }
                if (h.kind == 0)
                {
                    color = v3add(color, v3had(weight, h.em));  // :224-228
                    alive = false;
                    continue;
                }
                casts += 1;
                float att = v3dot(v3neg(d), h.n);
                if (att < 0.0f) { h.n = v3neg(h.n); att = 0.0f; }
                V3 pure = v3sub(d, v3mul(h.n, mulr(2.0f, v3dot(d, h.n))));  // :238-243
                pure = v3norm(pure);
                V3 rb;
                rb.x = pcg_bi(rng); rb.y = pcg_bi(rng); rb.z = pcg_bi(rng);   // :246
                rb = v3norm(v3add(rb, h.n));
                o = v3add(o, v3mul(d, h.t));                                   // :251
                d = v3norm(v3add(v3mul(v3sub(pure, rb), h.scatter), rb));      // lerp, PL_math.h:439-442
                color = v3add(color, v3had(weight, h.em));                     // :257-258
                weight = v3had(weight, v3mul(h.refl, att));
            }
            col = v3add(col, color);
        }
        if (active)
        {
            const float n = (float)cm.spp;
            const size_t idx = (size_t)p;
            if (out.bgrx) out.bgrx[idx] = pack_bgrx(__fdiv_rn(col.x, n), __fdiv_rn(col.y, n), __fdiv_rn(col.z, n));
            if (out.t) out.t[idx] = t_first;
            if (out.prim) out.prim[idx] = face_first;
        }
    }
    if (casts_total)
    {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) casts += __shfl_xor_sync(FULL_MASK, casts, off);
        if (lane == 0 && casts) atomicAdd(casts_total, casts);
    }
}

// ---- K0: microbenchmarks for the roofline denominators ------------------------------------------------------------
__global__ void k_fp32_issue(float* sink, int iters)
{
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; ++i)
    {
#pragma unroll
        for (int u = 0; u < 8; ++u)
        {
            a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
        }
    }
    const float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123456.789f) sink[0] = s;
}
__global__ void k_l2_read(const float4* __restrict__ buf, size_t n_vec, int passes, float* sink)
{
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int p = 0; p < passes; ++p)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride)
        {
            const float4 v = __ldcg(buf + i);  // L2 only (bypass L1) so the figure is an L2 read rate
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 123456.789f) sink[0] = acc;
}

// ---- launchers ------------------------------------------------------------------------------------------------------
LaunchCfg render_launch_cfg(int sm_count)
{
    LaunchCfg c;
    c.block = kBlock;
    c.grid = sm_count * 4;  // persistent: 4 resident CTAs of 8 warps per SM (B200: 148 SMs → 592 CTAs)
    return c;
}
int render_warps_in_flight(LaunchCfg cfg) { return cfg.grid * (cfg.block / 32); }

cudaError_t launch_render(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                          const DevScratch& scratch, unsigned int* tile_counter, DevCounters* counters, int fp_fast,
                          unsigned long long jitter_seed, LaunchCfg cfg, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(tile_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if (counters)
    {
        if (fp_fast) k_render<false, true><<<cfg.grid, cfg.block, 0, stream>>>(sc, cam, part, out, scratch, tile_counter, counters, jitter_seed);
        else k_render<true, true><<<cfg.grid, cfg.block, 0, stream>>>(sc, cam, part, out, scratch, tile_counter, counters, jitter_seed);
    }
    else
    {
        if (fp_fast) k_render<false, false><<<cfg.grid, cfg.block, 0, stream>>>(sc, cam, part, out, scratch, tile_counter, counters, jitter_seed);
        else k_render<true, false><<<cfg.grid, cfg.block, 0, stream>>>(sc, cam, part, out, scratch, tile_counter, counters, jitter_seed);
    }
    return cudaGetLastError();
}

DevSlabs make_slabs(const DevCamera& cam, const DevPartition& part, int n_slabs, const int* weights)
{
    DevSlabs s{};
    const int tiles_x = (cam.W + 7) >> 3;
    const int n_bands = (cam.H + part.band_rows - 1) / part.band_rows;
    const int my_bands = part.rank < n_bands ? (n_bands - part.rank + part.world - 1) / part.world : 0;
    const uint32_t tiles_per_band = (uint32_t)(tiles_x * (part.band_rows >> 2));
    int n = n_slabs < 1 ? 1 : (n_slabs > kMaxSlabs ? kMaxSlabs : n_slabs);
    if (n > my_bands) n = my_bands < 1 ? 1 : my_bands;
    s.n = n;
    // slab k covers the bands [band_of[k], band_of[k+1]): proportional to weights[k] (uniform without), at least one band
    int band_of[kMaxSlabs + 1];
    long long w_total = 0, w_acc = 0;
    for (int k = 0; k < n; ++k) w_total += weights && weights[k] > 0 ? weights[k] : 1;
    band_of[0] = 0;
    for (int k = 1; k <= kMaxSlabs; ++k)
    {
        if (k >= n) { band_of[k] = my_bands; continue; }
        w_acc += weights && weights[k - 1] > 0 ? weights[k - 1] : 1;
        int b = (int)(w_acc * my_bands / w_total);
        if (b <= band_of[k - 1]) b = band_of[k - 1] + 1;
        if (b > my_bands - (n - k)) b = my_bands - (n - k);   // leave a band for each later slab
        band_of[k] = b;
    }
    for (int k = 0; k <= kMaxSlabs; ++k) s.tile_begin[k] = (uint32_t)band_of[k] * tiles_per_band;
    for (int k = 0; k < kMaxSlabs; ++k)
    {
        uint32_t live = 0;
        for (int lb = band_of[k]; lb < band_of[k + 1]; ++lb)
        {
            const int y0 = (lb * part.world + part.rank) * part.band_rows;
            int rows = cam.H - y0;
            if (rows > part.band_rows) rows = part.band_rows;
            if (rows > 0) live += (uint32_t)(((rows + 3) >> 2) * tiles_x);
        }
        s.live_tiles[k] = live;
    }
    return s;
}

// Safety net behind the per-slab counters: once the render kernel has finished, every slab is final by definition.
__global__ void k_slabs_final(uint32_t* ctl, DevSlabs slabs)
{
    if (threadIdx.x < (unsigned)slabs.n) ctl[kCtlDone + threadIdx.x] = slabs.live_tiles[threadIdx.x];
}
cudaError_t launch_slabs_final(const DevView& view, const DevSlabs& slabs, cudaStream_t stream)
{
    k_slabs_final<<<1, 32, 0, stream>>>(view.busy_count, slabs);
    return cudaGetLastError();
}

struct SignalArgs { uint32_t* flags[kMaxPeers]; };
__global__ void k_view_signal_v(SignalArgs a, int n, int rank, uint32_t seq)
{
    if ((int)threadIdx.x < n)
    {
        __threadfence_system();
        *((volatile uint32_t*)(a.flags[threadIdx.x] + rank)) = seq;
    }
}
cudaError_t launch_view_signal(uint32_t* const* peer_flags, int n, int rank, uint32_t seq, cudaStream_t stream)
{
    SignalArgs a{};
    for (int i = 0; i < n && i < kMaxPeers; ++i) a.flags[i] = peer_flags[i];
    k_view_signal_v<<<1, 32, 0, stream>>>(a, n, rank, seq);
    return cudaGetLastError();
}
// One warp: lane q waits until rank q's flag in THIS GPU's flag row has reached `seq` (written by the peer GPU's
// k_view_signal_v over NVLink). The ranks run on different GPUs, so the wait never depends on a kernel of the same GPU. A
// peer that never signals (a dead rank) ends the wait after `timeout_ns` with *timed_out = seq instead of hanging the GPU.
__global__ void k_view_wait(const uint32_t* my_flags, int n, int rank, uint32_t seq, unsigned long long timeout_ns, uint32_t* timed_out)
{
    const int q = threadIdx.x;
    if (q >= n || q == rank) return;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    const volatile uint32_t* f = my_flags + q;
    for (;;)
    {
        if ((int32_t)(*f - seq) >= 0) break;
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > timeout_ns) { *timed_out = seq; break; }
        __nanosleep(100);
    }
    __threadfence_system();
}
cudaError_t launch_view_wait(const uint32_t* my_flags, int n, int rank, uint32_t seq, unsigned long long timeout_ns, uint32_t* timed_out,
                             cudaStream_t stream)
{
    k_view_wait<<<1, 32, 0, stream>>>(my_flags, n, rank, seq, timeout_ns, timed_out);
    return cudaGetLastError();
}

cudaError_t launch_view_bounds(const DevScene& sc, const ViewParams& vp, const DevView& view, const DevCamera& cam,
                               const DevPartition& part, const DevSlabs& slabs, cudaStream_t stream, cudaEvent_t after_reset,
                               bool finish, const ViewPeers* peers, const ViewWait* wait, cudaEvent_t after_bounds)
{
    ViewPeers pe{};
    if (peers) pe = *peers;
    else
    {
        pe.n = 1;
        pe.block_lo = 0;
        pe.block_hi = (sc.n_slots + kBlock - 1) / kBlock;
        pe.tri_px[0] = view.tri_px; pe.pkt_px[0] = view.pkt_px; pe.chunk_px[0] = view.chunk_px;
        pe.stamps[0] = view.stamps; pe.tile_cost[0] = view.tile_cost; pe.all_flag[0] = view.all_flag;
        if (const char* env = std::getenv("ATRAY_VIEW_SKIP")) pe.debug_skip = (uint32_t)std::atoi(env);   // profiling aid: 16 no costs, 32 no stamps
    }
    pe.inline_big = finish ? 0u : 1u;   // (a rasterised frame does not launch k_view_nodes)
    const int grid = (int)(pe.block_hi - pe.block_lo);
    k_view_bounds<<<grid < 1 ? 1 : grid, kBlock, 0, stream>>>(sc, vp, view, part, pe);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (after_reset)  // the per-slab counters of the previous frame are gone from here on
    {
        e = cudaEventRecord(after_reset, stream);
        if (e != cudaSuccess) return e;
    }
    if (after_bounds && (e = cudaEventRecord(after_bounds, stream)) != cudaSuccess) return e;
    if (!finish) return cudaSuccess;  // rasterised frame: no tile lists, no node ranges
    ViewWait vw{};
    if (wait) vw = *wait;
    if (pe.ranges_only)
    {
        // every rank stamps its own tiles from the complete range tables (waits for the peers' flags first)
        const uint32_t all_blocks = (sc.n_slots + kBlock - 1) / kBlock;
        k_view_stamp<<<all_blocks < 1u ? 1u : all_blocks, kBlock, 0, stream>>>(sc, vp, view, part, vw, pe.debug_skip);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        vw = ViewWait{};
    }

    // one warp per node + 296 compaction blocks of 32 registers: all resident at once
    uint32_t node_blocks = (sc.n_nodes + kBlock / 32 - 1) / (kBlock / 32);
    if (node_blocks > 296u) node_blocks = 296u;
    k_view_nodes<<<node_blocks + 296u, kBlock, 0, stream>>>(sc, view, cam, part, slabs, node_blocks, vw);
    return cudaGetLastError();
}

// The culled kernel is persistent: its grid is exactly what is resident at once (a slab completes only when every warp
// has done its static sky share, so no CTA may wait for an SM).
template <bool STRICT, bool COUNT, int MINB>
static cudaError_t launch_culled_m(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                                   const DevScratch& scratch, const DevView& view, const DevSlabs& slabs, DevCounters* counters,
                                   unsigned long long jitter_seed, LaunchCfg cfg, cudaStream_t stream)
{
    static int per_sm = 0;
    if (per_sm == 0)
    {
        int n = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render_culled<STRICT, COUNT, MINB>, kBlock, 0);
        if (e != cudaSuccess) return e;
        per_sm = n < 1 ? 1 : n;
    }
    const int sms = cfg.grid / 4;  // render_launch_cfg: 4 CTAs per SM is the upper bound the scratch areas are sized for
    int grid = sms * per_sm;
    if (grid > cfg.grid) grid = cfg.grid;
    k_render_culled<STRICT, COUNT, MINB><<<grid, cfg.block, 0, stream>>>(sc, cam, part, out, scratch, view, slabs, counters, jitter_seed);
    return cudaGetLastError();
}
template <bool STRICT, bool COUNT>
static cudaError_t launch_culled_t(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                                   const DevScratch& scratch, const DevView& view, const DevSlabs& slabs, DevCounters* counters,
                                   unsigned long long jitter_seed, LaunchCfg cfg, cudaStream_t stream)
{
    // tiles of this rank per resident warp at 3 CTAs per SM: above ~20 (a whole 4K frame: 73, half of one: 36) take the 4-CTA build
    static int forced = -1;
    if (forced < 0)
    {
        const char* env = std::getenv("ATRAY_RENDER_MIN_BLOCKS");
        forced = env ? std::atoi(env) : 0;
    }
    const long long tiles = (long long)((cam.W + 7) >> 3) * ((cam.H + 3) >> 2) / (part.world > 0 ? part.world : 1);
    const long long warps3 = (long long)(cfg.grid / 4) * 3 * (kBlock / 32);
    // (not for jittered frames: their sample loop pays for the spills — Deer 1080p 16 spp: 452 us at 3 CTAs, 541 us at 4)
    const bool four = forced ? forced >= 4 : (!cam.aa && tiles > 20 * warps3);
    if (four) return launch_culled_m<STRICT, COUNT, 4>(sc, cam, part, out, scratch, view, slabs, counters, jitter_seed, cfg, stream);
    return launch_culled_m<STRICT, COUNT, ATRAY_RENDER_MIN_BLOCKS>(sc, cam, part, out, scratch, view, slabs, counters, jitter_seed, cfg, stream);
}

cudaError_t launch_render_culled(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                                 const DevScratch& scratch, const DevView& view, const DevSlabs& slabs, DevCounters* counters,
                                 int fp_fast, unsigned long long jitter_seed, LaunchCfg cfg, cudaStream_t stream)
{
    if (counters)
    {
        if (fp_fast) return launch_culled_t<false, true>(sc, cam, part, out, scratch, view, slabs, counters, jitter_seed, cfg, stream);
        return launch_culled_t<true, true>(sc, cam, part, out, scratch, view, slabs, counters, jitter_seed, cfg, stream);
    }
    if (fp_fast) return launch_culled_t<false, false>(sc, cam, part, out, scratch, view, slabs, counters, jitter_seed, cfg, stream);
    return launch_culled_t<true, false>(sc, cam, part, out, scratch, view, slabs, counters, jitter_seed, cfg, stream);
}

// ---- rasterised frame (raster.cuh) ---------------------------------------------------------------------------------------
cudaError_t launch_film_tables(const DevCamera& cam, const DevRaster& rs, cudaStream_t stream)
{
    const int n = cam.W + cam.H;
    k_film_tables<<<(n + kBlock - 1) / kBlock, kBlock, 0, stream>>>(cam, rs);
    return cudaGetLastError();
}

template <bool STRICT, bool COUNT>
static cudaError_t launch_raster_t(const DevScene& sc, const ViewParams& vp, const DevCamera& cam, const DevPartition& part,
                                   const DevOut& out, const DevScratch& scratch, const DevView& view, const DevRaster& rs,
                                   const DevSlabs& slabs, DevCounters* counters, LaunchCfg cfg, cudaStream_t stream)
{
    static int per_sm = 0;
    if (per_sm == 0)
    {
        int n = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_resolve<STRICT, COUNT>, kBlock, 0);
        if (e != cudaSuccess) return e;
        per_sm = n < 1 ? 1 : n;
    }
    const int sms = cfg.grid / 4;
    const uint32_t groups = (sc.n_slots + 31u) / 32u;
    k_raster<STRICT, COUNT><<<(groups + 3u) / 4u, 128, 0, stream>>>(sc, vp, cam, part, view, rs, counters);
    k_raster_big<STRICT, COUNT><<<sms * 4, kBlock, 0, stream>>>(sc, vp, cam, part, view, rs, counters);
    k_hit_filter<<<sms * 4, kBlock, 0, stream>>>(rs);
    int grid = sms * per_sm;
    if (grid > cfg.grid) grid = cfg.grid;  // the spill areas are sized for cfg.grid CTAs
    k_resolve<STRICT, COUNT><<<grid, kBlock, 0, stream>>>(sc, cam, part, out, scratch, view, rs, slabs, counters);
    return cudaGetLastError();
}

cudaError_t launch_raster(const DevScene& sc, const ViewParams& vp, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                          const DevScratch& scratch, const DevView& view, const DevRaster& rs, const DevSlabs& slabs,
                          DevCounters* counters, int fp_fast, LaunchCfg cfg, cudaStream_t stream)
{
    if (counters)
    {
        if (fp_fast) return launch_raster_t<false, true>(sc, vp, cam, part, out, scratch, view, rs, slabs, counters, cfg, stream);
        return launch_raster_t<true, true>(sc, vp, cam, part, out, scratch, view, rs, slabs, counters, cfg, stream);
    }
    if (fp_fast) return launch_raster_t<false, false>(sc, vp, cam, part, out, scratch, view, rs, slabs, counters, cfg, stream);
    return launch_raster_t<true, false>(sc, vp, cam, part, out, scratch, view, rs, slabs, counters, cfg, stream);
}

cudaError_t launch_path(const DevScene& sc, const DevCamera& cam, const DevPartition& part, const DevOut& out,
                        const DevScratch& scratch, unsigned int* tile_counter, unsigned long long* casts, int fp_fast,
                        unsigned long long seed, LaunchCfg cfg, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(tile_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if (casts)
    {
        e = cudaMemsetAsync(casts, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
    }
    if (fp_fast) k_path<false><<<cfg.grid, cfg.block, 0, stream>>>(sc, cam, part, out, scratch, tile_counter, casts, seed);
    else k_path<true><<<cfg.grid, cfg.block, 0, stream>>>(sc, cam, part, out, scratch, tile_counter, casts, seed);
    return cudaGetLastError();
}

cudaError_t launch_trace(const DevScene& sc, const float* rays, unsigned long long n, uint32_t* prim, float* t,
                         const DevScratch& scratch, unsigned int* tile_counter, DevCounters* counters, int fp_fast,
                         LaunchCfg cfg, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(tile_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if (counters)
    {
        if (fp_fast) k_trace<false, true><<<cfg.grid, cfg.block, 0, stream>>>(sc, rays, n, prim, t, scratch, tile_counter, counters);
        else k_trace<true, true><<<cfg.grid, cfg.block, 0, stream>>>(sc, rays, n, prim, t, scratch, tile_counter, counters);
    }
    else
    {
        if (fp_fast) k_trace<false, false><<<cfg.grid, cfg.block, 0, stream>>>(sc, rays, n, prim, t, scratch, tile_counter, counters);
        else k_trace<true, false><<<cfg.grid, cfg.block, 0, stream>>>(sc, rays, n, prim, t, scratch, tile_counter, counters);
    }
    return cudaGetLastError();
}

// ray list with the hit's barycentrics (the single-ray query get_ray_kd_tree_intersection of the host layer): strict arithmetic only
cudaError_t launch_trace_uv(const DevScene& sc, const float* rays, unsigned long long n, uint32_t* prim, float* t, float* uv,
                            const DevScratch& scratch, unsigned int* tile_counter, LaunchCfg cfg, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(tile_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const unsigned long long groups = (n + 31ULL) >> 5, per_block = kBlock / 32;
    const unsigned long long want = (groups + per_block - 1) / per_block;
    const int grid = int(want < (unsigned long long)cfg.grid ? (want ? want : 1ULL) : (unsigned long long)cfg.grid);
    k_trace<true, false, true><<<grid, cfg.block, 0, stream>>>(sc, rays, n, prim, t, scratch, tile_counter, nullptr, uv);
    return cudaGetLastError();
}

// ---- ray-list trace without any allocation (include/atray_b200.h: atray_trace_rays_static) ---------------------------------------
// k_trace needs scratch memory: the work cursor its warps draw ray groups from (must start at zero) and, per resident warp, the
// overflow areas of the leaf list and the inner-node stack (used when a ray outgrows the per-lane local lists). The ordinary entry
// takes them from buffers cudaMalloc'ed at scene upload; this variant takes them from ONE file-scope device array of
// kStaticScratchBytes that every GPU loading the library holds until exit. Layout: [0, 256) cursor | leaves (warps x leaf_cap x
// uint2) | stack (warps x stack_cap x u32), both 256-byte aligned. In host code `g_static_scratch` is only a stand-in for the
// symbol: the device address comes from cudaGetSymbolAddress, once per device with that device current (static_scratch_address).
// Every call on a device shares the array (it keeps the last call's contents; only the cursor has to be cleared, which the
// launcher does on the stream): calls that overlap on one GPU race.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
static constexpr int kStaticMaxDevices = 64;
static void* g_static_scratch_addr[kStaticMaxDevices];
static std::mutex g_static_scratch_mutex;

size_t static_scratch_capacity() { return kStaticScratchBytes; }

static size_t static_round256(size_t v) { return (v + 255u) & ~size_t(255); }

// bytes of scratch one CTA of k_trace needs for a scene with these list capacities
size_t static_trace_bytes_per_block(uint32_t leaf_cap, uint32_t stack_cap)
{
    return size_t(kBlock / 32) * (size_t(leaf_cap) * sizeof(uint2) + size_t(stack_cap) * sizeof(uint32_t));
}

// CTAs that fit the static scratch (0: not even one: the scene is beyond the maximum), at most cfg.grid
int static_trace_blocks(uint32_t leaf_cap, uint32_t stack_cap, LaunchCfg cfg)
{
    const size_t per_block = static_trace_bytes_per_block(leaf_cap, stack_cap);
    const size_t room = kStaticScratchBytes - 768;   // cursor line + the two alignment gaps
    const size_t fit = per_block ? room / per_block : size_t(cfg.grid);
    return int(fit < size_t(cfg.grid) ? fit : size_t(cfg.grid));
}

// device address of the static scratch on `device`, which must be the current device; resolved once per device
cudaError_t static_scratch_address(int device, void** out)
{
    if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (!g_static_scratch_addr[device])
    {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_static_scratch_addr[device] = p;
    }
    *out = g_static_scratch_addr[device];
    return cudaSuccess;
}

cudaError_t launch_trace_static(const DevScene& sc, const float* rays, unsigned long long n, uint32_t* prim, float* t, uint32_t leaf_cap,
                                uint32_t stack_cap, int fp_fast, LaunchCfg cfg, int device, cudaStream_t stream)
{
    const int blocks = static_trace_blocks(leaf_cap, stack_cap, cfg);
    if (blocks < 1) return cudaErrorInvalidValue;   // (the C ABI checks this first and reports the limit)
    void* base = nullptr;
    cudaError_t e = static_scratch_address(device, &base);
    if (e != cudaSuccess) return e;
    unsigned char* p = static_cast<unsigned char*>(base);
    unsigned int* cursor = reinterpret_cast<unsigned int*>(p);
    const size_t warps = size_t(blocks) * (kBlock / 32);
    DevScratch scratch;
    scratch.leaf_cap = leaf_cap;
    scratch.stack_cap = stack_cap;
    scratch.leaves = reinterpret_cast<uint2*>(p + 256);
    scratch.stack = reinterpret_cast<uint32_t*>(p + 256 + static_round256(warps * leaf_cap * sizeof(uint2)));
    e = cudaMemsetAsync(cursor, 0, 256, stream);
    if (e != cudaSuccess) return e;
    if (fp_fast) k_trace<false, false><<<blocks, cfg.block, 0, stream>>>(sc, rays, n, prim, t, scratch, cursor, nullptr);
    else k_trace<true, false><<<blocks, cfg.block, 0, stream>>>(sc, rays, n, prim, t, scratch, cursor, nullptr);
    return cudaGetLastError();
}

cudaError_t launch_c5(const DevScene& sc, unsigned long long p0, unsigned long long p1, uint32_t spp, uint32_t chunk_px,
                      unsigned long long seed, const DevPartition& part, uint32_t* bgrx, uint32_t* prim, float* t, const DevScratch& scratch,
                      unsigned int* counter, DevCounters* counters, int fp_fast, LaunchCfg cfg, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    DevC5 job;
    job.p0 = p0; job.p1 = p1; job.spp = spp; job.chunk_px = chunk_px; job.seed = seed; job.bgrx = bgrx; job.prim = prim; job.t = t;
    if (counters)
    {
        if (fp_fast) k_c5<false, true><<<cfg.grid, cfg.block, 0, stream>>>(sc, job, part, scratch, counter, counters);
        else k_c5<true, true><<<cfg.grid, cfg.block, 0, stream>>>(sc, job, part, scratch, counter, counters);
    }
    else
    {
        if (fp_fast) k_c5<false, false><<<cfg.grid, cfg.block, 0, stream>>>(sc, job, part, scratch, counter, counters);
        else k_c5<true, false><<<cfg.grid, cfg.block, 0, stream>>>(sc, job, part, scratch, counter, counters);
    }
    return cudaGetLastError();
}

cudaError_t launch_microbench_fp32(float* sink, int iters, LaunchCfg cfg, cudaStream_t stream)
{
    k_fp32_issue<<<cfg.grid, cfg.block, 0, stream>>>(sink, iters);
    return cudaGetLastError();
}
cudaError_t launch_microbench_l2(const float4* buf, size_t n_vec, int passes, float* sink, LaunchCfg cfg, cudaStream_t stream)
{
    k_l2_read<<<cfg.grid, cfg.block, 0, stream>>>(buf, n_vec, passes, sink);
    return cudaGetLastError();
}

}  // namespace atray
