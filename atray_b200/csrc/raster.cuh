// Film-space rasteriser of the primary-ray frame (AA off, root of the tree is an inner node) — included by render.cu.
//
// Same result bits as the reference's per-ray walk (kd_tree.cpp:337-465 under renderer.cpp:314-365), computed triangle
// by triangle instead of ray by ray. For one camera ray the reference returns: among the leaves the ray lists (entry
// distance > 0, aabb.h:29-63), sorted by (entry distance, insertion order), the first one in which some triangle passes
// the float Möller–Trumbore test with t in (1e-4, MAX_FLOAT) (model.h:75-103, kd_tree.cpp:437-462), and inside that leaf
// the minimum of (t, position in the leaf). Both minima are order-independent, so they can be taken over (triangle slot,
// pixel) pairs in any order:
//   k_raster / k_raster_big  every slot visits the pixels of its film rectangle (view_bounds.h: a ray outside it cannot be
//                            accepted), regenerates the pixel's ray with the reference arithmetic, runs the reference
//                            triangle test, and for an accepted hit lowers key1[pixel] = (entry distance of the slot's
//                            leaf, static insertion rank of the leaf) and appends the hit to a list;
//   k_hit_filter             hits whose leaf won their pixel lower key2[pixel] = (t, position);
//   k_resolve                sky fill, decode of (key1, key2) into {t, face, BGRX}, keys restored to all-ones.
// What the per-ray walk does beyond "entry distance > 0" is covered as follows: a leaf that passes its own slab test
// passes the test of every ancestor when the ray is regular (finite origin, direction and reciprocals);
// the "at most 5 accepted children" cut (kd_tree.cpp:374) can only trigger at a node where two split-plane parameters of
// the ray coincide. A pixel with an accepted hit whose ray is irregular, or has such a tie at any ancestor of the hit's
// leaf, is flagged (key1 = 0) and traced by the exact per-ray walker inside k_resolve; so are pixels whose hits did not
// fit the list. The insertion rank is static: the walk examines children in index order and pops the last pushed inner
// node first, so the order in which a ray meets its leaves is the full-tree order restricted to them
// (host_model.cpp: FlatScene::leaf_seq).

static constexpr uint32_t kWarpPairs = 256;    // slots with more pixel pairs go to the medium / giant lists
static constexpr uint32_t kChunkPairs = 2048;  // pairs per CTA pass of k_raster_big
static constexpr unsigned long long kKeyEmpty = ~0ull;

// rows of this rank: band b (band_rows scanlines) belongs to rank b % world
struct RankRows
{
    int B, P, off;  // band rows, period, first row of the rank inside a period
    __device__ __forceinline__ RankRows(const DevPartition& p) : B(p.band_rows), P(p.band_rows * p.world), off(p.rank * p.band_rows) {}
    __device__ __forceinline__ int before(int y) const  // rows of this rank in [0, y)
    {
        const int q = y / P, r = y - q * P - off;
        return q * B + (r < 0 ? 0 : (r > B ? B : r));
    }
    __device__ __forceinline__ int row(int m) const  // m-th row of this rank
    {
        const int q = m / B;
        return q * P + off + (m - q * B);
    }
};

// Per-x and per-y halves of the camera ray (renderer.cpp:317,329,350: pixel_pos = (frame_center + camera_x*film_x) +
// camera_y*film_y) — the same float operations camera_ray() performs, split where the expression splits. Also resets the
// rasteriser's list counters for the frame.
__global__ void __launch_bounds__(kBlock) k_film_tables(DevCamera cm, DevRaster rs)
{
    const int i = blockIdx.x * kBlock + threadIdx.x;
    if (i < 4) rs.ctl[i] = 0u;
    if (i < cm.W)
    {
        const float fx = mulr(mulr(film_coord(i, cm.W), cm.h_fov), cm.aspect);
        rs.tabx[i] = make_float4(addr(cm.fc[0], mulr(cm.cx[0], fx)), addr(cm.fc[1], mulr(cm.cx[1], fx)), addr(cm.fc[2], mulr(cm.cx[2], fx)), fx);
    }
    else if (i < cm.W + cm.H)
    {
        const int y = i - cm.W;
        const float fy = film_coord(y, cm.H);
        rs.taby[y] = make_float4(mulr(cm.cy[0], fy), mulr(cm.cy[1], fy), mulr(cm.cy[2], fy), fy);
    }
}

// One (slot, pixel) pair. Called by whole warps (`valid` masks the lanes without a pair): the hit append is
// warp-aggregated.
template <bool STRICT, bool COUNT>
__device__ __forceinline__ void raster_pair(const DevScene& sc, const DevCamera& cm, const DevRaster& rs, int lane, bool valid,
                                            uint32_t slot, int x, int y, const float4 R, WorkCount& wc)
{
    bool emit = false, flag = false;
    uint32_t pix = 0, disb = 0, seq = 0, tb = 0, pos = 0;
    if (valid)
    {
        const float4 X = __ldg(&rs.tabx[x]), Y = __ldg(&rs.taby[y]);
        if (COUNT) wc.rect++;
        if (!(X.w < R.x || X.w > R.z || Y.w < R.y || Y.w > R.w))
        {
            // camera_ray(): pixel_pos − eye, normalize (ray.h:24-29, PL_math.h:387-392)
            RayL r;
            r.ox = cm.eye[0]; r.oy = cm.eye[1]; r.oz = cm.eye[2];
            const float dx = subr(addr(X.x, Y.x), r.ox), dy = subr(addr(X.y, Y.y), r.oy), dz = subr(addr(X.z, Y.z), r.oz);
            const float len2 = addr(addr(mulr(dx, dx), mulr(dy, dy)), mulr(dz, dz));
            const float inv = __fdiv_rn(1.0f, __fsqrt_rn(len2));
            r.dx = mulr(dx, inv); r.dy = mulr(dy, inv); r.dz = mulr(dz, inv);
            float t;
            uint32_t face;
            if (tri_accept<STRICT, COUNT>(sc.tris + (size_t)(slot >> 3) * 24 + (slot & 7), r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, t, pos, face, wc))
            {
                pix = (uint32_t)y * (uint32_t)cm.W + (uint32_t)x;
                const uint32_t leaf = __ldg(&sc.packet_leaf[slot >> 3]);
                const float ix = __frcp_rn(r.dx), iy = __frcp_rn(r.dy), iz = __frcp_rn(r.dz);  // Optimized_Ray, renderer.cpp:41-44
                const float chk = (r.ox - r.ox) + (r.oy - r.oy) + (r.oz - r.oz) + (ix - ix) + (iy - iy) + (iz - iz) +
                                  (r.dx - r.dx) + (r.dy - r.dy) + (r.dz - r.dz);
                if (COUNT) wc.box++;
                const float dis = box_entry(r, ix, iy, iz, __ldg(&sc.nodes[2 * leaf]), __ldg(&sc.nodes[2 * leaf + 1]));
                if (!(chk == 0.0f)) flag = true;  // irregular ray: the per-ray walker decides
                else if (dis > 0.0f)              // the ray lists this leaf (kd_tree.cpp:381-409)
                {
                    if (!box_check(r, ix, iy, iz, __ldg(&sc.nodes[0]), __ldg(&sc.nodes[1]))) flag = true;  // kd_tree.cpp:339 (implied; kept as a guard)
                    for (uint32_t a = __ldg(&sc.parent[leaf]); a != 0xffffffffu; a = __ldg(&sc.parent[a]))
                    {
                        // split point of node a = hi corner of its child 0 (kd_tree.cpp:120-148)
                        const uint32_t c0 = (uint32_t)__float_as_int(__ldg(&sc.nodes[2 * a]).w);
                        const float4 s = __ldg(&sc.nodes[2 * c0 + 1]);
                        const float tsx = mulr(subr(s.x, r.ox), ix), tsy = mulr(subr(s.y, r.oy), iy), tsz = mulr(subr(s.z, r.oz), iz);
                        if (COUNT) wc.box++;
                        if (tsx == tsy || tsy == tsz || tsx == tsz) flag = true;  // the ≤5 cut may trigger here
                    }
                    if (!flag)
                    {
                        disb = __float_as_uint(dis);
                        seq = __ldg(&sc.leaf_seq[leaf]);
                        tb = __float_as_uint(t);
                        const unsigned long long k1 = ((unsigned long long)disb << 32) | (unsigned long long)seq;
                        const unsigned long long old = atomicMin(&rs.key1[pix], k1);
                        emit = k1 <= old;  // a leaf that already lost this pixel can never win it back
                    }
                }
            }
        }
    }
    if (flag) atomicMin(&rs.key1[pix], 0ull);
    const unsigned m = __ballot_sync(FULL_MASK, emit);
    if (m)
    {
        const int leader = __ffs(m) - 1;
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&rs.ctl[0], (uint32_t)__popc(m));
        base = __shfl_sync(FULL_MASK, base, leader);
        if (emit)
        {
            const uint32_t at = base + (uint32_t)__popc(m & ((1u << lane) - 1u));
            if (at < rs.hit_cap)
            {
                rs.hit_a[at] = make_uint4(pix, disb, seq, tb);
                rs.hit_pos[at] = pos;
            }
            else
                atomicMin(&rs.key1[pix], 0ull);  // list full: exact walker
        }
    }
}

// pixel range of a slot's rectangle restricted to this rank's rows: x0, width, first own row (rank numbering), pairs
struct SlotSpan { int x0, w, m0; uint32_t cnt; };
__device__ __forceinline__ SlotSpan slot_span(const ViewParams& vp, const RankRows& rr, const float4 R)
{
    SlotSpan s;
    s.x0 = 0; s.w = 0; s.m0 = 0; s.cnt = 0;
    FilmRect fr;
    fr.xlo = R.x; fr.ylo = R.y; fr.xhi = R.z; fr.yhi = R.w;
    PxRect pr;
    if (!film_rect_px(vp, fr, pr)) return s;
    const int x0 = pr.x0, x1 = pr.x1, y0 = pr.y0, y1 = pr.y1;
    s.x0 = x0;
    s.w = x1 - x0 + 1;
    s.m0 = rr.before(y0);
    const int rows = rr.before(y1 + 1) - s.m0;
    s.cnt = (uint32_t)s.w * (uint32_t)(rows > 0 ? rows : 0);
    return s;
}

// One warp per 32 consecutive slots. Small rectangles are expanded into (slot, pixel) pairs across the warp (prefix sum
// of the pair counts, 32 pairs per step, owner found by binary search over the prefix); larger ones are queued.
template <bool STRICT, bool COUNT>
__global__ void __launch_bounds__(128) k_raster(DevScene sc, ViewParams vp, DevCamera cm, DevPartition part, DevView vw, DevRaster rs,
                                                DevCounters* counters)
{
    const int lane = threadIdx.x & 31;
    const uint32_t group = (blockIdx.x * 128u + threadIdx.x) >> 5;
    const uint32_t slot = group * 32u + (uint32_t)lane;
    if (group * 32u >= sc.n_slots) return;
    const RankRows rr(part);
    WorkCount wc;
    wc.clear();
    float4 R = make_float4(INFINITY, INFINITY, -INFINITY, -INFINITY);
    if (slot < sc.n_slots) R = vw.tri_rect[slot];
    SlotSpan sp = slot_span(vp, rr, R);
    if (sp.cnt > kWarpPairs)
    {
        const bool giant = sp.cnt > kChunkPairs;
        const uint32_t at = atomicAdd(&rs.ctl[giant ? 2 : 1], 1u);
        if (giant) rs.big[rs.big_cap - 1u - at] = make_uint2(slot, sp.cnt);  // giants grow down from the end
        else rs.big[at] = make_uint2(slot, sp.cnt);                           // medium entries grow up from 0
        sp.cnt = 0;
    }
    uint32_t incl = sp.cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1)
    {
        const uint32_t v = __shfl_up_sync(FULL_MASK, incl, off);
        if (lane >= off) incl += v;
    }
    const uint32_t excl = incl - sp.cnt;
    const uint32_t total = __shfl_sync(FULL_MASK, incl, 31);
    for (uint32_t base = 0; base < total; base += 32u)
    {
        const uint32_t j = base + (uint32_t)lane;
        const bool valid = j < total;
        // owner = last lane whose exclusive prefix is <= j (lanes without pairs share their successor's prefix and lose)
        int o = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1)
        {
            const uint32_t e = __shfl_sync(FULL_MASK, excl, o + step);  // o + step <= 31
            if (e <= j) o += step;
        }
        const int ow = __shfl_sync(FULL_MASK, sp.w, o), ox0 = __shfl_sync(FULL_MASK, sp.x0, o), om0 = __shfl_sync(FULL_MASK, sp.m0, o);
        const uint32_t oe = __shfl_sync(FULL_MASK, excl, o);
        const float4 oR = make_float4(__shfl_sync(FULL_MASK, R.x, o), __shfl_sync(FULL_MASK, R.y, o), __shfl_sync(FULL_MASK, R.z, o),
                                      __shfl_sync(FULL_MASK, R.w, o));
        int x = 0, y = 0;
        if (valid)
        {
            const uint32_t q = j - oe;
            const uint32_t k = q / (uint32_t)ow;
            x = ox0 + (int)(q - k * (uint32_t)ow);
            y = rr.row(om0 + (int)k);
        }
        raster_pair<STRICT, COUNT>(sc, cm, rs, lane, valid, group * 32u + (uint32_t)o, x, y, oR, wc);
    }
    if (COUNT) flush_counters(counters, wc, 0u);
}

// Queued slots: a medium entry (<= kChunkPairs pairs) is one CTA pass, entry e belongs to CTA e % gridDim; a giant entry
// is cut into chunks of kChunkPairs pairs that are dealt round-robin to the whole grid.
template <bool STRICT, bool COUNT>
__global__ void __launch_bounds__(kBlock) k_raster_big(DevScene sc, ViewParams vp, DevCamera cm, DevPartition part, DevView vw, DevRaster rs,
                                                       DevCounters* counters)
{
    const int lane = threadIdx.x & 31;
    const RankRows rr(part);
    WorkCount wc;
    wc.clear();
    const uint32_t n_med = min(__ldcg(&rs.ctl[1]), rs.big_cap), n_giant = min(__ldcg(&rs.ctl[2]), rs.big_cap);
    for (uint32_t e = blockIdx.x; e < n_med; e += gridDim.x)
    {
        const uint2 ent = __ldcg(&rs.big[e]);
        const float4 R = vw.tri_rect[ent.x];
        const SlotSpan sp = slot_span(vp, rr, R);
        for (uint32_t base = 0; base < sp.cnt; base += kBlock)
        {
            const uint32_t j = base + threadIdx.x;
            const bool valid = j < sp.cnt;
            int x = 0, y = 0;
            if (valid)
            {
                const uint32_t k = j / (uint32_t)sp.w;
                x = sp.x0 + (int)(j - k * (uint32_t)sp.w);
                y = rr.row(sp.m0 + (int)k);
            }
            raster_pair<STRICT, COUNT>(sc, cm, rs, lane, valid, ent.x, x, y, R, wc);
        }
    }
    for (uint32_t e = 0; e < n_giant; ++e)
    {
        const uint2 ent = __ldcg(&rs.big[rs.big_cap - 1u - e]);
        const uint32_t nch = (ent.y + kChunkPairs - 1u) / kChunkPairs;
        uint32_t c = (blockIdx.x + gridDim.x - (e * 37u) % gridDim.x) % gridDim.x;  // rotate so that short entries spread out
        if (c >= nch) continue;
        const float4 R = vw.tri_rect[ent.x];
        const SlotSpan sp = slot_span(vp, rr, R);
        for (; c < nch; c += gridDim.x)
        {
            const uint32_t lo = c * kChunkPairs, hi = min(sp.cnt, lo + kChunkPairs);
            for (uint32_t base = lo; base < hi; base += kBlock)
            {
                const uint32_t j = base + threadIdx.x;
                const bool valid = j < hi;
                int x = 0, y = 0;
                if (valid)
                {
                    const uint32_t k = j / (uint32_t)sp.w;
                    x = sp.x0 + (int)(j - k * (uint32_t)sp.w);
                    y = rr.row(sp.m0 + (int)k);
                }
                raster_pair<STRICT, COUNT>(sc, cm, rs, lane, valid, ent.x, x, y, R, wc);
            }
        }
    }
    if (COUNT) flush_counters(counters, wc, 0u);
}

// Hits whose leaf won their pixel compete on (t, position in the reference's leaf order).
__global__ void __launch_bounds__(kBlock) k_hit_filter(DevRaster rs)
{
    const uint32_t n = min(__ldcg(&rs.ctl[0]), rs.hit_cap);
    for (uint32_t i = blockIdx.x * kBlock + threadIdx.x; i < n; i += gridDim.x * kBlock)
    {
        const uint4 a = __ldcg(&rs.hit_a[i]);
        const unsigned long long k1 = ((unsigned long long)a.y << 32) | (unsigned long long)a.z;
        if (__ldcg(&rs.key1[a.x]) == k1) atomicMin(&rs.key2[a.x], ((unsigned long long)a.w << 32) | (unsigned long long)__ldcg(&rs.hit_pos[i]));
    }
}

// One busy 8x4 tile: decode the pixel keys, trace flagged pixels with the exact walker, shade, store, restore the keys.
template <bool STRICT, bool COUNT>
__device__ __forceinline__ void resolve_tile(const DevScene& sc, const DevCamera& cm, const DevRaster& rs, const DevScratch& scratch,
                                             const DevOut& out, int warp_slot, int lane, int tcol, int y0, uint32_t sky_px, uint32_t hit_px,
                                             uint2* leaves, uint32_t* stack, WorkCount& wc, unsigned int& rays_done)
{
    const int x = tcol * 8 + (lane & 7), y = y0 + (lane >> 3);
    const bool active = x < cm.W && y < cm.H;
    const size_t idx = (size_t)y * (size_t)cm.W + (size_t)x;
    unsigned long long k1 = kKeyEmpty;
    if (active) k1 = __ldcg(&rs.key1[idx]);
    float t = kMaxFloat;
    uint32_t face = 0xffffffffu;
    const bool fb = k1 == 0ull;
    if (k1 != kKeyEmpty)
    {
        const unsigned long long k2 = __ldcg(&rs.key2[idx]);
        rs.key1[idx] = kKeyEmpty;
        if (k2 != kKeyEmpty) rs.key2[idx] = kKeyEmpty;
        if (!fb)
        {
            const uint32_t leaf = __ldg(&sc.seq_leaf[(uint32_t)k1]);
            const uint32_t first = ~(uint32_t)__float_as_int(__ldg(&sc.nodes[2 * leaf]).w);
            t = __uint_as_float((uint32_t)(k2 >> 32));
            face = __ldg(&sc.pos_face[first + (uint32_t)k2]);
        }
    }
    if (__any_sync(FULL_MASK, fb))
    {
        const float film_y = film_coord(y, cm.H);
        const float film_x = mulr(mulr(film_coord(x, cm.W), cm.h_fov), cm.aspect);
        const RayL ray = camera_ray(cm, film_x, film_y);
        float t2;
        uint32_t f2;
        trace_warp<STRICT, COUNT>(sc, scratch, warp_slot, lane, fb, ray, leaves, stack, t2, f2, wc);
        if (fb) { t = t2; face = f2; if (COUNT) wc.ovf++; }
    }
    if (active)
    {
        if (out.bgrx) out.bgrx[idx] = face != 0xffffffffu ? hit_px : sky_px;
        if (out.t) out.t[idx] = t;
        if (out.prim) out.prim[idx] = face;
        rays_done += cm.spp;
    }
}

// Frame assembly, slab by slab (same per-slab counters as k_render_culled, so the host side can copy a slab out while
// later ones are still being assembled): sky tiles are written as 512-byte row segments, busy tiles are decoded.
template <bool STRICT, bool COUNT>
__global__ void __launch_bounds__(kBlock) k_resolve(DevScene sc, DevCamera cm, DevPartition part, DevOut out, DevScratch scratch, DevView vw,
                                                    DevRaster rs, DevSlabs slabs, DevCounters* counters)
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
    WorkCount wc;
    wc.clear();
    uint2 leaves[kLeafCap];
    uint32_t stack[kStackCap];
    unsigned int rays_done = 0;
    // pixel values: cast_ray at bounce_limit 1 adds the emission of the sky or of the model per sample (renderer.cpp:226,
    // 258, 353-358), then divide, clamp, truncate, pack
    float sr = 0.0f, sg = 0.0f, sb = 0.0f, hr = 0.0f, hg = 0.0f, hb = 0.0f;
    for (uint32_t s = 0; s < cm.spp; ++s)
    {
        sr = addr(sr, sc.sky[0]); sg = addr(sg, sc.sky[1]); sb = addr(sb, sc.sky[2]);
        hr = addr(hr, sc.mdl[0]); hg = addr(hg, sc.mdl[1]); hb = addr(hb, sc.mdl[2]);
    }
    const float nspp = (float)cm.spp;
    const uint32_t sky_px = pack_bgrx(__fdiv_rn(sr, nspp), __fdiv_rn(sg, nspp), __fdiv_rn(sb, nspp));
    const uint32_t hit_px = pack_bgrx(__fdiv_rn(hr, nspp), __fdiv_rn(hg, nspp), __fdiv_rn(hb, nspp));
    const bool vec_ok = (cm.W & 3) == 0 && (((size_t)out.bgrx | (size_t)out.t | (size_t)out.prim) & 15) == 0;
    for (int slab = 0; slab < slabs.n; ++slab)
    {
        const unsigned int tile_lo = slabs.tile_begin[slab], tile_hi = min(slabs.tile_begin[slab + 1], n_tiles);
        unsigned int finished = 0;
        const unsigned int strips_x = (unsigned int)(tiles_x + 31) >> 5;
        const unsigned int trow_lo = tile_lo / (unsigned int)tiles_x, trow_hi = tile_hi / (unsigned int)tiles_x;
        const unsigned int n_strips = (trow_hi - trow_lo) * strips_x;
        const uint4 px4 = make_uint4(sky_px, sky_px, sky_px, sky_px);
        const uint4 t4 = make_uint4(__float_as_uint(kMaxFloat), __float_as_uint(kMaxFloat), __float_as_uint(kMaxFloat), __float_as_uint(kMaxFloat));
        const uint4 m4 = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
        for (unsigned int st = (unsigned int)warp_slot; st < n_strips; st += n_warps)
        {
            const unsigned int srow = st / strips_x, scol = st - srow * strips_x;
            const unsigned int ltr = trow_lo + srow;  // local tile row
            const unsigned int lb = ltr / (unsigned int)rows_per_band, trow = ltr - lb * (unsigned int)rows_per_band;
            const int y0 = ((int)lb * part.world + part.rank) * part.band_rows + (int)trow * 4;
            if (y0 >= cm.H) continue;  // tile rows past the image in the last band (not counted in live_tiles either)
            const int tcol = (int)scol * 32 + lane;
            const bool in_img = tcol < tiles_x;
            const bool sky = in_img && !all_busy && __ldg(&vw.stamps[(y0 >> 2) * vw.tiles_x + tcol]) != vw.frame_id;
            const unsigned skym = __ballot_sync(FULL_MASK, sky);
            unsigned busym = __ballot_sync(FULL_MASK, in_img && !sky);
            finished += (unsigned int)(__popc(skym) + __popc(busym));
            if (COUNT && lane == 0) { wc.tsky += (unsigned int)__popc(skym); wc.tbusy += (unsigned int)__popc(busym); }
            if (skym && vec_ok)
            {
#pragma unroll
                for (int h = 0; h < 2; ++h)
                {
                    const int x = (int)scol * 256 + h * 128 + lane * 4;  // 4 pixels = half a tile row
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
            else if (skym)
            {
                unsigned m = skym;
                while (m)
                {
                    const int b = __ffs(m) - 1;
                    m &= m - 1;
                    const int x = ((int)scol * 32 + b) * 8 + (lane & 7), y = y0 + (lane >> 3);
                    if (x < cm.W && y < cm.H)
                    {
                        const size_t idx = (size_t)y * (size_t)cm.W + (size_t)x;
                        if (out.bgrx) out.bgrx[idx] = sky_px;
                        if (out.t) out.t[idx] = kMaxFloat;
                        if (out.prim) out.prim[idx] = 0xffffffffu;
                        rays_done += cm.spp;
                    }
                }
            }
            while (busym)
            {
                const int b = __ffs(busym) - 1;
                busym &= busym - 1;
                resolve_tile<STRICT, COUNT>(sc, cm, rs, scratch, out, warp_slot, lane, (int)scol * 32 + b, y0, sky_px, hit_px, leaves, stack, wc,
                                            rays_done);
            }
        }
        if (__any_sync(FULL_MASK, finished != 0u))
        {
            __threadfence();
            __syncwarp();
            if (lane == 0) atomicAdd(&vw.busy_count[kCtlDone + slab], finished);
        }
    }
    if (COUNT) flush_counters(counters, wc, rays_done);
}
