/* TEST INFRASTRUCTURE ONLY — CPU restatement (plain C) of ATRay's per-pixel render path.
 *
 * This is the ORACLE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may
 * compile, load or call it. The product (atray_b200/, include/) never links, loads or falls back to it.
 *
 * Parity status: PINNED against the reference's own code executed in this container — oracle/_ref
 * (libatray_ref.so, built by oracle/ref/build_ref.sh from /root/reference/Source) — on every mesh the
 * reference ships plus the procedural stand-ins (tests/test_oracle_vs_ref.py, tests/golden/). The reference
 * has no tests or golden vectors of its own, and one arithmetic boundary is project-defined:
 * normalize() uses SVML _mm_invsqrt_ps upstream (PL_math.h:387-392), defined here as 1.0f/sqrtf(x).
 *
 * Build: gcc -O2 -ffp-contract=off (no -ffast-math, no FMA contraction) — see oracle/Makefile.
 * Every function cites the reference file:line (relative to /root/reference/Source) it restates.
 */
#include <errno.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define OR_MAX_FLOAT 3.402823466e+38F /* PL/PL_base_defs.h:70 */
#define OR_TOL 0.0001f                /* engine/renderer/ray.h:5 */
#define OR_MISS 0xffffffffu

typedef struct { float x, y, z; } v3;

/* ---- PL/PL_math.h vector helpers (83-129, 330-338, 377-392, 416-422) ------------------------------------- */
static inline v3 v3_add(v3 a, v3 b) { v3 r = { a.x + b.x, a.y + b.y, a.z + b.z }; return r; }
static inline v3 v3_sub(v3 a, v3 b) { v3 r = { a.x - b.x, a.y - b.y, a.z - b.z }; return r; }
static inline v3 v3_mul(v3 a, float n) { v3 r = { a.x * n, a.y * n, a.z * n }; return r; }
static inline v3 v3_div(v3 a, float n) { v3 r = { a.x / n, a.y / n, a.z / n }; return r; }
static inline v3 v3_neg(v3 a) { v3 r = { -a.x, -a.y, -a.z }; return r; }
static inline float f_min(float a, float b) { return a > b ? b : a; } /* PL_math.h:330-333 */
static inline float f_max(float a, float b) { return a > b ? a : b; } /* PL_math.h:335-338 */
static inline float v3_dot(v3 p, v3 n) { return (p.x * n.x) + (p.y * n.y) + (p.z * n.z); }
static inline v3 v3_cross(v3 a, v3 b)
{
    v3 r = { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x };
    return r;
}
static inline float v3_mag2(v3 v) { return (v.x * v.x) + (v.y * v.y) + (v.z * v.z); }
/* normalize, PL_math.h:387-392: v * invsqrt(mag2(v)); invsqrt := 1.0f/sqrtf (project definition) */
static inline v3 v3_normalize(v3 v)
{
    float inv = 1.0f / sqrtf(v3_mag2(v));
    return v3_mul(v, inv);
}

/* ---- PCG32, PL_math.h:492-541 ------------------------------------------------------------------------------ */
typedef struct { uint64_t state, stream; } or_rng;
static inline uint32_t or_random_u32(or_rng* s)
{
    uint64_t old = s->state;
    s->state = old * 6364136223846793005ULL + (s->stream | 1);
    uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((0u - rot) & 31));
}
static inline float or_rand_uni(or_rng* s) { return (float)or_random_u32(s) * 2.328306437e-10F; }
static inline float or_rand_bi(or_rng* s) { return -1.0f + 2.0f * or_rand_uni(s); }

/* ---- number parser, utilities/parser.h:4-205 and the tables PL_math.h:6-41 -------------------------------- */
static const uint64_t OR_INT_POW10[20] = { 1ULL, 10ULL, 100ULL, 1000ULL, 10000ULL, 100000ULL, 1000000ULL, 10000000ULL,
    100000000ULL, 1000000000ULL, 10000000000ULL, 100000000000ULL, 1000000000000ULL, 10000000000000ULL,
    100000000000000ULL, 1000000000000000ULL, 10000000000000000ULL, 100000000000000000ULL, 1000000000000000000ULL,
    10000000000000000000ULL };
static const double OR_F64_POW10[48] = { 1.0e-28, 1.0e-27, 1.0e-26, 1.0e-25, 1.0e-24, 1.0e-23, 1.0e-22, 1.0e-21, 1.0e-20,
    1.0e-19, 1.0e-18, 1.0e-17, 1.0e-16, 1.0e-15, 1.0e-14, 1.0e-13, 1.0e-12, 1.0e-11, 1.0e-10, 1.0e-9, 1.0e-8, 1.0e-7,
    1.0e-6, 1.0e-5, 1.0e-4, 1.0e-3, 1.0e-2, 1.0e-1, 1.0e0, 1.0e1, 1.0e2, 1.0e3, 1.0e4, 1.0e5, 1.0e6, 1.0e7, 1.0e8,
    1.0e9, 1.0e10, 1.0e11, 1.0e12, 1.0e13, 1.0e14, 1.0e15, 1.0e16, 1.0e17, 1.0e18, 1.0e19 };
static inline int or_is_ws(char c) { return c == ' ' || c == '\t' || c == '\r'; }
static inline int or_is_digit(char c) { return c >= '0' && c <= '9'; }
static char* or_skip_ws(char* p) { while (or_is_ws(*p)) p++; return p; }

static char* or_parse_int(char* from, int32_t* val) /* parser.h:38-65 */
{
    from = or_skip_ws(from);
    int32_t sign = 1;
    if (*from == '+') from++;
    else if (*from == '-') { sign = -1; from++; }
    uint32_t v = 0; /* unsigned wrap == the two's-complement wrap the reference gets */
    while (or_is_digit(*from)) { v = v * 10u + (uint32_t)(*from - '0'); from++; }
    *val = (int32_t)v * sign;
    return from;
}

static char* or_parse_f64(char* from, double* val) /* parser.h:113-191 */
{
    double sign = 1.0;
    int32_t frac_prec = 0, exponent = 0, exponent_sign = 1;
    uint64_t fraction = 0, front = 0;
    from = or_skip_ws(from);
    if (*from == '+') from++;
    else if (*from == '-') { sign = -1.0; from++; }
    while (or_is_digit(*from)) { front = front * 10 + (uint64_t)(*from - '0'); from++; }
    if (*from == '.') from++;
    while (or_is_digit(*from)) { frac_prec++; fraction = fraction * 10 + (uint64_t)(*from - '0'); from++; }
    front *= OR_INT_POW10[frac_prec < 20 ? frac_prec : 19]; /* reference indexes out of bounds past 19 digits */
    front += fraction;
    if (*from == 'e' || *from == 'E')
    {
        from++;
        if (*from == '+') from++;
        else if (*from == '-') { exponent_sign = -1; from++; }
        while (or_is_digit(*from)) { exponent = 10 * exponent + (*from - '0'); from++; }
        exponent *= exponent_sign;
    }
    exponent -= frac_prec;
    double v = (double)front;
    v = v * sign;
    exponent = (exponent >= -28 && exponent <= 19) ? exponent : 0;
    v *= OR_F64_POW10[exponent + 28];
    *val = v;
    return from;
}

static char* or_parse_vec3f(char* pos, v3* out) /* parser.h:194-205 */
{
    double d;
    char* p = or_parse_f64(pos, &d); out->x = (float)d;
    p = or_parse_f64(p, &d); out->y = (float)d;
    p = or_parse_f64(p, &d); out->z = (float)d;
    return p;
}

/* ---- model data, engine/renderer/model.h:6-61,108-152 ------------------------------------------------------- */
typedef struct { v3 min, max; } or_box; /* aabb.h:5-17: bounds[0] = min, bounds[1] = max */
typedef struct { v3 a, b, c; } or_tri;
typedef struct { or_tri tri; uint32_t face; } or_prim; /* kd_tree.h:18-24 */

#define OR_VEC(T) struct { T* p; size_t n, cap; }
#define OR_PUSH(vec, val) do { if ((vec).n == (vec).cap) { (vec).cap = (vec).cap ? (vec).cap * 2 : 64; \
    (vec).p = realloc((vec).p, (vec).cap * sizeof(*(vec).p)); } (vec).p[(vec).n++] = (val); } while (0)

typedef struct { int32_t v[3], vt[3], vn[3]; } or_face;

typedef struct
{
    or_box box;
    int32_t children_start; /* 0 = leaf (union with has_children, kd_tree.h:30-34) */
    uint32_t count;
    or_prim* prims;
} or_node;

typedef struct
{
    OR_VEC(v3) vertices, normals, tex;
    OR_VEC(or_face) faces;
    or_box aabb;
    OR_VEC(or_node) nodes;
    v3 sky_emission, model_emission; /* app.cpp:91,96 */
    /* full materials + analytic primitives for the multi-bounce path (material.h:4-9, sphere.h:5-10, plane.h:5-10) */
    struct or_material { v3 emission, reflection; float scatter; } sky_mat, model_mat;
    OR_VEC(struct or_sphere { v3 center; float radius; struct or_material m; }) spheres;
    OR_VEC(struct or_plane { v3 normal; float distance; struct or_material m; }) planes;
    int no_tree; /* renderer.h:8: USE_KD_TREE undefined → the brute-force branch of get_intersection_data */
} or_scene;
typedef struct or_material or_material;
typedef struct or_sphere or_sphere;
typedef struct or_plane or_plane;

/* OBJ loader: parse_chunks (OBJ_loader.cpp:32-176) applied to the whole file; chunks are newline-aligned and
 * joined in order (190-227), so one sequential pass yields the same arrays. A '\n' sentinel is written just
 * past the last byte as load_model_data does (OBJ_loader.cpp:331-332). */
static char* or_parse_face_vertex(char* cursor, int32_t* v, int32_t* vt, int32_t* vn)
{
    cursor = or_parse_int(cursor, v);
    if (*cursor == '/')
    {
        cursor++;
        if (*cursor == '/') { cursor++; cursor = or_parse_int(cursor, vn); }
        else
        {
            cursor = or_parse_int(cursor, vt);
            if (*cursor == '/') { cursor++; cursor = or_parse_int(cursor, vn); }
        }
    }
    return cursor;
}

static int or_load_obj(or_scene* s, const char* path)
{
    FILE* f = fopen(path, "rb");
    if (!f) return -1;
    fseek(f, 0, SEEK_END);
    long size = ftell(f);
    fseek(f, 0, SEEK_SET);
    char* buf = calloc(1, (size_t)size + 2);
    if (fread(buf, 1, (size_t)size, f) != (size_t)size) { fclose(f); free(buf); return -1; }
    fclose(f);
    buf[size] = '\n';
    char* cursor = buf;
    char* end = buf + size;
    while (cursor < end)
    {
        v3 tmp;
        if (*cursor == 'v')
        {
            cursor++;
            if (*cursor == ' ') { cursor = or_parse_vec3f(cursor, &tmp); OR_PUSH(s->vertices, tmp); }
            else if (*cursor == 't') { cursor++; cursor = or_parse_vec3f(cursor, &tmp); OR_PUSH(s->tex, tmp); }
            else if (*cursor == 'n') { cursor++; cursor = or_parse_vec3f(cursor, &tmp); OR_PUSH(s->normals, tmp); }
        }
        else if (*cursor == 'f')
        {
            cursor++;
            or_face fc;
            memset(&fc, 0, sizeof fc);
            for (int k = 0; k < 3; k++) cursor = or_parse_face_vertex(cursor, &fc.v[k], &fc.vt[k], &fc.vn[k]);
            OR_PUSH(s->faces, fc);
        }
        while (*cursor != '\n') cursor++; /* skip_to_new_line, parser.h:19-26 */
        cursor++;
    }
    free(buf);
    /* prep_model_data, OBJ_loader.cpp:229-267: negative indices relative to the FINAL array sizes, then -1 */
    for (size_t i = 0; i < s->faces.n; i++)
        for (int j = 0; j < 3; j++)
        {
            or_face* fc = &s->faces.p[i];
            if (fc->vt[j] < 0) fc->vt[j] = (int32_t)s->tex.n + fc->vt[j] + 1;
            if (fc->vn[j] < 0) fc->vn[j] = (int32_t)s->normals.n + fc->vn[j] + 1;
            if (fc->v[j] < 0) fc->v[j] = (int32_t)s->vertices.n + fc->v[j] + 1;
            fc->vt[j]--; fc->v[j]--; fc->vn[j]--;
        }
    return 0;
}

static or_box or_get_aabb(or_scene* s) /* model.h:41-61 */
{
    float x_max = -OR_MAX_FLOAT, x_min = OR_MAX_FLOAT, y_max = -OR_MAX_FLOAT, y_min = OR_MAX_FLOAT,
          z_max = -OR_MAX_FLOAT, z_min = OR_MAX_FLOAT;
    for (size_t i = 0; i < s->vertices.n; i++)
    {
        v3 v = s->vertices.p[i];
        x_max = f_max(x_max, v.x); x_min = f_min(x_min, v.x);
        y_max = f_max(y_max, v.y); y_min = f_min(y_min, v.y);
        z_max = f_max(z_max, v.z); z_min = f_min(z_min, v.z);
    }
    or_box b;
    b.max.x = x_max + OR_TOL; b.max.y = y_max + OR_TOL; b.max.z = z_max + OR_TOL;
    b.min.x = x_min - OR_TOL; b.min.y = y_min - OR_TOL; b.min.z = z_min - OR_TOL;
    return b;
}

static void or_resize_scale(or_scene* s, float new_max_scale) /* model.h:108-121 */
{
    v3 range = v3_sub(s->aabb.max, s->aabb.min);
    float max_scale = f_max(f_max(range.x, range.y), range.z);
    float k = new_max_scale / max_scale;
    for (size_t i = 0; i < s->vertices.n; i++) s->vertices.p[i] = v3_mul(s->vertices.p[i], k);
    s->aabb.max = v3_mul(s->aabb.max, k);
    s->aabb.min = v3_mul(s->aabb.min, k);
}

static void or_translate_to(or_scene* s, v3 new_center) /* model.h:136-152 */
{
    v3 old_center = v3_add(s->aabb.min, v3_div(v3_sub(s->aabb.max, s->aabb.min), 2.0f));
    v3 tr = v3_sub(new_center, old_center);
    for (size_t i = 0; i < s->vertices.n; i++) s->vertices.p[i] = v3_add(s->vertices.p[i], tr);
    s->aabb.max = v3_add(s->aabb.max, tr);
    s->aabb.min = v3_add(s->aabb.min, tr);
}

/* ---- tree builder, engine/renderer/kd_tree.cpp:3-64 (build_KD_tree) and 67-288 (build_oct_kd_tree) ----------- */
static inline int or_pt_inside(v3 p, const or_box* b) /* aabb.h:19-27, inclusive bounds */
{
    return (p.x >= b->min.x) && (p.x <= b->max.x) && (p.y >= b->min.y) && (p.y <= b->max.y) && (p.z >= b->min.z) &&
           (p.z <= b->max.z);
}
static inline int or_tri_inside(const or_tri* t, const or_box* b) /* kd_tree.cpp:10-17: ANY vertex inside */
{
    return or_pt_inside(t->a, b) || or_pt_inside(t->b, b) || or_pt_inside(t->c, b);
}
static inline float or_tri_area(const or_tri* t) /* kd_tree.cpp:3-8 */
{
    v3 ab = v3_sub(t->a, t->b), ac = v3_sub(t->a, t->c);
    return sqrtf(v3_mag2(v3_cross(ac, ab))) / 2;
}

static int or_build_tree(or_scene* s, int method, uint32_t cap)
{
    or_node root;
    memset(&root, 0, sizeof root);
    root.box = or_get_aabb(s); /* kd_tree.cpp:23 — recomputed from the (transformed) vertices */
    root.count = (uint32_t)s->faces.n;
    root.prims = calloc(root.count ? root.count : 1, sizeof(or_prim));
    for (uint32_t i = 0; i < root.count; i++)
    {
        or_face* fc = &s->faces.p[i];
        root.prims[i].tri.a = s->vertices.p[fc->v[0]];
        root.prims[i].tri.b = s->vertices.p[fc->v[1]];
        root.prims[i].tri.c = s->vertices.p[fc->v[2]];
        root.prims[i].face = i;
    }
    OR_PUSH(s->nodes, root);
    OR_VEC(uint32_t) stack;
    memset(&stack, 0, sizeof stack);
    OR_PUSH(stack, 0u);
    size_t total_refs = root.count;
    while (stack.n > 0)
    {
        uint32_t cur_i = stack.p[--stack.n]; /* LIFO: the last child pushed is processed first */
        or_node cur = s->nodes.p[cur_i];
        if (cur.count <= cap) { s->nodes.p[cur_i].children_start = 0; continue; }
        v3 dp;
        if (method == 0) /* CENTER, kd_tree.cpp:85-92 */
        {
            dp = v3_div(v3_sub(cur.box.max, cur.box.min), 2.0f);
            dp = v3_add(dp, cur.box.min);
        }
        else /* "SAH" = area-weighted centroid of triangle centres, kd_tree.cpp:93-114 */
        {
            v3 sum = { 0, 0, 0 };
            double sum_of_areas = 0;
            for (uint32_t i = 0; i < cur.count; i++)
            {
                const or_tri* t = &cur.prims[i].tri;
                v3 c = v3_div(v3_add(v3_add(t->a, t->b), t->c), 3.0f);
                float area = or_tri_area(t);
                sum = v3_add(sum, v3_mul(c, area));
                sum_of_areas += area;
            }
            dp = v3_div(sum, (float)sum_of_areas);
            if (!or_pt_inside(dp, &cur.box)) { s->nodes.p[cur_i].children_start = 0; continue; } /* stays a leaf */
        }
        const or_box p = cur.box;
        or_box cb[8]; /* order bb_left, bf_left, tb_left, tf_left, bb_right, bf_right, tb_right, tf_right (120-148) */
        for (int c = 0; c < 8; c++)
        {
            int xh = (c >> 2) & 1, yh = (c >> 1) & 1, zh = c & 1;
            cb[c].min.x = xh ? dp.x : p.min.x; cb[c].max.x = xh ? p.max.x : dp.x;
            cb[c].min.y = yh ? dp.y : p.min.y; cb[c].max.y = yh ? p.max.y : dp.y;
            cb[c].min.z = zh ? dp.z : p.min.z; cb[c].max.z = zh ? p.max.z : dp.z;
        }
        OR_VEC(or_prim) lists[8];
        memset(lists, 0, sizeof lists);
        for (uint32_t i = 0; i < cur.count; i++)
            for (int c = 0; c < 8; c++)
                if (or_tri_inside(&cur.prims[i].tri, &cb[c])) OR_PUSH(lists[c], cur.prims[i]);
        free(cur.prims);
        total_refs -= cur.count;
        s->nodes.p[cur_i].prims = 0;
        s->nodes.p[cur_i].count = 0;
        s->nodes.p[cur_i].children_start = (int32_t)s->nodes.n; /* kd_tree.cpp:259 */
        uint32_t first = (uint32_t)s->nodes.n;
        for (int c = 0; c < 8; c++)
        {
            or_node ch;
            memset(&ch, 0, sizeof ch);
            ch.box = cb[c];
            ch.count = (uint32_t)lists[c].n;
            ch.prims = lists[c].p;
            total_refs += ch.count;
            OR_PUSH(s->nodes, ch);
        }
        for (uint32_t c = 0; c < 8; c++) OR_PUSH(stack, first + c);
        if (total_refs > 400000000u || s->nodes.n > 40000000u) { free(stack.p); return -2; } /* reference has no guard */
    }
    free(stack.p);
    return 0;
}

/* ---- rays and boxes: ray.h:8-29, renderer.cpp:41-44, aabb.h:29-93 ------------------------------------------- */
typedef struct { v3 o, d, inv; int sign[3]; } or_ray;

static inline void or_ray_prepare(or_ray* r) /* renderer.cpp:41-44 */
{
    r->inv.x = 1 / r->d.x; r->inv.y = 1 / r->d.y; r->inv.z = 1 / r->d.z;
    r->sign[0] = r->inv.x < 0; r->sign[1] = r->inv.y < 0; r->sign[2] = r->inv.z < 0;
}
static inline const v3* or_bound(const or_box* b, int i) { return i ? &b->max : &b->min; }

static inline int or_check_box(const or_ray* r, const or_box* bb) /* check_ray_AABB_intersection, aabb.h:65-93 */
{
    float tmin, tmax, tymin, tymax, tzmin, tzmax;
    tmin = (or_bound(bb, r->sign[0])->x - r->o.x) * r->inv.x;
    tmax = (or_bound(bb, 1 - r->sign[0])->x - r->o.x) * r->inv.x;
    tymin = (or_bound(bb, r->sign[1])->y - r->o.y) * r->inv.y;
    tymax = (or_bound(bb, 1 - r->sign[1])->y - r->o.y) * r->inv.y;
    if ((tmin > tymax) || (tymin > tmax)) return 0;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    tzmin = (or_bound(bb, r->sign[2])->z - r->o.z) * r->inv.z;
    tzmax = (or_bound(bb, 1 - r->sign[2])->z - r->o.z) * r->inv.z;
    if ((tmin > tzmax) || (tzmin > tmax)) return 0;
    return 1;
}

static inline float or_entry_dist(const or_ray* r, const or_box* bb) /* get_ray_AABB_intersection, aabb.h:29-63 */
{
    float tmin, tmax, tymin, tymax, tzmin, tzmax;
    tmin = (or_bound(bb, r->sign[0])->x - r->o.x) * r->inv.x;
    tmax = (or_bound(bb, 1 - r->sign[0])->x - r->o.x) * r->inv.x;
    tymin = (or_bound(bb, r->sign[1])->y - r->o.y) * r->inv.y;
    tymax = (or_bound(bb, 1 - r->sign[1])->y - r->o.y) * r->inv.y;
    if ((tmin > tymax) || (tymin > tmax)) return 0;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    tzmin = (or_bound(bb, r->sign[2])->z - r->o.z) * r->inv.z;
    tzmax = (or_bound(bb, 1 - r->sign[2])->z - r->o.z) * r->inv.z;
    if ((tmin > tzmax) || (tzmin > tmax)) return 0;
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    if (tmin > 0) return tmin;
    else if (tmax > 0) return tmax;
    return 0;
}

/* get_triangle_ray_intersection_culled, model.h:75-103 */
static inline float or_tri_hit(const or_ray* r, const or_tri* tri, float* u, float* v)
{
    v3 ab = v3_sub(tri->b, tri->a), ac = v3_sub(tri->c, tri->a);
    v3 pvec = v3_cross(r->d, ac);
    float det = v3_dot(ab, pvec);
    if (det < OR_TOL) return 0;
    float det_inv = 1 / det;
    v3 tvec = v3_sub(r->o, tri->a);
    *u = v3_dot(tvec, pvec) * det_inv;
    if (*u < 0 || *u > 1) return 0;
    v3 qvec = v3_cross(tvec, ab);
    *v = v3_dot(r->d, qvec) * det_inv;
    if (*v < 0 || *u + *v > 1) return 0;
    return v3_dot(qvec, ac) * det_inv;
}

/* ---- traversal: traverse_oct_tree_new, kd_tree.cpp:337-465 ------------------------------ */
typedef struct { uint32_t node; float dist; } or_leaf;
typedef struct { uint32_t* hit_stack; or_leaf* leaves; /* leaves[-1] is the -MAX_FLOAT sentinel */ } or_scratch;
typedef struct { uint64_t box_tests, tri_tests, pass_det, pass_u, pass_v, inner_pops, leaves_listed, leaves_scanned; } or_counters;

static inline void or_scan_leaf(const or_ray* r, const or_node* n, float* closest, uint32_t* face, float* bu, float* bv,
                                int* hit, or_counters* ct)
{
    for (uint32_t i = 0; i < n->count; i++)
    {
        float u = 0, v = 0;
        float d = or_tri_hit(r, &n->prims[i].tri, &u, &v);
        if (ct) ct->tri_tests++;
        if (d < *closest && d > OR_TOL) { *closest = d; *face = n->prims[i].face; *bu = u; *bv = v; *hit = 1; }
    }
}

static float or_traverse(const or_scene* s, const or_ray* r, or_scratch* sc, uint32_t* face, float* bu, float* bv, or_counters* ct)
{
    const or_node* N = s->nodes.p;
    float closest = OR_MAX_FLOAT;
    int hit = 0;
    if (ct) ct->box_tests++;
    if (!or_check_box(r, &N[0].box)) return closest;
    if (N[0].children_start == 0) { or_scan_leaf(r, &N[0], &closest, face, bu, bv, &hit, ct); return closest; }
    int nleaf = 0, nstack = 1;
    sc->hit_stack[0] = 0;
    while (nstack > 0)
    {
        uint32_t cur = sc->hit_stack[--nstack];
        if (ct) ct->inner_pops++;
        uint32_t c0 = (uint32_t)N[cur].children_start;
        int nodes_hit = 0;
        for (int i = 0; i < 8 && nodes_hit <= 4; i++) /* at most 5 accepted children, kd_tree.cpp:374 */
        {
            const or_node* ch = &N[c0 + i];
            if (ct) ct->box_tests++;
            if (ch->children_start)
            {
                if (or_check_box(r, &ch->box)) { nodes_hit++; sc->hit_stack[nstack++] = c0 + i; }
            }
            else
            {
                float dis = or_entry_dist(r, &ch->box);
                if (dis > 0.0f)
                {
                    nodes_hit++;
                    or_leaf* end = sc->leaves + nleaf - 1; /* stable insertion from the back, kd_tree.cpp:389-409 */
                    while (dis < end->dist) { *(end + 1) = *end; end--; }
                    (end + 1)->node = c0 + i;
                    (end + 1)->dist = dis;
                    nleaf++;
                    if (ct) ct->leaves_listed++;
                }
            }
        }
    }
    for (int j = 0; j < nleaf; j++) /* front-to-back, stop after the first leaf that produced a hit (437-462) */
    {
        if (ct) ct->leaves_scanned++;
        or_scan_leaf(r, &N[sc->leaves[j].node], &closest, face, bu, bv, &hit, ct);
        if (hit) break;
    }
    return closest;
}

/* get_intersection_data, model branch only (renderer.cpp:34-57): returns 1 on a triangle hit */
/* The #else branch of renderer.cpp:49-82 (USE_KD_TREE undefined): model box first, then every face in file order;
 * a face is accepted when t > tolerance && t < the best so far (which starts at MAX_FLOAT, renderer.cpp:36). */
static float or_brute(const or_scene* s, const or_ray* r, uint32_t* face, float* bu, float* bv)
{
    float best = OR_MAX_FLOAT;
    if (or_entry_dist(r, &s->aabb))                                   /* renderer.cpp:59: non-zero float = true */
    {
        for (size_t j = 0; j < s->faces.n; j++)                       /* :61-79 */
        {
            or_tri tri;
            tri.a = s->vertices.p[s->faces.p[j].v[0]];
            tri.b = s->vertices.p[s->faces.p[j].v[1]];
            tri.c = s->vertices.p[s->faces.p[j].v[2]];
            float u, v;
            float t = or_tri_hit(r, &tri, &u, &v);
            if (t > OR_TOL && t < best) { best = t; *bu = u; *bv = v; *face = (uint32_t)j; }
        }
    }
    return best;
}

/* model query of get_intersection_data: the tree (USE_KD_TREE, the shipped configuration) or the brute-force branch */
static inline float or_model_query(const or_scene* s, const or_ray* r, or_scratch* sc, uint32_t* face, float* bu, float* bv,
                                   or_counters* ct)
{
    return s->no_tree ? or_brute(s, r, face, bu, bv) : or_traverse(s, r, sc, face, bu, bv, ct);
}

static inline int or_intersect(const or_scene* s, v3 o, v3 d, or_scratch* sc, float* t_out, uint32_t* face, float* u,
                               float* v, or_counters* ct)
{
    or_ray r;
    r.o = o; r.d = d;
    or_ray_prepare(&r);
    uint32_t f = OR_MISS; float bu = 0, bv = 0;
    float t = or_model_query(s, &r, sc, &f, &bu, &bv, ct);
    if (t > OR_TOL && t < OR_MAX_FLOAT) { *t_out = t; *face = f; *u = bu; *v = bv; return 1; }
    *t_out = OR_MAX_FLOAT; *face = OR_MISS; *u = 0; *v = 0;
    return 0;
}

/* ---- camera, camera.h:23-45 ---------------------------------------------------------------------------------- */
typedef struct { float aspect; v3 cz, cx, cy, eye, fc; float h_fov, hpw, hph; int W, H; } or_camera_t;

static void or_set_camera(or_camera_t* cm, int W, int H, v3 eye, v3 facing, float h_fov)
{
    cm->W = W; cm->H = H;
    cm->h_fov = h_fov;
    cm->aspect = W / (float)H;
    cm->eye = eye;
    facing = v3_normalize(facing);
    cm->fc = v3_add(eye, facing);
    cm->cz = v3_neg(facing);
    v3 y_axis = { 0.f, 1.f, 0.f };
    cm->cx = v3_normalize(v3_cross(y_axis, cm->cz));
    cm->cy = v3_normalize(v3_cross(cm->cz, cm->cx));
    cm->hpw = (0.5f * h_fov) / (float)W;
    cm->hph = 0.5f / (float)H;
}

static inline v3 or_pixel_dir(const or_camera_t* cm, float fx, float fy) /* renderer.cpp:342-343/350-351 + SetRay ray.h:24-29 */
{
    v3 pos = v3_add(v3_add(cm->fc, v3_mul(cm->cx, fx)), v3_mul(cm->cy, fy));
    return v3_normalize(v3_sub(pos, cm->eye));
}

static inline uint32_t or_pack(v3 c) /* clamp PL_math.h:430-436, rgb_float_to_byte :489, Set_Pixel texture.h:27-38 */
{
    c.x = f_max(0.0f, f_min(c.x, 1.0f)); c.y = f_max(0.0f, f_min(c.y, 1.0f)); c.z = f_max(0.0f, f_min(c.z, 1.0f));
    uint8_t r = (uint8_t)(c.x * 255.0f), g = (uint8_t)(c.y * 255.0f), b = (uint8_t)(c.z * 255.0f);
    return (uint32_t)b | ((uint32_t)g << 8) | ((uint32_t)r << 16);
}

/* ---- public C interface (same shapes as oracle/ref/ref_harness.cpp so one Python binding serves both) ------ */
typedef struct { int threads; int n; void (*fn)(void*, int, int, or_scratch*, or_counters*); void* arg; const or_scene* s;
                 volatile int next; or_counters total; pthread_mutex_t mu; } or_job;

static or_scratch or_scratch_new(const or_scene* s)
{
    or_scratch sc;
    sc.hit_stack = malloc((s->nodes.n + 8) * sizeof(uint32_t));
    or_leaf* base = malloc((s->nodes.n + 8) * sizeof(or_leaf));
    base[0].node = 0; base[0].dist = -OR_MAX_FLOAT; /* sentinel, renderer.cpp:386-387 */
    sc.leaves = base + 1;
    return sc;
}
static void or_scratch_free(or_scratch* sc) { free(sc->hit_stack); free(sc->leaves - 1); }

static void* or_worker(void* p)
{
    or_job* j = p;
    or_scratch sc = or_scratch_new(j->s);
    or_counters ct;
    memset(&ct, 0, sizeof ct);
    for (;;)
    {
        int i = __sync_fetch_and_add(&j->next, 16);
        if (i >= j->n) break;
        j->fn(j->arg, i, i + 16 > j->n ? j->n : i + 16, &sc, &ct);
    }
    or_scratch_free(&sc);
    pthread_mutex_lock(&j->mu);
    uint64_t* a = (uint64_t*)&j->total; uint64_t* b = (uint64_t*)&ct;
    for (size_t k = 0; k < sizeof(or_counters) / 8; k++) a[k] += b[k];
    pthread_mutex_unlock(&j->mu);
    return 0;
}
static or_counters or_parallel(const or_scene* s, int n, int threads, void (*fn)(void*, int, int, or_scratch*, or_counters*), void* arg)
{
    or_job j;
    memset(&j, 0, sizeof j);
    j.threads = threads < 1 ? 1 : threads; j.n = n; j.fn = fn; j.arg = arg; j.s = s;
    pthread_mutex_init(&j.mu, 0);
    if (j.threads == 1) or_worker(&j);
    else
    {
        pthread_t* th = malloc(sizeof(pthread_t) * j.threads);
        for (int t = 0; t < j.threads; t++) pthread_create(&th[t], 0, or_worker, &j);
        for (int t = 0; t < j.threads; t++) pthread_join(th[t], 0);
        free(th);
    }
    return j.total;
}

void* or_scene_create(const char* obj_path, int loader_threads, float rescale, float tx, float ty, float tz, int method, unsigned cap)
{
    (void)loader_threads;
    or_scene* s = calloc(1, sizeof(or_scene));
    if (or_load_obj(s, obj_path) != 0) { free(s); return 0; }
    s->aabb = or_get_aabb(s);                       /* app.cpp:68 */
    if (rescale > 0.0f) or_resize_scale(s, rescale); /* app.cpp:69 (commented out upstream; used for the small meshes) */
    v3 c = { tx, ty, tz };
    or_translate_to(s, c);                          /* app.cpp:72 */
    v3 sky = { 0.3f, 0.4f, 0.5f }, mdl = { 0.4f, 0.2f, 0.2f };
    s->sky_emission = sky; s->model_emission = mdl;
    v3 sky_r = { 0.2f, 0.3f, 0.4f }, mdl_r = { 0.92f, 0.5f, 0.0f };   /* app.cpp:91,96 */
    s->sky_mat.emission = sky; s->sky_mat.reflection = sky_r; s->sky_mat.scatter = 0.3f;
    s->model_mat.emission = mdl; s->model_mat.reflection = mdl_r; s->model_mat.scatter = 0.3f;
    if (or_build_tree(s, method, cap) != 0) { return 0; }
    return s;
}

/* 0 → every query takes the no-tree branch (the reference compiled without USE_KD_TREE, renderer.h:8) */
void or_scene_use_kd_tree(void* h, int use) { ((or_scene*)h)->no_tree = !use; }

void or_scene_destroy(void* h)
{
    or_scene* s = h;
    for (size_t i = 0; i < s->nodes.n; i++) free(s->nodes.p[i].prims);
    free(s->nodes.p); free(s->vertices.p); free(s->normals.p); free(s->tex.p); free(s->faces.p);
    free(s->spheres.p); free(s->planes.p); free(s);
}

void or_counts(void* h, unsigned* c)
{
    or_scene* s = h;
    c[0] = (unsigned)s->vertices.n; c[1] = (unsigned)s->normals.n; c[2] = (unsigned)s->tex.n; c[3] = (unsigned)s->faces.n;
    c[4] = (unsigned)s->nodes.n;
    unsigned refs = 0, inner = 0, maxleaf = 0;
    for (size_t i = 0; i < s->nodes.n; i++)
    {
        if (s->nodes.p[i].children_start) inner++;
        else { refs += s->nodes.p[i].count; if (s->nodes.p[i].count > maxleaf) maxleaf = s->nodes.p[i].count; }
    }
    c[5] = refs; c[6] = inner; c[7] = maxleaf;
}

void or_model_dump(void* h, float* vertices, float* normals, float* tex, int* fv, int* fvt, int* fvn, float* aabb6)
{
    or_scene* s = h;
    if (vertices) memcpy(vertices, s->vertices.p, s->vertices.n * 12);
    if (normals) memcpy(normals, s->normals.p, s->normals.n * 12);
    if (tex) memcpy(tex, s->tex.p, s->tex.n * 12);
    for (size_t i = 0; i < s->faces.n; i++)
        for (int j = 0; j < 3; j++)
        {
            if (fv) fv[3 * i + j] = s->faces.p[i].v[j];
            if (fvt) fvt[3 * i + j] = s->faces.p[i].vt[j];
            if (fvn) fvn[3 * i + j] = s->faces.p[i].vn[j];
        }
    if (aabb6) memcpy(aabb6, &s->aabb, 24);
}

void or_tree_dump(void* h, float* boxes, int* children_start, unsigned* prim_count, unsigned* prim_first, float* tri9, unsigned* face)
{
    or_scene* s = h;
    unsigned off = 0;
    for (size_t i = 0; i < s->nodes.n; i++)
    {
        const or_node* n = &s->nodes.p[i];
        memcpy(boxes + 6 * i, &n->box, 24);
        children_start[i] = n->children_start;
        unsigned cnt = n->children_start ? 0u : n->count;
        prim_count[i] = cnt; prim_first[i] = off;
        for (unsigned k = 0; k < cnt; k++) { memcpy(tri9 + 9 * (size_t)(off + k), &n->prims[k].tri, 36); face[off + k] = n->prims[k].face; }
        off += cnt;
    }
}

void or_camera(int W, int H, const float* eye, const float* facing, float hfov, float* out19)
{
    or_camera_t cm;
    v3 e = { eye[0], eye[1], eye[2] }, f = { facing[0], facing[1], facing[2] };
    or_set_camera(&cm, W, H, e, f, hfov);
    memcpy(out19, &cm.aspect, sizeof(float) * 19);
}

typedef struct { const or_scene* s; or_camera_t cm; unsigned* face; float* t; float* uv; unsigned* bgrx; } or_primary_arg;
static void or_primary_rows(void* a_, int y0, int y1, or_scratch* sc, or_counters* ct)
{
    or_primary_arg* a = a_;
    const or_camera_t* cm = &a->cm;
    for (int y = y0; y < y1; y++)
    {
        float film_y = -1.0f + 2.0f * ((float)y / (float)cm->H); /* renderer.cpp:317 */
        for (int x = 0; x < cm->W; x++)
        {
            float film_x = (-1.0f + 2.0f * ((float)x / (float)cm->W)) * cm->h_fov * cm->aspect; /* :329 */
            v3 d = or_pixel_dir(cm, film_x, film_y);
            float t, u, v; uint32_t f;
            int hit = or_intersect(a->s, cm->eye, d, sc, &t, &f, &u, &v, ct);
            size_t p = (size_t)y * cm->W + x;
            if (a->face) a->face[p] = f;
            if (a->t) a->t[p] = t;
            if (a->uv) { a->uv[2 * p] = u; a->uv[2 * p + 1] = v; }
            if (a->bgrx) a->bgrx[p] = or_pack(hit ? a->s->model_emission : a->s->sky_emission); /* cast_ray @ bounce 1 */
        }
    }
}

/* counters_out (8 x u64, may be NULL): box_tests, tri_tests, -, -, -, inner_pops, leaves_listed, leaves_scanned */
void or_trace_primary_ex(void* h, int W, int H, const float* eye, const float* facing, float hfov, int threads,
                         unsigned* face, float* t, float* uv, unsigned* bgrx, unsigned long long* counters_out)
{
    or_primary_arg a;
    a.s = h; a.face = face; a.t = t; a.uv = uv; a.bgrx = bgrx;
    v3 e = { eye[0], eye[1], eye[2] }, f = { facing[0], facing[1], facing[2] };
    or_set_camera(&a.cm, W, H, e, f, hfov);
    or_counters c = or_parallel(a.s, H, threads, or_primary_rows, &a);
    if (counters_out) memcpy(counters_out, &c, sizeof c);
}
void or_trace_primary(void* h, int W, int H, const float* eye, const float* facing, float hfov, int threads,
                      unsigned* face, float* t, float* uv, unsigned* bgrx)
{
    or_trace_primary_ex(h, W, H, eye, facing, hfov, threads, face, t, uv, bgrx, 0);
}

/* Verification aid for the product's film-space culling (tests/test_view_bounds.py): every (pixel, leaf, position in
 * the leaf) whose triangle test is ACCEPTED (model.h:75-103 returns t, and t > tolerance, t < MAX_FLOAT) in ANY leaf the
 * traversal lists for the pixel's ray (kd_tree.cpp:363-435) — without the early exit of :457-460, so it is a superset
 * of what the reference's scan can ever accept. */
typedef struct { const or_scene* s; or_camera_t cm; long long cap; long long* count; unsigned* pixel; unsigned* node; unsigned* pos; } or_acc_arg;
static void or_accepted_rows(void* a_, int y0, int y1, or_scratch* sc, or_counters* ct)
{
    (void)ct;
    or_acc_arg* a = a_;
    const or_camera_t* cm = &a->cm;
    const or_node* N = a->s->nodes.p;
    for (int y = y0; y < y1; y++)
    {
        float film_y = -1.0f + 2.0f * ((float)y / (float)cm->H);
        for (int x = 0; x < cm->W; x++)
        {
            float film_x = (-1.0f + 2.0f * ((float)x / (float)cm->W)) * cm->h_fov * cm->aspect;
            or_ray r;
            r.o = cm->eye; r.d = or_pixel_dir(cm, film_x, film_y);
            or_ray_prepare(&r);
            if (!or_check_box(&r, &N[0].box)) continue;
            int nleaf = 0;
            if (N[0].children_start == 0) { sc->leaves[0].node = 0; nleaf = 1; }
            else
            {
                int nstack = 1;
                sc->hit_stack[0] = 0;
                while (nstack > 0)
                {
                    uint32_t cur = sc->hit_stack[--nstack];
                    uint32_t c0 = (uint32_t)N[cur].children_start;
                    int nodes_hit = 0;
                    for (int i = 0; i < 8 && nodes_hit <= 4; i++)
                    {
                        const or_node* ch = &N[c0 + i];
                        if (ch->children_start) { if (or_check_box(&r, &ch->box)) { nodes_hit++; sc->hit_stack[nstack++] = c0 + i; } }
                        else if (or_entry_dist(&r, &ch->box) > 0.0f) { nodes_hit++; sc->leaves[nleaf++].node = c0 + i; }
                    }
                }
            }
            for (int j = 0; j < nleaf; j++)
            {
                const or_node* n = &N[sc->leaves[j].node];
                for (uint32_t i = 0; i < n->count; i++)
                {
                    float u = 0, v = 0;
                    float d = or_tri_hit(&r, &n->prims[i].tri, &u, &v);
                    if (d < OR_MAX_FLOAT && d > OR_TOL)
                    {
                        long long k = __sync_fetch_and_add(a->count, 1);
                        if (k < a->cap) { a->pixel[k] = (unsigned)((size_t)y * cm->W + x); a->node[k] = sc->leaves[j].node; a->pos[k] = i; }
                    }
                }
            }
        }
    }
}
long long or_accepted_hits(void* h, int W, int H, const float* eye, const float* facing, float hfov, int threads, long long cap,
                           unsigned* pixel, unsigned* node, unsigned* pos)
{
    or_acc_arg a;
    long long count = 0;
    a.s = h; a.cap = cap; a.count = &count; a.pixel = pixel; a.node = node; a.pos = pos;
    v3 e = { eye[0], eye[1], eye[2] }, f = { facing[0], facing[1], facing[2] };
    or_set_camera(&a.cm, W, H, e, f, hfov);
    or_parallel(a.s, H, threads, or_accepted_rows, &a);
    return count;
}

typedef struct { const or_scene* s; const float* rays; long long n; unsigned* face; float* t; float* uv; } or_rays_arg;
static void or_rays_blocks(void* a_, int b0, int b1, or_scratch* sc, or_counters* ct)
{
    or_rays_arg* a = a_;
    for (long long i = (long long)b0 * 4096; i < (long long)b1 * 4096 && i < a->n; i++)
    {
        v3 o = { a->rays[6 * i], a->rays[6 * i + 1], a->rays[6 * i + 2] };
        v3 d = { a->rays[6 * i + 3], a->rays[6 * i + 4], a->rays[6 * i + 5] };
        float u, v;
        if (or_intersect(a->s, o, d, sc, &a->t[i], &a->face[i], &u, &v, ct) && a->uv) { a->uv[2 * i] = u; a->uv[2 * i + 1] = v; }
    }
}
void or_trace_rays(void* h, long long n, const float* rays, int threads, unsigned* face, float* t)
{
    or_rays_arg a = { h, rays, n, face, t, 0 };
    or_parallel(a.s, (int)((n + 4095) / 4096), threads, or_rays_blocks, &a);
}
/* same with TriangleIntersectionData's u, v of the winning triangle (model.h:18-22, kd_tree.cpp:447-452); a miss leaves uv untouched */
void or_trace_rays_uv(void* h, long long n, const float* rays, int threads, unsigned* face, float* t, float* uv)
{
    or_rays_arg a = { h, rays, n, face, t, uv };
    or_parallel(a.s, (int)((n + 4095) / 4096), threads, or_rays_blocks, &a);
}

static inline uint64_t or_splitmix64(uint64_t x)
{
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}

/* Jittered pixel loop, renderer.cpp:334-347,358-365, bounce_limit = 1 only: per sample x-jitter then y-jitter,
 * then cast_ray (213-262), which at bounce 1 returns the sky emission (no draws) or the model emission after
 * three rand_bi draws (renderer.cpp:246). Per-pixel stream as defined in SURVEY §8(d) C3. */
typedef struct { const or_scene* s; or_camera_t cm; unsigned spp; uint64_t seed; unsigned* bgrx; unsigned* hits; } or_jit_arg;
static void or_jitter_rows(void* a_, int y0, int y1, or_scratch* sc, or_counters* ct)
{
    or_jit_arg* a = a_;
    const or_camera_t* cm = &a->cm;
    for (int y = y0; y < y1; y++)
    {
        float film_y = -1.0f + 2.0f * ((float)y / (float)cm->H);
        for (int x = 0; x < cm->W; x++)
        {
            uint64_t p = (uint64_t)y * cm->W + x;
            or_rng rng = { or_splitmix64(a->seed ^ p), 2 * p + 1 };
            float film_x = (-1.0f + 2.0f * ((float)x / (float)cm->W)) * cm->h_fov * cm->aspect;
            v3 col = { 0, 0, 0 };
            unsigned hits = 0;
            for (unsigned i = 0; i < a->spp; i++)
            {
                float x_off = or_rand_bi(&rng) * cm->hpw + film_x;
                float y_off = or_rand_bi(&rng) * cm->hph + film_y;
                v3 d = or_pixel_dir(cm, x_off, y_off);
                float t, u, v; uint32_t f;
                if (or_intersect(a->s, cm->eye, d, sc, &t, &f, &u, &v, ct))
                {
                    or_rand_bi(&rng); or_rand_bi(&rng); or_rand_bi(&rng);
                    col = v3_add(col, a->s->model_emission);
                    hits++;
                }
                else col = v3_add(col, a->s->sky_emission);
            }
            col = v3_div(col, (float)a->spp);
            a->bgrx[p] = or_pack(col);
            if (a->hits) a->hits[p] = hits;
        }
    }
}
void or_render_jitter(void* h, int W, int H, unsigned spp, int bounce, unsigned long long seed, const float* eye,
                      const float* facing, float hfov, int threads, unsigned* bgrx, unsigned* hits)
{
    (void)bounce;
    or_jit_arg a;
    a.s = h; a.spp = spp; a.seed = seed; a.bgrx = bgrx; a.hits = hits;
    v3 e = { eye[0], eye[1], eye[2] }, f = { facing[0], facing[1], facing[2] };
    or_set_camera(&a.cm, W, H, e, f, hfov);
    or_parallel(a.s, H, threads, or_jitter_rows, &a);
}

/* ---- C5 (BASELINE.json configs[4]): incoherent random-direction rays, `spp` per pixel index, traced through the
 * model query of get_intersection_data (renderer.cpp:41-57). The reference has no such workload of its own; the harness
 * defines the rays from the reference's functions: ray i = p*spp + s draws from the PCG32 stream (PL_math.h:506-541)
 * {state = splitmix64(seed ^ i), stream = 2i + 1} six rand_bi values b0..b5 (in that order) and is
 *     origin o = c + normalize((b0,b1,b2)) * R          (normalize: PL_math.h:387-392)
 *     target   = c + (0.5 * extent) (had) (b3,b4,b5)    (hadamard: PL_math.h:419)
 *     SetRay(o, target)                                  (ray.h:24-29)
 * with c = (lo + hi) * 0.5, extent = hi - lo of Model::surrounding_aabb and R = 1.5 * mag(extent) (PL_math.h:381-385), all in
 * f32 in the order written. Pixel colour = sequential f32 sum of the samples' flat colours (sky / model emission) / spp,
 * clamp, truncate, pack (renderer.cpp:358-365). */
typedef struct { const or_scene* s; long long p0, p1; unsigned spp; uint64_t seed; unsigned* face; float* t; unsigned* bgrx; float* rays; } or_c5_arg;
static void or_c5_ray(const or_scene* s, uint64_t seed, uint64_t i, v3* o, v3* d)
{
    or_rng rng = { or_splitmix64(seed ^ i), 2 * i + 1 };
    v3 b, q;
    b.x = or_rand_bi(&rng); b.y = or_rand_bi(&rng); b.z = or_rand_bi(&rng);
    q.x = or_rand_bi(&rng); q.y = or_rand_bi(&rng); q.z = or_rand_bi(&rng);
    const v3 lo = s->aabb.min, hi = s->aabb.max;
    const v3 c = v3_mul(v3_add(lo, hi), 0.5f), ext = v3_sub(hi, lo);
    const float R = 1.5f * sqrtf((ext.x * ext.x + ext.y * ext.y) + ext.z * ext.z);
    *o = v3_add(c, v3_mul(v3_normalize(b), R));
    v3 he = v3_mul(ext, 0.5f), tg;
    tg.x = c.x + he.x * q.x; tg.y = c.y + he.y * q.y; tg.z = c.z + he.z * q.z;
    *d = v3_normalize(v3_sub(tg, *o));
}
static void or_c5_blocks(void* a_, int b0, int b1, or_scratch* sc, or_counters* ct)
{
    or_c5_arg* a = a_;
    for (long long p = a->p0 + (long long)b0 * 64; p < a->p0 + (long long)b1 * 64 && p < a->p1; p++)
    {
        v3 col = { 0, 0, 0 };
        for (unsigned k = 0; k < a->spp; k++)
        {
            const uint64_t i = (uint64_t)p * a->spp + k;
            const long long j = (p - a->p0) * a->spp + k;
            v3 o, d;
            or_c5_ray(a->s, a->seed, i, &o, &d);
            if (a->rays) { float* r = a->rays + 6 * j; r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = d.x; r[4] = d.y; r[5] = d.z; }
            float t, u, v; uint32_t f;
            const int hit = or_intersect(a->s, o, d, sc, &t, &f, &u, &v, ct);
            if (a->face) a->face[j] = f;
            if (a->t) a->t[j] = t;
            col = v3_add(col, hit ? a->s->model_emission : a->s->sky_emission);
        }
        if (a->bgrx) a->bgrx[p - a->p0] = or_pack(v3_div(col, (float)a->spp));
    }
}
/* pixels [p0, p1): face / t / rays per ray ((p1-p0)*spp entries, any may be NULL), bgrx per pixel */
void or_trace_c5(void* h, long long p0, long long p1, unsigned spp, unsigned long long seed, int threads, unsigned* face, float* t,
                 unsigned* bgrx, float* rays)
{
    or_c5_arg a = { h, p0, p1, spp, seed, face, t, bgrx, rays };
    or_parallel(a.s, (int)((p1 - p0 + 63) / 64), threads, or_c5_blocks, &a);
}

/* ---- multi-bounce path: get_intersection_data (renderer.cpp:34-160) with spheres and planes, cast_ray (213-262) ---- */
static or_material or_mat7(const float* m7)
{
    or_material m;
    m.emission.x = m7[0]; m.emission.y = m7[1]; m.emission.z = m7[2];
    m.reflection.x = m7[3]; m.reflection.y = m7[4]; m.reflection.z = m7[5];
    m.scatter = m7[6];
    return m;
}
int or_scene_add_sphere(void* h, const float* center, float radius, const float* m7)
{
    or_scene* s = h;
    or_sphere sp;
    sp.center.x = center[0]; sp.center.y = center[1]; sp.center.z = center[2];
    sp.radius = radius; sp.m = or_mat7(m7);
    OR_PUSH(s->spheres, sp);
    return 0;
}
int or_scene_add_plane(void* h, const float* normal, float distance, const float* m7)
{
    or_scene* s = h;
    or_plane pl;
    v3 n = { normal[0], normal[1], normal[2] };
    pl.normal = v3_normalize(n);               /* prep_scene, renderer.cpp:267-270 */
    pl.distance = distance; pl.m = or_mat7(m7);
    OR_PUSH(s->planes, pl);
    return 0;
}

static inline float or_sphere_hit(v3 o, v3 d, const or_sphere* sp) /* get_sphere_ray_intersection, sphere.h:12-39 */
{
    v3 p_to_c = v3_sub(o, sp->center);
    float p_to_c_sqr = v3_mag2(p_to_c);
    float b = 2 * v3_dot(d, p_to_c);
    float b_sqr = b * b;
    float c = p_to_c_sqr - sp->radius * sp->radius;
    float dmt = b_sqr - (4 * c);
    if (dmt < 0) return 0;
    float ta = (-b + sqrtf(dmt)) * 0.5f;        /* sqroot = _mm_sqrt_ss, PL_math.h:352-356: IEEE */
    float tb = (-b - sqrtf(dmt)) * 0.5f;
    if (ta <= 0 && tb <= 0) return 0;
    if (tb > 0) return tb;
    return ta;
}
static inline float or_plane_hit(v3 o, v3 d, const or_plane* pl) /* get_plane_ray_intersection, plane.h:12-22 */
{
    float denom = v3_dot(pl->normal, d);
    if (denom > -OR_TOL && denom < OR_TOL) return 0;
    return (pl->distance - v3_dot(o, pl->normal)) / denom;
}

/* kinds: 0 sky, 1 triangle, 2 sphere, 3 plane (ObjectType, renderer.cpp:12-15) */
typedef struct { int kind; float t; v3 normal; const or_material* mat; uint32_t face; } or_hit;

static void or_intersect_full(const or_scene* s, v3 o, v3 d, or_scratch* sc, or_hit* out, or_counters* ct)
{
    float dist = OR_MAX_FLOAT;
    const or_sphere* near_sphere = 0; const or_plane* near_plane = 0; int near_model = 0;
    uint32_t face = OR_MISS; float u = 0, v = 0;
    {
        or_ray r; r.o = o; r.d = d; or_ray_prepare(&r);
        uint32_t f = OR_MISS; float bu = 0, bv = 0;
        float t = or_model_query(s, &r, sc, &f, &bu, &bv, ct);
        if (t > OR_TOL && t < dist) { dist = t; face = f; u = bu; v = bv; near_model = 1; }      /* :52-57 */
    }
    for (size_t i = 0; i < s->spheres.n; i++)                                                     /* :86-96 */
    {
        float t = or_sphere_hit(o, d, &s->spheres.p[i]);
        if (t > OR_TOL && t < dist) { dist = t; near_sphere = &s->spheres.p[i]; }
    }
    for (size_t i = 0; i < s->planes.n; i++)                                                      /* :97-107 */
    {
        float t = or_plane_hit(o, d, &s->planes.p[i]);
        if (t > OR_TOL && t < dist) { near_plane = &s->planes.p[i]; dist = t; }
    }
    out->t = dist; out->face = OR_MISS;
    v3 zero = { 0, 0, 0 };
    out->normal = zero;
    if (near_plane) { out->kind = 3; out->normal = near_plane->normal; out->mat = &near_plane->m; }             /* :111-116 */
    else if (near_sphere)                                                                                     /* :117-122 */
    {
        out->kind = 2;
        out->normal = v3_sub(v3_add(o, v3_mul(d, dist)), near_sphere->center);   /* ray.at(t) - center, ray.h:11-14 */
        out->mat = &near_sphere->m;
    }
    else if (near_model)                                                                                      /* :124-148 */
    {
        out->kind = 1; out->face = face;
        const or_face* fc = &s->faces.p[face];
        if (s->normals.n > 0)
        {
            v3 na = s->normals.p[fc->vn[0]], nb = s->normals.p[fc->vn[1]], nc = s->normals.p[fc->vn[2]];
            out->normal = v3_add(v3_add(v3_mul(na, 1 - u - v), v3_mul(nb, u)), v3_mul(nc, v));
        }
        else
        {
            v3 ab = v3_sub(s->vertices.p[fc->v[0]], s->vertices.p[fc->v[1]]);
            v3 ac = v3_sub(s->vertices.p[fc->v[0]], s->vertices.p[fc->v[2]]);
            out->normal = v3_cross(ab, ac);
        }
        out->mat = &s->model_mat;
    }
    else { out->kind = 0; out->mat = &s->sky_mat; }                                                           /* :149-155 */
    out->normal = v3_normalize(out->normal);                                                                  /* :157 */
}

static inline v3 v3_had(v3 a, v3 b) { v3 r = { a.x * b.x, a.y * b.y, a.z * b.z }; return r; } /* PL_math.h:419 */

static v3 or_cast_ray(const or_scene* s, v3 o, v3 d, int bounce_limit, long long* casts, or_rng* rng, or_scratch* sc, or_counters* ct)
{
    v3 color = { 0, 0, 0 }, weight = { 1.0f, 1.0f, 1.0f };
    int i;
    for (i = 0; i < bounce_limit; i++)
    {
        or_hit id;
        or_intersect_full(s, o, d, sc, &id, ct);
        if (id.kind == 0) { color = v3_add(color, v3_had(weight, id.mat->emission)); break; }   /* :224-228 */
        float attenuation = v3_dot(v3_neg(d), id.normal);
        if (attenuation < 0) { id.normal = v3_neg(id.normal); attenuation = 0; }
        v3 pure = v3_sub(d, v3_mul(id.normal, 2 * v3_dot(d, id.normal)));                       /* :238-243 */
        pure = v3_normalize(pure);
        v3 rb;
        rb.x = or_rand_bi(rng); rb.y = or_rand_bi(rng); rb.z = or_rand_bi(rng);                /* :246, braces: left to right */
        rb = v3_add(rb, id.normal);
        rb = v3_normalize(rb);
        o = v3_add(o, v3_mul(d, id.t));                                                         /* :251 */
        d = v3_add(v3_mul(v3_sub(pure, rb), id.mat->scatter), rb);                              /* lerp, PL_math.h:439-442 */
        d = v3_normalize(d);
        color = v3_add(color, v3_had(weight, id.mat->emission));                                /* :257-258 */
        weight = v3_had(weight, v3_mul(id.mat->reflection, attenuation));
    }
    *casts += i;
    return color;
}

typedef struct { const or_scene* s; or_camera_t cm; unsigned spp; int aa, bounce; uint64_t seed; unsigned* bgrx; float* t;
                 unsigned* prim; unsigned* kind; long long casts; pthread_mutex_t mu; } or_path_arg;
static void or_path_rows(void* a_, int y0, int y1, or_scratch* sc, or_counters* ct)
{
    or_path_arg* a = a_;
    const or_camera_t* cm = &a->cm;
    long long casts = 0;
    for (int y = y0; y < y1; y++)
    {
        float film_y = -1.0f + 2.0f * ((float)y / (float)cm->H);
        for (int x = 0; x < cm->W; x++)
        {
            uint64_t p = (uint64_t)y * cm->W + x;
            or_rng rng = { or_splitmix64(a->seed ^ p), 2 * p + 1 };
            float film_x = (-1.0f + 2.0f * ((float)x / (float)cm->W)) * cm->h_fov * cm->aspect;
            v3 col = { 0, 0, 0 };
            v3 d = { 0, 0, 0 };
            if (a->aa)
                for (unsigned i = 0; i < a->spp; i++)                                           /* renderer.cpp:336-347 */
                {
                    float x_off = or_rand_bi(&rng) * cm->hpw + film_x;
                    float y_off = or_rand_bi(&rng) * cm->hph + film_y;
                    d = or_pixel_dir(cm, x_off, y_off);
                    col = v3_add(col, or_cast_ray(a->s, cm->eye, d, a->bounce, &casts, &rng, sc, ct));
                }
            else
            {
                d = or_pixel_dir(cm, film_x, film_y);                                          /* :349-356 */
                for (unsigned i = 0; i < a->spp; i++) col = v3_add(col, or_cast_ray(a->s, cm->eye, d, a->bounce, &casts, &rng, sc, ct));
            }
            col = v3_div(col, (float)a->spp);
            a->bgrx[p] = or_pack(col);
            if (a->t || a->prim || a->kind)
            {
                or_hit id;
                or_intersect_full(a->s, cm->eye, d, sc, &id, ct);
                if (a->t) a->t[p] = id.t;
                if (a->prim) a->prim[p] = id.kind == 1 ? id.face : OR_MISS;
                if (a->kind) a->kind[p] = (unsigned)id.kind;
            }
        }
    }
    pthread_mutex_lock(&a->mu);
    a->casts += casts;
    pthread_mutex_unlock(&a->mu);
}
void or_render_path(void* h, int W, int H, unsigned spp, int aa, int bounce, unsigned long long seed, const float* eye,
                    const float* facing, float hfov, int threads, unsigned* bgrx, float* t, unsigned* prim, unsigned* kind,
                    long long* casts_out)
{
    or_path_arg a;
    memset(&a, 0, sizeof a);
    a.s = h; a.spp = spp; a.aa = aa; a.bounce = bounce; a.seed = seed; a.bgrx = bgrx; a.t = t; a.prim = prim; a.kind = kind;
    pthread_mutex_init(&a.mu, 0);
    v3 e = { eye[0], eye[1], eye[2] }, f = { facing[0], facing[1], facing[2] };
    or_set_camera(&a.cm, W, H, e, f, hfov);
    or_parallel(a.s, H, threads, or_path_rows, &a);
    if (casts_out) *casts_out = a.casts;
}

void or_pcg32(unsigned long long state, unsigned long long stream, int n, unsigned* out)
{
    or_rng s = { state, stream };
    for (int i = 0; i < n; i++) out[i] = or_random_u32(&s);
}

/* BMP_FILE_FORMAT::Write_To_File (engine/tools/texture.cpp:66-115) under Write_To_File(Texture&, name) (:147-158):
 * 14-byte packed file header + 56-byte DIB header (compression 3 = bit fields, ppm 197 x 39, masks R/G/B/0) + the
 * BGRX rows bottom-up exactly as they sit in the bitmap buffer. The file is "<file_name>_<id>.bmp" for the first id
 * whose file does not exist yet (pl_create_file = CreateFileA(CREATE_NEW), PL/Win32/pl_utils_win32.cpp:143-158); the
 * name is formatted into length + 8 characters (texture.cpp:101). Returns 1 like the reference (0 on a bad argument). */
int or_write_bmp(const unsigned* pixels, unsigned W, unsigned H, const char* file_name)
{
    if (!pixels || !file_name) return 0;
    unsigned char h[70];
    memset(h, 0, sizeof h);
    const unsigned size = 4u * W * H;                      /* BitmapBuffer::size(), texture.h:17 */
    const unsigned total = 14u + 56u + size, off = 14u + 56u, hdr = 56u;
    const unsigned short planes = 1, bpp = 32;
    const unsigned comp = 3, ppx = 197, ppy = 39, rm = 0x00FF0000u, gm = 0x0000FF00u, bm = 0x000000FFu;
    h[0] = 'B'; h[1] = 'M';
    memcpy(h + 2, &total, 4); memcpy(h + 10, &off, 4);
    memcpy(h + 14, &hdr, 4); memcpy(h + 18, &W, 4); memcpy(h + 22, &H, 4); memcpy(h + 26, &planes, 2); memcpy(h + 28, &bpp, 2);
    memcpy(h + 30, &comp, 4); memcpy(h + 34, &size, 4); memcpy(h + 38, &ppx, 4); memcpy(h + 42, &ppy, 4);
    memcpy(h + 54, &rm, 4); memcpy(h + 58, &gm, 4); memcpy(h + 62, &bm, 4);
    const size_t len = strlen(file_name);
    if (len + 8 > 1024) return 0;
    char name[1024];
    FILE* f = 0;
    for (unsigned id = 0; !f; id++)
    {
        snprintf(name, len + 8, "%s%c%u%s", file_name, '_', id, ".bmp");
        f = fopen(name, "wbx");
        if (!f && errno != EEXIST) return 1;               /* reference: no handle → nothing written, still TRUE */
    }
    fwrite(h, 1, 70, f);
    fwrite(pixels, 1, size, f);
    fclose(f);
    return 1;
}
