/*
 * terra_oracle.c — TEST INFRASTRUCTURE ONLY. See terra_oracle.h for what this is and how it is
 * pinned. Plain C11, gcc -O2, x86-64 baseline (no FMA contraction), libm only.
 * Citations are to the reference checkout (src/..., include/...).
 */
#include "terra_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ small vector helpers
 * Operation order follows include/TerraMath.inl; min/max are the reference's ternaries. */
typedef TerraFloat3 f3;
#define OPI  3.1416926535f          /* terra_PI, include/TerraMath.h:17 */
#define OEPS 1e-4                   /* terra_Epsilon (double), :18 */

static f3 mk(float x, float y, float z) { f3 r = { x, y, z }; return r; }
static f3 add(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
static f3 sub(f3 a, f3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
static f3 scl(f3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
static f3 had(f3 a, f3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
static f3 neg(f3 a) { return mk(-a.x, -a.y, -a.z); }
static float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static f3 cross(f3 a, f3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static float len(f3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
static float sqlen(f3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
static f3 unit(f3 a) { float l = len(a); return mk(a.x / l, a.y / l, a.z / l); }       /* TerraMath.inl:146-153 */
static float fmx(float a, float b) { return a > b ? a : b; }                            /* :171-173 */
static float fmn(float a, float b) { return a < b ? a : b; }                            /* :175-177 */
static float comp(f3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
static uint32_t bits_of(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float float_of(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* rows of the 4x4 are (tangent.i, normal.i, bitangent.i, 0): TerraMath.inl:251-272.
 * The tangent is MULTIPLIED by the sqrt (not divided) — reproduced, see SURVEY Q19. */
static TerraFloat4x4 frame_from_normal(f3 n) {
    f3 t;
    if (fabs(n.x) > fabs(n.y)) { t = mk(n.z, 0.f, -n.x); t = scl(t, sqrtf(n.x * n.x + n.z * n.z)); }
    else                       { t = mk(0.f, -n.z, n.y); t = scl(t, sqrtf(n.y * n.y + n.z * n.z)); }
    f3 b = cross(n, t);
    TerraFloat4x4 m;
    m.rows[0] = (TerraFloat4){ t.x, n.x, b.x, 0.f };
    m.rows[1] = (TerraFloat4){ t.y, n.y, b.y, 0.f };
    m.rows[2] = (TerraFloat4){ t.z, n.z, b.z, 0.f };
    m.rows[3] = (TerraFloat4){ 0.f, 0.f, 0.f, 1.f };
    return m;
}
static f3 xform(const TerraFloat4x4* m, f3 v) {                                          /* TerraMath.inl:234-238 */
    return mk(dot(mk(m->rows[0].x, m->rows[0].y, m->rows[0].z), v),
              dot(mk(m->rows[1].x, m->rows[1].y, m->rows[1].z), v),
              dot(mk(m->rows[2].x, m->rows[2].y, m->rows[2].z), v));
}

/* ------------------------------------------------------------------ rays
 * terra_ray (Terra.c:1702-1710) + Wald-2013 per-ray setup (TerraGeometry.c:98-134). */
typedef struct { f3 o, d, inv; int kx, ky, kz; float sx, sy, sz; } ORay;

static ORay ray_make(f3 o, f3 d) {
    ORay r;
    r.o = o; r.d = d;
    r.inv = mk(1.f / d.x, 1.f / d.y, 1.f / d.z);
    f3 a = mk((float)fabs(d.x), (float)fabs(d.y), (float)fabs(d.z));
    int kz = a.x > a.y ? (a.x > a.z ? 0 : 2) : (a.y > a.z ? 1 : 2);                     /* TerraMath.inl:155-169 */
    int kx = kz + 1 == 3 ? 0 : kz + 1;
    int ky = kx + 1 == 3 ? 0 : kx + 1;
    if (comp(d, kz) < 0.f) { int t = kx; kx = ky; ky = t; }
    r.sz = 1.f / comp(d, kz);
    r.sx = comp(d, kx) * r.sz;
    r.sy = comp(d, ky) * r.sz;
    r.kx = kx; r.ky = ky; r.kz = kz;
    return r;
}
/* terra_scene_raycast moves the origin 0.001 along the direction and re-derives the state: Terra.c:1628-1632 */
static ORay ray_for_raycast(f3 o, f3 d) { return ray_make(add(o, scl(d, 0.001f)), d); }

/* terra_ray_aabb_intersection, Terra.c:851-878 */
static int hit_box(const ORay* r, const TerraAABB* b, float* tmin_out) {
    float t1 = (b->min.x - r->o.x) * r->inv.x, t2 = (b->max.x - r->o.x) * r->inv.x;
    float tmin = fmn(t1, t2), tmax = fmx(t1, t2);
    t1 = (b->min.y - r->o.y) * r->inv.y; t2 = (b->max.y - r->o.y) * r->inv.y;
    tmin = fmx(tmin, fmn(t1, t2)); tmax = fmn(tmax, fmx(t1, t2));
    t1 = (b->min.z - r->o.z) * r->inv.z; t2 = (b->max.z - r->o.z) * r->inv.z;
    tmin = fmx(tmin, fmn(t1, t2)); tmax = fmn(tmax, fmx(t1, t2));
    if (tmin_out) *tmin_out = tmin;
    return tmax > fmx(tmin, 0.f);
}

/* terra_ray_triangle_intersection_query, TerraGeometry.c:159-259 */
static int hit_triangle(const ORay* r, const TerraTriangle* tri, float* t, float* u, float* v) {
    f3 A = sub(tri->a, r->o), B = sub(tri->b, r->o), C = sub(tri->c, r->o);
    float Az_ = comp(A, r->kz), Bz_ = comp(B, r->kz), Cz_ = comp(C, r->kz);
    float Ax = comp(A, r->kx) - r->sx * Az_, Ay = comp(A, r->ky) - r->sy * Az_;
    float Bx = comp(B, r->kx) - r->sx * Bz_, By = comp(B, r->ky) - r->sy * Bz_;
    float Cx = comp(C, r->kx) - r->sx * Cz_, Cy = comp(C, r->ky) - r->sy * Cz_;
    float U = Cx * By - Cy * Bx, V = Ax * Cy - Ay * Cx, W = Bx * Ay - By * Ax;
    if (U == 0.f || V == 0.f || W == 0.f) {                                             /* :204-208 fp64 retry */
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    uint32_t sgn = bits_of(U) & 0x80000000u;
    if (sgn != (bits_of(V) & 0x80000000u) || sgn != (bits_of(W) & 0x80000000u)) return 0;
    float det = U + V + W;
    if (det == 0.f) return 0;
    float depth = U * (r->sz * Az_) + V * (r->sz * Bz_) + W * (r->sz * Cz_);
    if (float_of(bits_of(depth) ^ sgn) < 0.f) return 0;                                 /* :243 */
    float inv_det = 1.f / det;
    *u = U * inv_det; *v = V * inv_det; *t = depth * inv_det;
    return 1;
}

/* ------------------------------------------------------------------ BVH (TerraBVH.c)
 * 64-byte node: two child boxes, two indices, two types (-1 inner / 1 leaf), TerraBVH.h:13-17 */
typedef struct { TerraAABB box[2]; int32_t index[2]; int32_t type[2]; } ONode;
typedef struct { TerraAABB box; uint32_t index; int32_t type; } OVolume;                /* TerraBVH.c:11-15 */

static TerraAABB box_empty(void) {
    TerraAABB b = { { FLT_MAX, FLT_MAX, FLT_MAX }, { -FLT_MAX, -FLT_MAX, -FLT_MAX } };
    return b;
}
static void box_add_triangle(TerraAABB* b, const TerraTriangle* t) {                    /* Terra.c:972-997 */
    const f3* v[3] = { &t->a, &t->b, &t->c };
    for (int i = 0; i < 3; ++i) b->min.x = fmn(b->min.x, v[i]->x);
    for (int i = 0; i < 3; ++i) b->min.y = fmn(b->min.y, v[i]->y);
    for (int i = 0; i < 3; ++i) b->min.z = fmn(b->min.z, v[i]->z);
    b->min.x -= OEPS; b->min.y -= OEPS; b->min.z -= OEPS;
    for (int i = 0; i < 3; ++i) b->max.x = fmx(b->max.x, v[i]->x);
    for (int i = 0; i < 3; ++i) b->max.y = fmx(b->max.y, v[i]->y);
    for (int i = 0; i < 3; ++i) b->max.z = fmx(b->max.z, v[i]->z);
    b->max.x += OEPS; b->max.y += OEPS; b->max.z += OEPS;
}
static void box_add_box(TerraAABB* b, const TerraAABB* o) {                             /* TerraBVH.c:70-77 */
    b->min.x = fmn(b->min.x, o->min.x); b->min.y = fmn(b->min.y, o->min.y); b->min.z = fmn(b->min.z, o->min.z);
    b->max.x = fmx(b->max.x, o->max.x) + OEPS;
    b->max.y = fmx(b->max.y, o->max.y) + OEPS;
    b->max.z = fmx(b->max.z, o->max.z) + OEPS;
}
static float box_surface(const TerraAABB* b) {                                          /* TerraBVH.c:24-29 */
    float w = b->max.x - b->min.x, h = b->max.y - b->min.y, d = b->max.z - b->min.z;
    return 2 * (w * d + w * h + d * h);
}
static int by_centre_x(const void* l, const void* r) {                                  /* TerraBVH.c:37-46 (bool, not 3-way) */
    const TerraAABB* a = &((const OVolume*)l)->box;
    const TerraAABB* b = &((const OVolume*)r)->box;
    return (a->min.x + a->max.x) / 2 < (b->min.x + b->max.x) / 2;
}
static int split_sah(OVolume* v, int n, const TerraAABB* container, float* la, float* ra) {   /* TerraBVH.c:79-126 */
    float carea = container ? box_surface(container) : FLT_MAX;
    qsort(v, (size_t)n, sizeof *v, by_centre_x);
    TerraAABB acc = box_empty();
    for (int i = 0; i < n; ++i) { box_add_box(&acc, &v[i].box); la[i] = box_surface(&acc); }
    acc = box_empty();
    for (int i = n - 1; i >= 0; --i) { box_add_box(&acc, &v[i].box); ra[i] = box_surface(&acc); }
    float best = FLT_MAX; int best_i = -1;
    for (int i = 0; i < n; ++i) {
        int lc = i + 1, rc = n - lc;
        float cost = lc * la[i] / carea + rc * ra[i] / carea;
        if (cost < best) { best = cost; best_i = i; }
    }
    return best_i;
}
typedef struct { int lo, hi, node; const TerraAABB* box; } OTask;

static ONode* bvh_build(const TerraObject* objects, int n_objects, int* n_nodes) {      /* TerraBVH.c:128-244 */
    TerraAABB scene = box_empty();
    int total = 0;
    for (int j = 0; j < n_objects; ++j) total += (int)objects[j].triangles_count;
    OVolume* vol = malloc(sizeof *vol * (size_t)total);
    int p = 0;
    for (int j = 0; j < n_objects; ++j)
        for (int i = 0; i < (int)objects[j].triangles_count; ++i, ++p) {
            box_add_triangle(&scene, &objects[j].triangles[i]);
            vol[p].box = box_empty();
            box_add_triangle(&vol[p].box, &objects[j].triangles[i]);
            vol[p].type = 1;
            vol[p].index = (uint32_t)(j | (i << 8));
        }
    ONode* nodes = malloc(sizeof *nodes * (size_t)total * 2);
    OTask* stack = malloc(sizeof *stack * (size_t)total * 2);
    float* la = malloc(sizeof(float) * (size_t)total);
    float* ra = malloc(sizeof(float) * (size_t)total);
    int count = 1, sp = 0;
    stack[sp++] = (OTask){ 0, total, 0, &scene };
    while (sp > 0) {
        OTask t = stack[--sp];
        int split = split_sah(vol + t.lo, t.hi - t.lo, t.box, la, ra) + t.lo;
        ONode* nd = &nodes[t.node];
        if (split == t.lo) {
            nd->type[0] = vol[t.lo].type; nd->box[0] = vol[t.lo].box; nd->index[0] = (int32_t)vol[t.lo].index;
        } else {
            TerraAABB b = box_empty();
            for (int i = t.lo; i < split + 1; ++i) box_add_box(&b, &vol[i].box);
            nd->type[0] = -1; nd->box[0] = b; nd->index[0] = count;
            stack[sp++] = (OTask){ t.lo, split + 1, count, &nd->box[0] };
            ++count;
        }
        if (split == t.hi - 2) {
            nd->type[1] = vol[t.hi - 1].type; nd->box[1] = vol[t.hi - 1].box; nd->index[1] = (int32_t)vol[t.hi - 1].index;
        } else {
            TerraAABB b = box_empty();
            for (int i = split + 1; i < t.hi; ++i) box_add_box(&b, &vol[i].box);
            nd->type[1] = -1; nd->box[1] = b; nd->index[1] = count;
            stack[sp++] = (OTask){ split + 1, t.hi, count, &nd->box[1] };
            ++count;
        }
    }
    free(vol); free(stack); free(la); free(ra);
    *n_nodes = count;
    return nodes;
}

/* ------------------------------------------------------------------ scene */
typedef struct { f3 power; float area; int object; float* tri_area; } OLight;          /* TerraPrivate.h:24-29 */

struct OracleScene {
    const TerraObject* objects;
    int                n_objects;
    int                env_lit;      /* extension mode: Terra.c:1056 un-commented */
    TerraSceneOptions  opts;
    ONode*             nodes;
    int                n_nodes;
    OLight*            lights;
    int                n_lights;
    size_t             lights_triangles_count;
};

static f3 attr_eval(const TerraAttribute* a, const void* uv, const f3* xyz) {           /* Terra.c:1804-1810 */
    if (a->state != NULL) return a->eval(a->state, uv, xyz);
    return a->value;
}
static float tri_area(const TerraTriangle* t) {                                         /* Terra.c:1833-1838 */
    return len(cross(sub(t->b, t->a), sub(t->c, t->a))) / 2;
}

OracleScene* oracle_scene_create(const TerraObject* objects, int n_objects, const TerraSceneOptions* opts) {
    OracleScene* s = calloc(1, sizeof *s);
    s->objects = objects; s->n_objects = n_objects; s->opts = *opts;
    s->nodes = bvh_build(objects, n_objects, &s->n_nodes);
    s->lights = calloc((size_t)n_objects + 1, sizeof *s->lights);
    for (int i = 0; i < n_objects; ++i) {                                               /* Terra.c:194-231 */
        TerraFloat2 mid = { 0.5f, 0.5f };
        f3 e = attr_eval(&objects[i].material.emissive, &mid, NULL);
        if (e.x == 0 && e.y == 0 && e.z == 0) continue;
        OLight* l = &s->lights[s->n_lights++];
        l->tri_area = malloc(sizeof(float) * objects[i].triangles_count);
        float area = 0;
        for (size_t j = 0; j < objects[i].triangles_count; ++j) {
            float a = tri_area(&objects[i].triangles[j]);
            l->tri_area[j] = a;
            area += a;
        }
        l->power = scl(e, area * OPI);
        l->area = area;
        l->object = i;
        s->lights_triangles_count += objects[i].triangles_count;
    }
    return s;
}
void oracle_scene_destroy(OracleScene* s) {
    for (int i = 0; i < s->n_lights; ++i) free(s->lights[i].tri_area);
    free(s->lights); free(s->nodes); free(s);
}
const void* oracle_bvh_nodes(const OracleScene* s, int* n_nodes) { *n_nodes = s->n_nodes; return s->nodes; }

/* ------------------------------------------------------------------ traversal
 * Reference walk, TerraBVH.c:250-310: LIFO stack, both children of every popped node handled in
 * order 0,1; leaves are tested without a box test; strict < keeps the first-found triangle. */
typedef struct { int found; uint32_t obj, tri; float t, u, v; f3 point; } OClosest;

static OClosest closest_reference(const OracleScene* s, const ORay* r) {
    int stack[4096], sp = 0;          /* the reference has 64 entries and no overflow check */
    OClosest best = { 0, 0, 0, FLT_MAX, 0.f, 0.f, { FLT_MAX, FLT_MAX, FLT_MAX } };
    stack[sp++] = 0;
    while (sp > 0) {
        const ONode* nd = &s->nodes[stack[--sp]];
        for (int c = 0; c < 2; ++c) {
            if (nd->type[c] == -1) {
                if (hit_box(r, &nd->box[c], NULL)) stack[sp++] = nd->index[c];
            } else {
                int obj = nd->index[c] & 0xff, tri = nd->index[c] >> 8;
                float t, u, v;
                if (hit_triangle(r, &s->objects[obj].triangles[tri], &t, &u, &v) && t < best.t) {
                    best.found = 1; best.obj = (uint32_t)obj; best.tri = (uint32_t)tri;
                    best.t = t; best.u = u; best.v = v;
                    best.point = add(r->o, scl(r->d, t));                                /* terra_ray_pos, TerraGeometry.c:9-12 */
                }
            }
        }
    }
    return best;
}

static OracleHit as_hit(const OClosest* c) {
    OracleHit h = { -1.f, 0xFFFFFFFFu, 0.f, 0.f };
    if (c->found) { h.t = c->t; h.prim = c->obj | (c->tri << 8); h.u = c->u; h.v = c->v; }
    return h;
}

void oracle_trace_rays(const OracleScene* s, size_t n, const float* o, const float* d, OracleHit* out) {
    for (size_t i = 0; i < n; ++i) {
        ORay r = ray_for_raycast(mk(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
        OClosest c = closest_reference(s, &r);
        out[i] = as_hit(&c);
    }
}

/* Canonical counter: near-first ordered walk with t culling over any tree in the 64-byte format.
 * Leaf children are box-tested first (their box is already in the fetched node). Exact t ties keep
 * the smaller object|tri<<8 id. This is the minimum-work closest-hit walk whose visit counts define
 * the algorithmic bytes per ray: 64 B per node fetched, 36 B per triangle tested. */
void oracle_trace_counted(const OracleScene* s, const void* nodes_v, size_t n, const float* o, const float* d,
                          OracleHit* out, uint64_t counters[2]) {
    const ONode* nodes = nodes_v ? (const ONode*)nodes_v : s->nodes;
    for (size_t i = 0; i < n; ++i) {
        ORay r = ray_for_raycast(mk(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
        struct { int node; float tmin; } stack[512];
        int sp = 0, node = 0;
        OracleHit best = { -1.f, 0xFFFFFFFFu, 0.f, 0.f };
        float best_t = FLT_MAX, cull = FLT_MAX;
        while (node >= 0) {
            const ONode* nd = &nodes[node];
            counters[0]++;
            float tm[2]; int hit[2];
            for (int c = 0; c < 2; ++c) hit[c] = hit_box(&r, &nd->box[c], &tm[c]) && !(tm[c] > cull);
            for (int c = 0; c < 2; ++c) {
                if (!hit[c] || nd->type[c] == -1 || tm[c] > cull) continue;
                int obj = nd->index[c] & 0xff, tri = nd->index[c] >> 8;
                float t, u, v;
                counters[1]++;
                if (hit_triangle(&r, &s->objects[obj].triangles[tri], &t, &u, &v)) {
                    uint32_t prim = (uint32_t)nd->index[c];
                    if (t < best_t || (t == best_t && best.prim != 0xFFFFFFFFu && prim < best.prim)) {
                        best_t = t; best.t = t; best.prim = prim; best.u = u; best.v = v;
                        cull = best_t + best_t * 1e-5f;
                    }
                }
            }
            int in0 = hit[0] && nd->type[0] == -1 && !(tm[0] > cull);
            int in1 = hit[1] && nd->type[1] == -1 && !(tm[1] > cull);
            if (in0 && in1) {
                int near1 = tm[1] < tm[0];
                stack[sp].node = near1 ? nd->index[0] : nd->index[1];
                stack[sp].tmin = near1 ? tm[0] : tm[1];
                ++sp;
                node = near1 ? nd->index[1] : nd->index[0];
            } else if (in0) node = nd->index[0];
            else if (in1) node = nd->index[1];
            else {
                node = -1;
                while (sp > 0) { --sp; if (!(stack[sp].tmin > cull)) { node = stack[sp].node; break; } }
            }
        }
        if (out) out[i] = best;
    }
}

/* ------------------------------------------------------------------ camera (Terra.c:1770-1799) */
static TerraFloat4x4 camera_frame(const TerraCamera* c) {
    f3 z = unit(c->direction), x = unit(cross(c->up, z)), y = cross(z, x);
    TerraFloat4x4 m;
    m.rows[0] = (TerraFloat4){ x.x, y.x, z.x, 0.f };
    m.rows[1] = (TerraFloat4){ x.y, y.y, z.y, 0.f };
    m.rows[2] = (TerraFloat4){ x.z, y.z, z.z, 0.f };
    m.rows[3] = (TerraFloat4){ 0.f, 0.f, 0.f, 1.f };
    return m;
}
static f3 camera_sample(const TerraCamera* c, size_t fbw, size_t fbh, size_t px, size_t py, float jitter, float r1, float r2) {
    float dx = -jitter + 2 * r1 * jitter, dy = -jitter + 2 * r2 * jitter;
    float ndc_x = (px + 0.5f + dx) / fbw, ndc_y = (py + 0.5f + dy) / fbh;
    float sx = 2 * ndc_x - 1, sy = 1 - 2 * ndc_y;
    float aspect = (float)fbw / (float)fbh;
    float fx = sx * aspect * (float)tan((c->fov * 0.0174533f) / 2);
    float fy = sy * (float)tan((c->fov * 0.0174533f) / 2);
    return unit(mk(fx, fy, 1.f));
}

void oracle_primary_hits(const OracleScene* s, const TerraCamera* cam, size_t w, size_t h, OracleHit* out, float* dirs_out) {
    TerraFloat4x4 rot = camera_frame(cam);
    for (size_t i = 0; i < h; ++i)
        for (size_t j = 0; j < w; ++j) {
            f3 d = xform(&rot, camera_sample(cam, w, h, j, i, 0.f, 0.f, 0.f));
            if (dirs_out) memcpy(dirs_out + 3 * (i * w + j), &d, 12);
            ORay r = ray_for_raycast(cam->position, d);
            OClosest c = closest_reference(s, &r);
            out[i * w + j] = as_hit(&c);
        }
}

/* ------------------------------------------------------------------ surface (Terra.c:1726-1764) */
static void surface_init(TerraShadingSurface* sf, const TerraTriangle* tri, const TerraMaterial* m,
                         const TerraTriangleProperties* pr, const f3* point) {
    f3 e0 = sub(tri->b, tri->a), e1 = sub(tri->c, tri->a), p = sub(*point, tri->a);
    float d00 = dot(e0, e0), d11 = dot(e1, e1), d01 = dot(e0, e1), dp0 = dot(p, e0), dp1 = dot(p, e1);
    float div = d00 * d11 - d01 * d01;
    float ux = (d11 * dp0 - d01 * dp1) / div, uy = (d00 * dp1 - d01 * dp0) / div;
    sf->normal = unit(add(add(scl(pr->normal_c, uy), scl(pr->normal_b, ux)), scl(pr->normal_a, 1 - ux - uy)));
    TerraFloat2 tc;
    tc.x = (pr->texcoord_c.x * uy + pr->texcoord_b.x * ux) + pr->texcoord_a.x * (1 - ux - uy);
    tc.y = (pr->texcoord_c.y * uy + pr->texcoord_b.y * ux) + pr->texcoord_a.y * (1 - ux - uy);
    for (int i = 0; i < (int)m->attributes_count; ++i) sf->attributes[i] = attr_eval(&m->attributes[i], &tc, point);
    sf->emissive = attr_eval(&m->emissive, &tc, point);
    sf->transform = frame_from_normal(sf->normal);
}

/* terra_scene_raycast, Terra.c:1623-1657. Returns object index or -1. */
static int raycast(const OracleScene* s, f3 o, f3 d, TerraShadingSurface* sf, f3* point, size_t* tri) {
    ORay r = ray_for_raycast(o, d);
    OClosest c = closest_reference(s, &r);
    *point = c.point;
    if (!c.found) return -1;
    if (tri) *tri = c.tri;
    const TerraObject* ob = &s->objects[c.obj];
    surface_init(sf, &ob->triangles[c.tri], &ob->material, &ob->properties[c.tri], point);
    return (int)c.obj;
}

int oracle_raycast_surface(const OracleScene* s, const float* origin, const float* dir,
                           TerraShadingSurface* surf, float* point, int* tri) {
    f3 p; size_t t = 0;
    int obj = raycast(s, mk(origin[0], origin[1], origin[2]), mk(dir[0], dir[1], dir[2]), surf, &p, &t);
    if (obj < 0) return -1;
    memcpy(point, &p, 12);
    *tri = (int)t;
    return obj;
}

/* KAT helper: the surface record a BSDF callback sees for a given shading normal and attribute
 * values (frame = terra_f4x4_basis(normal), TerraMath.inl:251-272). */
void oracle_surface_from_normal(const float* normal, const float* attributes24, TerraShadingSurface* sf) {
    memset(sf, 0, sizeof *sf);
    sf->normal = mk(normal[0], normal[1], normal[2]);
    sf->transform = frame_from_normal(sf->normal);
    for (int i = 0; i < TERRA_MATERIAL_MAX_ATTRIBUTES; ++i)
        sf->attributes[i] = mk(attributes24[3 * i], attributes24[3 * i + 1], attributes24[3 * i + 2]);
}

/* ------------------------------------------------------------------ presets
 * Diffuse (TerraPresets.c:34-60) and Phong (:66-146): live reference code restated.
 * Slots: diffuse [0]=albedo; phong [0]=specular colour [1]=albedo [2].x=exponent [3].x=lobe pick. */
static f3 o_diffuse_sample(const TerraShadingSurface* s, float e1, float e2, float e3, const f3* wo) {
    (void)e3; (void)wo;
    float r = sqrtf(e1), theta = 2 * OPI * e2;
    f3 wi = mk(r * cosf(theta), sqrtf(fmx(0.f, 1 - e1)), r * sinf(theta));
    return unit(xform(&s->transform, wi));
}
static float o_diffuse_pdf(const TerraShadingSurface* s, const f3* wi, const f3* wo) {
    (void)wo;
    return fmx(0.f, dot(s->normal, *wi)) / OPI;
}
static f3 o_diffuse_eval(const TerraShadingSurface* s, const f3* wi, const f3* wo) {
    (void)wi; (void)wo;
    return scl(s->attributes[0], 1. / OPI);
}
static void phong_kd_ks(const TerraShadingSurface* s, float* kd, float* ks) {           /* :66-82 */
    float diffuse = fmx(s->attributes[1].x + s->attributes[1].y + s->attributes[1].z, OEPS);
    float specular = s->attributes[0].x + s->attributes[0].y + s->attributes[0].z;
    if (specular > diffuse) { *kd = 0.5f * diffuse / specular; *ks = 1.f - *kd; }
    else                    { *ks = 0.5f * specular / diffuse; *kd = 1.f - *ks; }
}
static f3 mirror_dir(const TerraShadingSurface* s, const f3* wo) {                      /* :96-97 */
    return sub(scl(s->normal, 2.f * dot(*wo, s->normal)), *wo);
}
static f3 o_phong_sample(const TerraShadingSurface* s, float e1, float e2, float e3, const f3* wo) {
    float kd, ks;
    phong_kd_ks(s, &kd, &ks);
    float* pick = &((TerraShadingSurface*)s)->attributes[3].x;     /* the reference writes through const too, :87 */
    if (e3 < kd) { *pick = 1.f; return o_diffuse_sample(s, e1, e2, e3, wo); }
    *pick = -1.f;
    TerraFloat4x4 fr = frame_from_normal(mirror_dir(s, wo));
    float phi = 2 * OPI * e1;
    float theta = acosf(powf(1.f - e2, 1.f / (s->attributes[2].x + 1)));
    float st = sinf(theta);
    return unit(xform(&fr, mk(st * cosf(phi), cosf(theta), st * sinf(phi))));
}
static float o_phong_pdf(const TerraShadingSurface* s, const f3* wi, const f3* wo) {    /* :108-123 */
    float pick = s->attributes[3].x;
    if (pick == 1.f) return o_diffuse_pdf(s, wi, wo);
    if (pick == -1.f) {
        float ca = dot(*wi, mirror_dir(s, wo)), n = s->attributes[2].x;
        return (n + 1) / (2 * OPI) * powf(ca, n);
    }
    return 0.f;
}
static f3 o_phong_eval(const TerraShadingSurface* s, const f3* wi, const f3* wo) {      /* :125-140 */
    float kd, ks;
    phong_kd_ks(s, &kd, &ks);
    float n = s->attributes[2].x;
    f3 dterm = scl(s->attributes[1], kd * 1.f / OPI);
    float ca = dot(*wi, mirror_dir(s, wo));
    float cn = powf(ca, n);
    f3 sterm = scl(s->attributes[0], ks * cn * (n + 2) / (2 * OPI));
    return add(dterm, sterm);
}

/* GGX rough dielectric — dead code in the reference (TerraPresets.c:298-394), old
 * {sample, weight, shade} API. Restated for {sample, pdf, eval}: pdf := weight, eval := shade
 * without its trailing *NoL; the cached half vector is recomputed as normalize(wi + wo).
 * Slots: [0]=albedo [1]=(roughness, metalness, ior). Pinned against the dead code compiled through
 * ref_dead_presets.c: directions bit-identical, weight / shade within 1e-4 (the recomputed half vector). */
static float ggx_chi(float v) { return v <= 0.f ? 0.f : 1.f; }
static float ggx_G1(f3 v, f3 n, f3 h, float a2) {                                       /* :307-314 */
    float VoH = dot(v, h), VoN = dot(v, n);
    float chi = ggx_chi(VoH / VoN), VoH2 = VoH * VoH;
    if (VoH2 == 0.f || VoN == 0.f) return 0.f;        /* restatement guard against 0/0, 0*inf */
    float tan2 = (1.f - VoH2) / VoH2;
    return (chi * 2) / (sqrtf(1 + a2 * tan2) + 1);
}
static float ggx_D(float NoH, float a2) {                                               /* :316-320 */
    float NoH2 = NoH * NoH, den = NoH2 * a2 + (1 - NoH2);
    return (ggx_chi(NoH) * a2) / (OPI * den * den);
}
static int ggx_half(const f3* wi, const f3* wo, f3* h) {      /* restatement guard: wi == -wo has no half vector */
    f3 hs = add(*wi, *wo);
    if (!(sqlen(hs) > 0.f)) return 0;
    *h = unit(hs);
    return 1;
}
static f3 o_ggx_sample(const TerraShadingSurface* s, float e1, float e2, float e3, const f3* wo) {   /* :322-345 */
    const float pd = 1.f - s->attributes[1].y;
    if (e3 <= pd) return o_diffuse_sample(s, e1, e2, e3, wo);
    float alpha = s->attributes[1].x;
    float theta = atanf((alpha * sqrtf(e1)) / sqrtf(1 - e1));
    float phi = 2 * OPI * e2, st = sinf(theta);
    f3 h = unit(xform(&s->transform, mk(st * cosf(phi), cosf(theta), st * sinf(phi))));
    float HoV = fmx(0.f, dot(h, *wo));
    return sub(scl(h, 2 * HoV), *wo);
}
static float o_ggx_pdf(const TerraShadingSurface* s, const f3* wi, const f3* wo) {      /* :347-356 */
    f3 h;
    if (!ggx_half(wi, wo, &h)) return 0.f;
    float alpha = s->attributes[1].x, a2 = alpha * alpha;
    float NoH = dot(s->normal, h);
    float ws = ggx_D(NoH, a2) * NoH, wd = o_diffuse_pdf(s, wi, wo);
    const float pd = 1.f - s->attributes[1].y, ps = 1.f - pd;
    return wd * pd + ws * ps;
}
static float lerp1(float a, float b, float t) { return a + (b - a) * t; }
static f3 o_ggx_eval(const TerraShadingSurface* s, const f3* wi, const f3* wo) {        /* :358-388 */
    f3 h;
    if (!ggx_half(wi, wo, &h)) return mk(0.f, 0.f, 0.f);
    float metal = s->attributes[1].y, ior = s->attributes[1].z;
    f3 alb = s->attributes[0];
    float f0 = (1.f - ior) / (1.f + ior);                                               /* :21-27 */
    f0 *= f0;
    f3 F0 = mk(lerp1(f0, alb.x, metal), lerp1(f0, alb.y, metal), lerp1(f0, alb.z, metal));
    float ct = fmx(0.f, dot(*wo, h));                                                   /* :14-19 */
    float p5 = powf(1 - ct, 5);
    f3 ks = add(scl(mk(1 - F0.x, 1 - F0.y, 1 - F0.z), p5), F0);
    float NoL = fmx(dot(s->normal, *wi), 0.f), NoV = fmx(dot(s->normal, *wo), 0.f), NoH = fmx(dot(s->normal, h), 0.f);
    float alpha = s->attributes[1].x, a2 = alpha * alpha;
    float D = ggx_D(NoH, a2);
    float G = ggx_G1(*wo, s->normal, h, a2) * ggx_G1(*wi, s->normal, h, a2);
    float den = fmn((4 * NoL * NoV) + 0.05f, 1.f);                                      /* :375 */
    f3 spec = scl(ks, G * D / den);
    f3 diff = scl(alb, 1. / OPI);
    const float pd = 1.f - metal, ps = 1.f - pd;
    f3 dfac = scl(mk(1.f - ks.x, 1.f - ks.y, 1.f - ks.z), (1.f - metal) * pd);
    return add(had(diff, dfac), scl(spec, ps));
}

/* Perfect glass — dead code in the reference (TerraPresets.c:396-464). sample() leaves the
 * fresnel weight in attributes[2].x (the way Phong leaves its lobe pick); pdf returns it; eval =
 * albedo * weight / (n.wi), so that terra_trace's f/pdf*(n.wi) yields the old shade/weight = albedo
 * for both the reflected and the transmitted (n.wi < 0) direction. Slots: [0]=albedo [1].x=ior
 * [2].x=state (reset to its constant 0 at every hit by terra_surface_init). Pinned against the dead code
 * compiled through ref_dead_presets.c: directions and weights bit-identical, from outside and from inside. */
static f3 o_glass_sample(const TerraShadingSurface* s, float e1, float e2, float e3, const f3* wo) {
    (void)e1; (void)e2;
    float* state = &((TerraShadingSurface*)s)->attributes[2].x;
    f3 n = s->normal, in = neg(*wo);
    float n1, n2, cos_i = dot(n, in), ior = s->attributes[1].x;
    if (cos_i > 0.f) { n1 = ior; n2 = 1.f; n = neg(n); }
    else             { n1 = 1.f; n2 = ior; cos_i = -cos_i; }
    f3 refl = sub(in, scl(n, 2 * dot(n, in)));
    float nni = n1 / n2;
    float cos_t = 1.f - nni * nni * (1.f - cos_i * cos_i);
    if (cos_t < 0.f) { *state = 1.f; return refl; }                                     /* total internal reflection */
    cos_t = sqrtf(cos_t);
    float t = 1.f - (n1 <= n2 ? cos_i : cos_t);
    float R0 = (n1 - n2) / (n1 + n2);
    R0 *= R0;
    float R = R0 + (1 - R0) * (t * t * t * t * t);
    if (e3 < R) { *state = R; return refl; }
    f3 tr = unit(add(scl(n, (n1 / n2) * cos_i - cos_t), scl(in, n1 / n2)));
    *state = 1 - R;
    return tr;
}
static float o_glass_pdf(const TerraShadingSurface* s, const f3* wi, const f3* wo) { (void)wi; (void)wo; return s->attributes[2].x; }
static f3 o_glass_eval(const TerraShadingSurface* s, const f3* wi, const f3* wo) {
    (void)wo;
    float NoL = dot(s->normal, *wi);
    if (NoL == 0.f) return mk(0.f, 0.f, 0.f);
    return scl(s->attributes[0], s->attributes[2].x / NoL);
}

void oracle_set_env_lit(OracleScene* s, int on) { s->env_lit = on; }
void terra_oracle_bsdf_diffuse_init(TerraBSDF* b) { b->sample = o_diffuse_sample; b->pdf = o_diffuse_pdf; b->eval = o_diffuse_eval; }
void terra_oracle_bsdf_phong_init(TerraBSDF* b)   { b->sample = o_phong_sample;   b->pdf = o_phong_pdf;   b->eval = o_phong_eval; }
void terra_oracle_bsdf_ggx_init(TerraBSDF* b)     { b->sample = o_ggx_sample;     b->pdf = o_ggx_pdf;     b->eval = o_ggx_eval; }
void terra_oracle_bsdf_glass_init(TerraBSDF* b)   { b->sample = o_glass_sample;   b->pdf = o_glass_pdf;   b->eval = o_glass_eval; }

/* ------------------------------------------------------------------ lights + integrators */
static float randf(void) { return (float)rand() / RAND_MAX; }                            /* Terra.c:115 */

static f3 surface_ray_origin(const TerraShadingSurface* s, const f3* p) {               /* Terra.c:1715-1724, sign +1 */
    return add(*p, scl(s->normal, 0.0001f * 1.f));
}
static const OLight* pick_light(const OracleScene* s, float e, float* pdf) {            /* Terra.c:1592-1621 */
    size_t i = (size_t)(e * (double)s->n_lights);
    *pdf = 1.f / s->lights_triangles_count;
    return &s->lights[i];
}
static void sample_light_triangle(const OracleScene* s, const OLight* l, size_t ti, float e1, float e2, f3* pos, f3* nrm) {
    const TerraObject* ob = &s->objects[l->object];                                      /* Terra.c:1668-1697 */
    const TerraTriangle* tri = &ob->triangles[ti];
    const TerraTriangleProperties* pr = &ob->properties[ti];
    float sq = sqrtf(e1), a = 1 - sq, b = e2 * sq, c = 1 - a - b;
    *pos = add(add(scl(tri->a, a), scl(tri->b, b)), scl(tri->c, c));
    *nrm = unit(add(add(scl(pr->normal_a, a), scl(pr->normal_b, b)), scl(pr->normal_c, c)));
}

static f3 integrate_direct(const OracleScene* s, const TerraObject* ob, const TerraShadingSurface* sf,
                           const f3* p, const f3* wo, const f3* thr, size_t bounce) {    /* Terra.c:1349-1426 */
    f3 Lo = mk(0, 0, 0);
    if (bounce == 0 && dot(*wo, sf->normal) > 0) Lo = add(Lo, sf->emissive);
    float pick_pdf;
    float e = randf() - OEPS;
    const OLight* l = pick_light(s, e, &pick_pdf);
    size_t ti = (size_t)(randf() * s->objects[l->object].triangles_count);               /* :1662-1666 */
    f3 lp, ln;
    { float e1 = randf(); float e2 = randf(); sample_light_triangle(s, l, ti, e1, e2, &lp, &ln); }
    f3 to_light = sub(lp, *p), wi = unit(to_light);
    TerraShadingSurface lsf;
    f3 hp; size_t htri = 0;
    int hobj = raycast(s, surface_ray_origin(sf, p), wi, &lsf, &hp, &htri);
    if (hobj == l->object && htri == ti) {
        float cs = dot(neg(wi), ln);
        if (cs > 0) {
            f3 f = ob->material.bsdf.eval(sf, &wi, wo);
            float pdf = sqlen(to_light) / fabsf(cs * l->tri_area[htri]);
            f3 Ld = scl(had(lsf.emissive, f), dot(wi, sf->normal) / (pdf * pick_pdf));
            Lo = add(Lo, Ld);
        }
    }
    return had(Lo, *thr);
}

static f3 integrate_direct_mis(const OracleScene* s, const TerraObject* ob, const TerraShadingSurface* sf,
                               const f3* p, const f3* wo, const f3* thr, size_t bounce) { /* Terra.c:1428-1587 */
    f3 Lo = mk(0, 0, 0);
    if (bounce == 0 && dot(*wo, sf->normal) > 0) Lo = add(Lo, sf->emissive);
    f3 bs;
    { float e1 = randf(); float e2 = randf(); float e3 = randf(); bs = ob->material.bsdf.sample(sf, e1, e2, e3, wo); }
    float pick_pdf;
    float e = randf() - OEPS;
    const OLight* l = pick_light(s, e, &pick_pdf);
    size_t ti = (size_t)(randf() * s->objects[l->object].triangles_count);
    f3 lp, ln;
    { float e1 = randf(); float e2 = randf(); sample_light_triangle(s, l, ti, e1, e2, &lp, &ln); }
    {
        f3 to_light = sub(lp, *p), wi = unit(to_light);
        TerraShadingSurface lsf;
        f3 hp; size_t htri = 0;
        int hobj = raycast(s, surface_ray_origin(sf, p), wi, &lsf, &hp, &htri);
        if (hobj == l->object && htri == ti) {
            float cs = dot(ln, neg(wi));
            if (cs > 0) {
                float bpdf = ob->material.bsdf.pdf(sf, &wi, wo);
                float lpdf = sqlen(to_light) / fabsf(cs * l->tri_area[htri]);
                float wgt = (lpdf * lpdf) / (lpdf * lpdf + bpdf * bpdf);
                if (lpdf != 0) {
                    f3 f = ob->material.bsdf.eval(sf, &wi, wo);
                    Lo = add(Lo, scl(had(lsf.emissive, f), dot(wi, sf->normal) * wgt / (lpdf * pick_pdf)));
                }
            }
        }
    }
    {
        f3 wi = bs;
        f3 f = ob->material.bsdf.eval(sf, &wi, wo);
        float bpdf = ob->material.bsdf.pdf(sf, &wi, wo);
        TerraShadingSurface lsf;
        f3 hp; size_t htri = 0;
        int hobj = raycast(s, surface_ray_origin(sf, p), wi, &lsf, &hp, &htri);
        if (hobj == l->object) {
            float NoW = dot(lsf.normal, neg(wi));
            if (NoW > 0) {
                f3 dl = sub(*p, hp);
                float dist = dot(dl, dl);                                                /* terra_sqdistf3(&hp, p) */
                float lpdf = dist / (NoW * tri_area(&s->objects[hobj].triangles[htri]));
                float wgt = (bpdf * bpdf) / (lpdf * lpdf + bpdf * bpdf);
                if (bpdf != 0) Lo = add(Lo, scl(had(lsf.emissive, f), dot(wi, sf->normal) * wgt / bpdf));
            }
        }
    }
    return had(Lo, *thr);
}

/* Debug views, Terra.c:1140-1337. All return zero past bounce 0. */
static f3 integrate_debug_depth(const f3* ray_origin, const f3* p) {                    /* :1148-1157 */
    const float far_plane = 500.f;
    f3 ba = sub(*ray_origin, *p);
    float d = sqrtf(dot(ba, ba)) / far_plane;
    return mk(d, d, d);
}
static f3 integrate_debug_normals(const TerraShadingSurface* sf) {                       /* :1159-1198, terra_clampf3 TerraMath.inl:290-299 */
    f3 nn = sf->normal, p, n, c = mk(0, 0, 0);
    p.x = nn.x > 0.f ? nn.x : 0.f; p.y = nn.y > 0.f ? nn.y : 0.f; p.z = nn.z > 0.f ? nn.z : 0.f;
    p.x = p.x < 1.f ? p.x : 1.f;   p.y = p.y < 1.f ? p.y : 1.f;   p.z = p.z < 1.f ? p.z : 1.f;
    n.x = nn.x > -1.f ? nn.x : -1.f; n.y = nn.y > -1.f ? nn.y : -1.f; n.z = nn.z > -1.f ? nn.z : -1.f;
    n.x = n.x < 0.f ? n.x : 0.f;   n.y = n.y < 0.f ? n.y : 0.f;   n.z = n.z < 0.f ? n.z : 0.f;
    n = scl(n, -1.f);
    c = add(c, scl(mk(1, 0, 0), p.x)); c = add(c, scl(mk(0, 1, 0), p.y)); c = add(c, scl(mk(0, 0, 1), p.z));
    c = add(c, scl(mk(0, 1, 1), n.x)); c = add(c, scl(mk(1, 0, 1), n.y)); c = add(c, scl(mk(1, 1, 0), n.z));
    return c;
}
static f3 integrate_debug_mis_weight(const OracleScene* s, const TerraObject* ob, const TerraShadingSurface* sf,
                                     const f3* p, const f3* wo, const f3* thr) {          /* :1200-1337 */
    f3 Lo = mk(0, 0, 0);
    f3 bs;
    { float e1 = randf(); float e2 = randf(); float e3 = randf(); bs = ob->material.bsdf.sample(sf, e1, e2, e3, wo); }
    float pick_pdf;
    float e = randf() - OEPS;
    const OLight* l = pick_light(s, e, &pick_pdf);
    size_t ti = (size_t)(randf() * s->objects[l->object].triangles_count);
    f3 lp, ln;
    { float e1 = randf(); float e2 = randf(); sample_light_triangle(s, l, ti, e1, e2, &lp, &ln); }
    {
        f3 to_light = sub(lp, *p), wi = unit(to_light);
        TerraShadingSurface lsf;
        f3 hp; size_t htri = 0;
        int hobj = raycast(s, surface_ray_origin(sf, p), wi, &lsf, &hp, &htri);
        if (hobj == l->object && htri == ti) {
            float cs = dot(ln, neg(wi));
            if (cs > 0) {
                float bpdf = ob->material.bsdf.pdf(sf, &wi, wo);
                float lpdf = sqlen(to_light) / fabsf(cs * l->tri_area[htri]);
                Lo = add(Lo, mk(0, 0, (bpdf * bpdf) / (lpdf * lpdf + bpdf * bpdf)));
            }
        }
    }
    {
        f3 wi = bs;
        (void)ob->material.bsdf.eval(sf, &wi, wo);
        float bpdf = ob->material.bsdf.pdf(sf, &wi, wo);
        TerraShadingSurface lsf;
        f3 hp; size_t htri = 0;
        int hobj = raycast(s, surface_ray_origin(sf, p), wi, &lsf, &hp, &htri);
        if (hobj == l->object) {
            float NoW = dot(lsf.normal, neg(wi));
            if (NoW > 0) {
                f3 dl = sub(*p, hp);
                float dist = dot(dl, dl);
                float lpdf = dist / (NoW * tri_area(&s->objects[hobj].triangles[htri]));
                Lo = add(Lo, mk((bpdf * bpdf) / (lpdf * lpdf + bpdf * bpdf), 0, 0));
            }
        }
    }
    return had(Lo, *thr);
}

static f3 integrate(const OracleScene* s, const f3* ray_origin, const TerraObject* ob, const TerraShadingSurface* sf,
                    const f3* p, const f3* wo, const f3* thr, size_t bounce) {           /* Terra.c:1099-1138 */
    switch (s->opts.integrator) {
        case kTerraIntegratorSimple:                                                     /* :1339-1347 */
            return dot(*wo, sf->normal) > 0 ? had(sf->emissive, *thr) : mk(0, 0, 0);
        case kTerraIntegratorDirect:    return integrate_direct(s, ob, sf, p, wo, thr, bounce);
        case kTerraIntegratorDirectMis: return integrate_direct_mis(s, ob, sf, p, wo, thr, bounce);
        case kTerraIntegratorDebugMono:       return bounce != 0 ? mk(0, 0, 0) : mk(1, 1, 1);   /* :1140-1146 */
        case kTerraIntegratorDebugDepth:      return bounce != 0 ? mk(0, 0, 0) : integrate_debug_depth(ray_origin, p);
        case kTerraIntegratorDebugNormals:    return bounce != 0 ? mk(0, 0, 0) : integrate_debug_normals(sf);
        case kTerraIntegratorDebugMisWeights: return bounce != 0 ? mk(0, 0, 0) : integrate_debug_mis_weight(s, ob, sf, p, wo, thr);
        default:                        return mk(0, 0, 0);
    }
}

/* terra_trace, Terra.c:1039-1097 */
static f3 trace_path(const OracleScene* s, f3 o, f3 d) {
    f3 Lo = mk(0, 0, 0), thr = mk(1, 1, 1);
    for (size_t bounce = 0; bounce <= s->opts.bounces; ++bounce) {
        TerraShadingSurface sf;
        f3 p;
        int obj = raycast(s, o, d, &sf, &p, NULL);
        if (obj < 0) {            /* the environment lookup's result is discarded by the reference (:1054-1057) */
            if (s->env_lit) {     /* "patched oracle": the commented-out `Lo = Lo + throughput` of :1056 restored */
                f3 env = attr_eval(&s->opts.environment_map, &d, &p);
                thr = had(thr, env);
                Lo = add(Lo, thr);
            }
            break;
        }
        const TerraObject* ob = &s->objects[obj];
        f3 wo = neg(d);
        Lo = add(Lo, integrate(s, &o, ob, &sf, &p, &wo, &thr, bounce));
        float e0 = randf(); float e1 = randf(); float e2 = randf();
        f3 wi = ob->material.bsdf.sample(&sf, e0, e1, e2, &wo);
        float pdf = fmx(ob->material.bsdf.pdf(&sf, &wi, &wo), OEPS);
        f3 f = scl(ob->material.bsdf.eval(&sf, &wi, &wo), 1.f / pdf);
        thr = had(thr, f);
        thr = scl(thr, dot(sf.normal, wi));
        float pr = fmx(thr.x, fmx(thr.y, thr.z));
        float e3 = (float)rand() / RAND_MAX;
        if (e3 > pr) break;
        thr = scl(thr, 1.f / (pr + OEPS));
        o = surface_ray_origin(&sf, &p);
        d = wi;
    }
    return Lo;
}

/* ------------------------------------------------------------------ render (Terra.c:512-635) */
typedef struct { uint64_t state, inc; } OPcg;                                            /* Terra.c:678-701 */
static float pcg_next(OPcg* g) {
    uint64_t old = g->state;
    g->state = old * 6364136223846793005ULL + g->inc;
    uint32_t xs = (uint32_t)(((old >> 18) ^ old) >> 27), rot = (uint32_t)(old >> 59);
    uint32_t r = (xs >> rot) | (xs << ((-rot) & 31));
    return r * (1.f / ((uint64_t)1 << 32));
}
static f3 uncharted2(f3 x) {                                                             /* :1815-1828 */
    const float A = 0.15f, B = 0.5f, C = 0.1f, D = 0.2f, E = 0.02f, F = 0.3f;
    return mk(((x.x * (A * x.x + C * B) + D * E) / (x.x * (A * x.x + B) + D * F)) - E / F,
              ((x.y * (A * x.y + C * B) + D * E) / (x.y * (A * x.y + B) + D * F)) - E / F,
              ((x.z * (A * x.z + C * B) + D * E) / (x.z * (A * x.z + B) + D * F)) - E / F);
}
static f3 pow3(f3 c, float e) { return mk(powf(c.x, e), powf(c.y, e), powf(c.z, e)); }

void oracle_render(const OracleScene* s, const TerraCamera* cam, const TerraFramebuffer* fb,
                   size_t x, size_t y, size_t w, size_t h, uint32_t pcg_seed) {
    TerraFloat4x4 rot = camera_frame(cam);
    size_t spp = s->opts.samples_per_pixel;
    if (s->opts.sampling_method == kTerraSamplingMethodStratified) {                     /* :519-527 */
        size_t cur = s->opts.strata * s->opts.strata;
        while (spp > cur) cur *= cur;
        spp = cur;
    }
    OPcg g = { 0, 1 };
    pcg_next(&g); g.state += pcg_seed; pcg_next(&g);
    for (size_t i = y; i < y + h; ++i)
        for (size_t j = x; j < x + w; ++j) {
            f3 acc = mk(0, 0, 0);
            for (size_t k = 0; k < spp; ++k) {
                float r1 = pcg_next(&g), r2 = pcg_next(&g);
                f3 d = xform(&rot, camera_sample(cam, fb->width, fb->height, j, i, s->opts.subpixel_jitter, r1, r2));
                acc = add(acc, trace_path(s, cam->position, d));
            }
            TerraRawIntegrationResult* part = &fb->results[i * fb->width + j];
            part->acc = add(acc, part->acc);
            part->samples += (int)spp;
            f3 c = scl(mk(part->acc.x / (float)part->samples, part->acc.y / (float)part->samples, part->acc.z / (float)part->samples),
                       s->opts.manual_exposure);
            switch (s->opts.tonemapping_operator) {                                      /* :578-627 */
                case kTerraTonemappingOperatorLinear: c = pow3(c, 1.f / s->opts.gamma); break;
                case kTerraTonemappingOperatorReinhard:
                    c = mk(c.x / (1.f + c.x), c.y / (1.f + c.y), c.z / (1.f + c.z));
                    c = pow3(c, 1.f / s->opts.gamma);
                    break;
                case kTerraTonemappingOperatorFilmic: {
                    f3 q = mk(fmx(0.f, c.x - 0.004f), fmx(0.f, c.y - 0.004f), fmx(0.f, c.z - 0.004f));
                    c.x = (q.x * (6.2f * q.x + 0.5f)) / (q.x * (6.2f * q.x + 1.7f) + 0.06f);
                    c.y = (q.y * (6.2f * q.y + 0.5f)) / (q.y * (6.2f * q.y + 1.7f) + 0.06f);
                    c.x = (q.z * (6.2f * q.z + 0.5f)) / (q.z * (6.2f * q.z + 1.7f) + 0.06f);   /* sic, SURVEY Q17 */
                    break;
                }
                case kTerraTonemappingOperatorUncharted2: {
                    f3 ws = uncharted2(mk(11.2f, 11.2f, 11.2f));
                    ws = mk(1.f / ws.x, 1.f / ws.y, 1.f / ws.z);
                    c = had(uncharted2(scl(c, 2.f)), ws);
                    c = pow3(c, 1.f / s->opts.gamma);
                    break;
                }
                default: break;
            }
            fb->pixels[i * fb->width + j] = c;
        }
}

/* ================================================================================ discrete distributions
 * terra_distribution_1d_init (Terra.c:760-779): running float sum in index order -> unnormalised cdf; every
 * entry divided by the total afterwards. */
float oracle_distribution_1d_init(const float* f, size_t n, float* cdf_out) {
    float integral = 0;
    for (size_t i = 0; i < n; ++i) {
        integral += f[i];
        cdf_out[i] = integral;
    }
    for (size_t i = 0; i < n; ++i) cdf_out[i] /= integral;
    return integral;
}

/* terra_distribution_1d_sample (Terra.c:781-809): first bucket with e < cdf[i]; pdf = f[i] / integral; the
 * returned coordinate interpolates inside the bucket, (i + (e - prev) / (cdf[i] - prev)) / n. No bucket
 * (e >= cdf[n-1]): the reference asserts, and returns FLT_MAX under NDEBUG. */
float oracle_distribution_1d_sample(const float* f, const float* cdf, float integral, size_t n, float e,
                                    float* pdf_out, size_t* idx_out) {
    float prev = 0;
    for (size_t i = 0; i < n; ++i) {
        float curr = cdf[i];
        if (e < curr) {
            if (pdf_out) *pdf_out = f[i] / integral;
            if (idx_out) *idx_out = i;
            float d = e - prev;
            d /= curr - prev;
            return (i + d) / n;
        }
        prev = curr;
    }
    return FLT_MAX;
}

/* terra_distribution_2d_init (Terra.c:811-833): one conditional per row; the marginal is the 1-D distribution
 * of the row integrals (built by the same running sum). */
OracleDistribution2D* oracle_distribution_2d_create(const float* f, size_t nx, size_t ny) {
    OracleDistribution2D* d = (OracleDistribution2D*)calloc(1, sizeof *d);
    d->nx = nx; d->ny = ny;
    d->f = (float*)malloc(sizeof(float) * nx * ny);
    d->cdf = (float*)malloc(sizeof(float) * nx * ny);
    d->row_integral = (float*)malloc(sizeof(float) * ny);
    d->marginal_cdf = (float*)malloc(sizeof(float) * ny);
    memcpy(d->f, f, sizeof(float) * nx * ny);
    for (size_t i = 0; i < ny; ++i) d->row_integral[i] = oracle_distribution_1d_init(f + nx * i, nx, d->cdf + nx * i);
    d->integral = oracle_distribution_1d_init(d->row_integral, ny, d->marginal_cdf);
    return d;
}

void oracle_distribution_2d_destroy(OracleDistribution2D* d) {
    if (!d) return;
    free(d->f); free(d->cdf); free(d->row_integral); free(d->marginal_cdf); free(d);
}

/* terra_distribution_2d_sample (Terra.c:835-846): row from the marginal with e1, column from that row's
 * conditional with e2, pdf = product of the two bucket probabilities. */
void oracle_distribution_2d_sample(const OracleDistribution2D* d, size_t n, const float* e1, const float* e2, float* out) {
    for (size_t k = 0; k < n; ++k) {
        float pdfs[2] = { 0.f, 0.f };
        size_t i = 0;
        float s1 = oracle_distribution_1d_sample(d->row_integral, d->marginal_cdf, d->integral, d->ny, e1[k], &pdfs[0], &i);
        float s2 = oracle_distribution_1d_sample(d->f + d->nx * i, d->cdf + d->nx * i, d->row_integral[i], d->nx, e2[k], &pdfs[1], NULL);
        out[4 * k] = s1; out[4 * k + 1] = s2; out[4 * k + 2] = pdfs[0] * pdfs[1]; out[4 * k + 3] = 0.f;
    }
}

/* Importance of a lat-long texel for the env-NEE extension (no reference counterpart: the reference never
 * builds the distribution): luminance with the reference's weights (Terra.c:1036) times the solid-angle
 * factor of the row, sin(theta) at the row centre, theta = terra_PI * (y + 0.5) / ny as in
 * terra_texture_sample_latlong's v = theta / terra_PI * height (Terra.c:468-477). */
void oracle_env_importance(const float* rgb, size_t nx, size_t ny, float* f_out) {
    for (size_t y = 0; y < ny; ++y) {
        const float w = sinf(OPI * (((float)y + 0.5f) / (float)ny));
        for (size_t x = 0; x < nx; ++x) {
            const float* p = rgb + (y * nx + x) * 3;
            const float lum = 0.212671f * p[0] + 0.715160f * p[1] + 0.072169f * p[2];
            f_out[y * nx + x] = lum * w;
        }
    }
}
