/*
 * ref_shim.c — TEST INFRASTRUCTURE ONLY.  Compiled together with the UNMODIFIED reference
 * sources (see Makefile) into oracle/_ref/libterra_ref.so.  It adds nothing to the reference's
 * arithmetic: it only loops the reference's own (externally linked) functions over batches so
 * that tests/ and bench.py's cpu_baseline / --impl reference legs can drive them from Python,
 * and counts terra_bvh_traverse calls through the linker's --wrap (no source change).
 *
 * Reference functions driven from here:
 *   terra_camera_to_world_frame / terra_camera_perspective_sample / terra_ray   Terra.c:1770-1799,1702
 *   terra_scene_raycast                                                          Terra.c:1623-1657
 *   terra_bvh_create / terra_bvh_traverse                                        TerraBVH.c:128-310
 *   terra_ray_state_init / terra_ray_triangle_intersection_query                 TerraGeometry.c:15,159
 *   terra_render                                                                 Terra.c:512-635
 */
#define _GNU_SOURCE
#include <Terra.h>
#include "TerraPrivate.h"
#include "TerraBVH.h"

#include <string.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>
#include <unistd.h>
#include <sys/mman.h>
#include <sys/wait.h>

/* non-static in Terra.c, not declared in any header */
TerraObject*  terra_scene_raycast(void* scene, const TerraRay* ray, const TerraRayState* state,
                                  TerraShadingSurface* surface_out, TerraFloat3* point, size_t* triangle);
TerraFloat3   terra_camera_perspective_sample(const TerraCamera* camera, const TerraFramebuffer* frame,
                                              size_t x, size_t y, float jitter, float r1, float r2);
TerraFloat4x4 terra_camera_to_world_frame(const TerraCamera* camera);
TerraRay      terra_ray(const TerraFloat3* origin, const TerraFloat3* direction);

typedef struct { float t; uint32_t prim; float u, v; } RefHit;

static const RefHit k_miss = { -1.f, 0xFFFFFFFFu, 0.f, 0.f };

/* t,u,v of the winning triangle: re-run the reference's own query on it. */
static RefHit ref_requery(const TerraObject* objects, const TerraRay* offset_ray, uint32_t obj, uint32_t tri) {
    TerraRayState st;
    terra_ray_state_init(offset_ray, &st);
    TerraRayIntersectionQuery q;
    TerraRayIntersectionResult r;
    q.ray = (TerraRay*)offset_ray;
    q.state = &st;
    q.primitive.triangle = objects[obj].triangles + tri;
    RefHit h = k_miss;
    if (terra_ray_triangle_intersection_query(&q, &r)) {
        h.t = r.ray_depth; h.u = r.u; h.v = r.v; h.prim = obj | (tri << 8);
    }
    return h;
}

static TerraRay ref_offset_ray(const TerraRay* ray) {        /* Terra.c:1628-1630 */
    TerraRay o = *ray;
    TerraFloat3 off = terra_mulf3(&o.direction, 0.001f);
    o.origin = terra_addf3(&o.origin, &off);
    return o;
}

/* ---- BVH straight from the reference builder ------------------------------------------- */
void* ref_bvh_build(const TerraObject* objects, int objects_count, int* nodes_count) {
    TerraBVH bvh;
    terra_bvh_create(&bvh, objects, objects_count);
    *nodes_count = bvh.nodes_count;
    return bvh.nodes;
}

void ref_bvh_free(void* nodes) { terra_free(nodes); }

/* n rays (xyz origins, xyz directions) through terra_bvh_traverse, with the raycast offset. */
void ref_bvh_trace(void* nodes, int nodes_count, const TerraObject* objects, size_t n,
                   const float* origins, const float* dirs, RefHit* out) {
    TerraBVH bvh = { (TerraBVHNode*)nodes, nodes_count };
    for (size_t i = 0; i < n; ++i) {
        TerraFloat3 o = terra_f3_setv(origins + 3 * i), d = terra_f3_setv(dirs + 3 * i);
        TerraRay ray = terra_ray(&o, &d);
        TerraRay oray = ref_offset_ray(&ray);
        TerraRayState st;
        terra_ray_state_init(&oray, &st);
        TerraFloat3 p;
        TerraPrimitiveRef prim;
        if (terra_bvh_traverse(&bvh, objects, &oray, &st, &p, &prim)) {
            out[i] = ref_requery(objects, &oray, prim.object_idx, prim.triangle_idx);
        } else {
            out[i] = k_miss;
        }
    }
}

/* Same, timing only (no re-query): returns seconds; hits counted so the loop is not elided. */
double ref_bvh_trace_timed(void* nodes, int nodes_count, const TerraObject* objects, size_t n,
                           const float* origins, const float* dirs, size_t* hit_count) {
    TerraBVH bvh = { (TerraBVHNode*)nodes, nodes_count };
    struct timespec t0, t1;
    size_t hits = 0;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (size_t i = 0; i < n; ++i) {
        TerraFloat3 o = terra_f3_setv(origins + 3 * i), d = terra_f3_setv(dirs + 3 * i);
        TerraRay ray = terra_ray(&o, &d);
        TerraRay oray = ref_offset_ray(&ray);
        TerraRayState st;
        terra_ray_state_init(&oray, &st);
        TerraFloat3 p;
        TerraPrimitiveRef prim;
        hits += terra_bvh_traverse(&bvh, objects, &oray, &st, &p, &prim) ? 1 : 0;
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (hit_count) *hit_count = hits;
    return (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
}

/* ---- primary hits through the committed scene (jitter 0 => deterministic rays) ---------- */
void ref_primary_hits(const TerraCamera* camera, void* scene, const TerraObject* objects_base,
                      size_t width, size_t height, RefHit* out, float* dirs_out) {
    TerraFramebuffer fb;
    memset(&fb, 0, sizeof fb);
    fb.width = width; fb.height = height;
    TerraFloat4x4 rot = terra_camera_to_world_frame(camera);
    for (size_t i = 0; i < height; ++i)
        for (size_t j = 0; j < width; ++j) {
            TerraFloat3 dir = terra_camera_perspective_sample(camera, &fb, j, i, 0.f, 0.f, 0.f);
            dir = terra_transformf3(&rot, &dir);
            TerraRay ray = terra_ray(&camera->position, &dir);
            if (dirs_out) memcpy(dirs_out + 3 * (i * width + j), &dir, 12);
            TerraRayState st;
            terra_ray_state_init(&ray, &st);
            TerraShadingSurface surf;
            TerraFloat3 p;
            size_t tri = 0;
            TerraObject* obj = terra_scene_raycast(scene, &ray, &st, &surf, &p, &tri);
            if (!obj) { out[i * width + j] = k_miss; continue; }
            TerraRay oray = ref_offset_ray(&ray);
            out[i * width + j] = ref_requery(objects_base, &oray, (uint32_t)(obj - objects_base), (uint32_t)tri);
        }
}

/* Surface record at a ray's closest hit (KAT for terra_surface_init). Returns object index or -1. */
int ref_raycast_surface(void* scene, const TerraObject* objects_base, const float* origin, const float* dir,
                        TerraShadingSurface* surf, float* point, int* tri_out) {
    TerraFloat3 o = terra_f3_setv(origin), d = terra_f3_setv(dir);
    TerraRay ray = terra_ray(&o, &d);
    TerraRayState st;
    terra_ray_state_init(&ray, &st);
    TerraFloat3 p;
    size_t tri = 0;
    TerraObject* obj = terra_scene_raycast(scene, &ray, &st, surf, &p, &tri);
    if (!obj) return -1;
    memcpy(point, &p, 12);
    *tri_out = (int)tri;
    return (int)(obj - objects_base);
}

/* ---- ray counter via -Wl,--wrap=terra_bvh_traverse -------------------------------------- */
static uint64_t g_ray_count;
bool __real_terra_bvh_traverse(TerraBVH*, const TerraObject*, const TerraRay*, const TerraRayState*,
                               TerraFloat3*, TerraPrimitiveRef*);
bool __wrap_terra_bvh_traverse(TerraBVH* bvh, const TerraObject* objects, const TerraRay* ray,
                               const TerraRayState* st, TerraFloat3* p, TerraPrimitiveRef* prim) {
    ++g_ray_count;
    return __real_terra_bvh_traverse(bvh, objects, ray, st, p, prim);
}
uint64_t ref_ray_count(void) { return g_ray_count; }
void     ref_ray_count_reset(void) { g_ray_count = 0; }

/* ---- multi-process tile farm: the reference's "one terra_render per tile from N workers"
 *      (satellite/src/Renderer.cpp:316-349) with processes instead of threads, because glibc
 *      rand() serialises threads on a lock.  The scene must be committed.
 *      Renders only tiles whose linear index t satisfies t % tile_stride == 0 (bounded sample).
 *      results_out / pixels_out: width*height records (may be NULL).  Returns wall seconds;
 *      *rays_out = total terra_bvh_traverse calls, *paths_out = pixel-samples rendered. */
double ref_render_farm(const TerraCamera* camera, void* scene, size_t width, size_t height,
                       size_t tile, int nproc, unsigned seed, size_t tile_stride,
                       TerraRawIntegrationResult* results_out, TerraFloat3* pixels_out,
                       uint64_t* rays_out, uint64_t* paths_out, size_t spp) {
    size_t npx = width * height;
    size_t bytes_res = npx * sizeof(TerraRawIntegrationResult), bytes_pix = npx * sizeof(TerraFloat3);
    size_t bytes_cnt = sizeof(uint64_t) * 2 * (size_t)nproc;
    char* shm = mmap(NULL, bytes_res + bytes_pix + bytes_cnt, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (shm == MAP_FAILED) return -1.0;
    memset(shm, 0, bytes_res + bytes_pix + bytes_cnt);
    TerraFramebuffer fb;
    fb.width = width; fb.height = height;
    fb.results = (TerraRawIntegrationResult*)shm;
    fb.pixels = (TerraFloat3*)(shm + bytes_res);
    uint64_t* counters = (uint64_t*)(shm + bytes_res + bytes_pix);
    if (tile_stride == 0) tile_stride = 1;
    size_t tx = (width + tile - 1) / tile, ty = (height + tile - 1) / tile;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int r = 0; r < nproc; ++r) {
        pid_t pid = fork();
        if (pid == 0) {
            srand(seed + (unsigned)r + 1u);
            g_ray_count = 0;
            uint64_t paths = 0;
            size_t k = 0;
            for (size_t t = 0; t < tx * ty; ++t) {
                if (t % tile_stride) continue;
                if ((k++ % (size_t)nproc) != (size_t)r) continue;
                size_t x = (t % tx) * tile, y = (t / tx) * tile;
                size_t w = x + tile > width ? width - x : tile, h = y + tile > height ? height - y : tile;
                terra_render(camera, scene, &fb, x, y, w, h);
                paths += (uint64_t)w * h * spp;
            }
            counters[2 * r] = g_ray_count;
            counters[2 * r + 1] = paths;
            _exit(0);
        }
    }
    for (int r = 0; r < nproc; ++r) { int st; wait(&st); }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    uint64_t rays = 0, paths = 0;
    for (int r = 0; r < nproc; ++r) { rays += counters[2 * r]; paths += counters[2 * r + 1]; }
    if (rays_out) *rays_out = rays;
    if (paths_out) *paths_out = paths;
    if (results_out) memcpy(results_out, fb.results, bytes_res);
    if (pixels_out) memcpy(pixels_out, fb.pixels, bytes_pix);
    munmap(shm, bytes_res + bytes_pix + bytes_cnt);
    return (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
}
