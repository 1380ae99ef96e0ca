// tb_host.cpp — host side of libterra_b200.so: the reference's C API (include/terra_b200.h) on top
// of the CUDA layer in tb_render.cu. Scene container, the reference's BVH build restated, the
// flatten into the GPU node/triangle layout, light table, textures, and the thin terra_render.
//
// Mirrors (behaviour, argument meaning, error conventions) of reference src/Terra.c:130-507,
// 640-663 and src/TerraBVH.c:24-244. Deliberate host-side fixes, none of them arithmetic:
//   - object/light arrays grow by element count (the reference reallocs cap*2 BYTES, Terra.c:144,221)
//   - terra_texture_init* return true (the reference falls off the end, Terra.c:350-366)
//   - terra_scene_commit is the host->device crossing: build, flatten, upload.
#include <assert.h>
#include <float.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <chrono>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include <vector_types.h>

#include "tb_internal.h"
#include "tb_bsdf.h"

// =============================================================================== system API
extern "C" void* terra_malloc(size_t size) { return malloc(size); }                 // Terra.c:641-651
extern "C" void* terra_realloc(void* p, size_t size) { return realloc(p, size); }
extern "C" void  terra_free(void* p) { free(p); }
extern "C" void  terra_log(const char* fmt, ...) {                                  // Terra.c:657-662
    va_list args;
    va_start(args, fmt);
    vfprintf(stdout, fmt, args);
    va_end(args);
    fflush(stdout);
}

[[noreturn]] static void tb_fatal(const char* what) {
    terra_log("terra_b200: fatal: %s\n", what);
    abort();
}

// =============================================================================== presets (host pointers)
// The init functions install these; commit recognises a material by the address of `sample`.
static TbSurface surface_from_abi(const TerraShadingSurface* s, int state_slot) {
    TbSurface t;
    const TerraFloat4* r = s->transform.rows;
    t.xf.r0 = v3(r[0].x, r[0].y, r[0].z);
    t.xf.r1 = v3(r[1].x, r[1].y, r[1].z);
    t.xf.r2 = v3(r[2].x, r[2].y, r[2].z);
    t.n = v3(s->normal.x, s->normal.y, s->normal.z);
    t.emissive = v3(s->emissive.x, s->emissive.y, s->emissive.z);
    t.a0 = v3(s->attributes[0].x, s->attributes[0].y, s->attributes[0].z);
    t.a1 = v3(s->attributes[1].x, s->attributes[1].y, s->attributes[1].z);
    t.a2 = v3(s->attributes[2].x, s->attributes[2].y, s->attributes[2].z);
    t.state = state_slot >= 0 ? s->attributes[state_slot].x : 0.f;
    return t;
}
static TerraFloat3 abi3(V3 v) { TerraFloat3 r = { v.x, v.y, v.z }; return r; }
static V3 v3abi(const TerraFloat3* v) { return v3(v->x, v->y, v->z); }
// The reference's Phong writes its lobe pick through the const surface pointer (TerraPresets.c:87-95).
static void store_state(const TerraShadingSurface* s, int slot, float v) {
    const_cast<TerraShadingSurface*>(s)->attributes[slot].x = v;
}

static TerraFloat3 h_diffuse_sample(const TerraShadingSurface* s, float e1, float e2, float, const TerraFloat3*) {
    TbSurface t = surface_from_abi(s, -1);
    return abi3(tb_diffuse_sample(t, e1, e2));
}
static float h_diffuse_pdf(const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3*) {
    TbSurface t = surface_from_abi(s, -1);
    return tb_diffuse_pdf(t, v3abi(wi));
}
static TerraFloat3 h_diffuse_eval(const TerraShadingSurface* s, const TerraFloat3*, const TerraFloat3*) {
    return abi3(tb_diffuse_eval(v3abi(&s->attributes[TERRA_DIFFUSE_ALBEDO])));
}
static TerraFloat3 h_phong_sample(const TerraShadingSurface* s, float e1, float e2, float e3, const TerraFloat3* wo) {
    TbSurface t = surface_from_abi(s, TERRA_PHONG_SAMPLE_PICK);
    V3 wi = tb_phong_sample(t, e1, e2, e3, v3abi(wo));
    store_state(s, TERRA_PHONG_SAMPLE_PICK, t.state);
    return abi3(wi);
}
static float h_phong_pdf(const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo) {
    TbSurface t = surface_from_abi(s, TERRA_PHONG_SAMPLE_PICK);
    return tb_phong_pdf(t, v3abi(wi), v3abi(wo));
}
static TerraFloat3 h_phong_eval(const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo) {
    TbSurface t = surface_from_abi(s, TERRA_PHONG_SAMPLE_PICK);
    return abi3(tb_phong_eval(t, v3abi(wi), v3abi(wo)));
}
static TerraFloat3 h_ggx_sample(const TerraShadingSurface* s, float e1, float e2, float e3, const TerraFloat3* wo) {
    TbSurface t = surface_from_abi(s, -1);
    return abi3(tb_ggx_sample(t, e1, e2, e3, v3abi(wo)));
}
static float h_ggx_pdf(const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo) {
    TbSurface t = surface_from_abi(s, -1);
    return tb_ggx_pdf(t, v3abi(wi), v3abi(wo));
}
static TerraFloat3 h_ggx_eval(const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo) {
    TbSurface t = surface_from_abi(s, -1);
    return abi3(tb_ggx_eval(t, v3abi(wi), v3abi(wo)));
}
static TerraFloat3 h_glass_sample(const TerraShadingSurface* s, float, float, float e3, const TerraFloat3* wo) {
    TbSurface t = surface_from_abi(s, TERRA_GLASS_STATE);
    V3 wi = tb_glass_sample(t, e3, v3abi(wo));
    store_state(s, TERRA_GLASS_STATE, t.state);
    return abi3(wi);
}
static float h_glass_pdf(const TerraShadingSurface* s, const TerraFloat3*, const TerraFloat3*) {
    return s->attributes[TERRA_GLASS_STATE].x;
}
static TerraFloat3 h_glass_eval(const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3*) {
    TbSurface t = surface_from_abi(s, TERRA_GLASS_STATE);
    return abi3(tb_glass_eval(t, v3abi(wi)));
}

extern "C" void terra_bsdf_diffuse_init(TerraBSDF* b) { b->sample = h_diffuse_sample; b->pdf = h_diffuse_pdf; b->eval = h_diffuse_eval; }
extern "C" void terra_bsdf_phong_init(TerraBSDF* b)   { b->sample = h_phong_sample;   b->pdf = h_phong_pdf;   b->eval = h_phong_eval; }
extern "C" void terra_bsdf_ggx_init(TerraBSDF* b)     { b->sample = h_ggx_sample;     b->pdf = h_ggx_pdf;     b->eval = h_ggx_eval; }
extern "C" void terra_bsdf_glass_init(TerraBSDF* b)   { b->sample = h_glass_sample;   b->pdf = h_glass_pdf;   b->eval = h_glass_eval; }

static int classify_bsdf(const TerraBSDF* b) {
    if (b->sample == h_diffuse_sample) return TB_PRESET_DIFFUSE;
    if (b->sample == h_phong_sample)   return TB_PRESET_PHONG;
    if (b->sample == h_ggx_sample)     return TB_PRESET_GGX;
    if (b->sample == h_glass_sample)   return TB_PRESET_GLASS;
    return -1;
}

// =============================================================================== attributes
extern "C" void terra_attribute_init_constant(TerraAttribute* attr, const TerraFloat3* value) {   // Terra.c:287-292
    attr->eval = NULL;
    attr->state = NULL;
    attr->finalize = NULL;
    attr->value = *value;
}
extern "C" void terra_attribute_init_texture(TerraAttribute* attr, TerraTexture* texture) {       // :294-298
    attr->state = texture;
    attr->eval = terra_texture_sample;
    attr->finalize = terra_texture_finalize;
}
extern "C" void terra_attribute_init_cubemap(TerraAttribute* attr, TerraTexture* texture) {       // :300-304
    attr->state = texture;
    attr->eval = terra_texture_sample_latlong;
    attr->finalize = terra_texture_finalize;
}

// =============================================================================== framebuffer
// Framebuffer memory is page-locked when a CUDA device is present (the copies inside terra_render then run at
// PCIe speed); the struct stays the reference's, and a caller-provided (pageable) buffer works just the same.
static std::mutex g_pinned_lock;
static std::vector<void*> g_pinned;
static void* fb_alloc(size_t bytes) {
    void* p = tb_host_alloc_pinned(bytes);
    if (p) { std::lock_guard<std::mutex> g(g_pinned_lock); g_pinned.push_back(p); return p; }
    return terra_malloc(bytes);
}
static void fb_free(void* p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> g(g_pinned_lock);
        auto it = std::find(g_pinned.begin(), g_pinned.end(), p);
        if (it != g_pinned.end()) { g_pinned.erase(it); tb_host_free_pinned(p); return; }
    }
    terra_free(p);
}
extern "C" bool terra_framebuffer_create(TerraFramebuffer* fb, size_t width, size_t height) {     // :309-326
    if (width == 0 || height == 0) return false;
    fb->width = width;
    fb->height = height;
    fb->pixels = (TerraFloat3*)fb_alloc(sizeof(TerraFloat3) * width * height);
    fb->results = (TerraRawIntegrationResult*)fb_alloc(sizeof(TerraRawIntegrationResult) * width * height);
    terra_framebuffer_clear(fb);
    return true;
}
extern "C" void terra_framebuffer_clear(TerraFramebuffer* fb) {                                    // :328-336
    size_t n = fb->width * fb->height;
    memset(fb->pixels, 0, n * sizeof(TerraFloat3));
    memset(fb->results, 0, n * sizeof(TerraRawIntegrationResult));
}
extern "C" void terra_framebuffer_destroy(TerraFramebuffer* fb) {                                  // :338-345
    if (fb == NULL) return;
    fb_free(fb->results);
    fb_free(fb->pixels);
}

// =============================================================================== textures
#include "tb_texture.h"

static TbDevTexture texture_view(const TerraTexture* t) {
    TbDevTexture v;
    v.pixels = t->pixels;
    v.width = t->width; v.height = t->height;
    v.components = t->components; v.depth = t->depth;
    v.filter = t->filter; v.address_mode = t->address_mode;
    v.pad_ = 0;
    return v;
}
static bool texture_init_common(TerraTexture* texture, size_t width, size_t height, size_t components,
                                const void* data, size_t depth) {                                 // Terra.c:350-366
    size_t bytes = depth * width * height * components;
    texture->pixels = terra_malloc(bytes);
    memcpy(texture->pixels, data, bytes);
    texture->width = (uint16_t)width;
    texture->height = (uint16_t)height;
    texture->components = (uint8_t)components;
    texture->depth = (uint8_t)depth;
    return true;    // filter / address_mode stay the caller's job, as in the reference
}
extern "C" bool terra_texture_init(TerraTexture* t, size_t w, size_t h, size_t c, const void* data) {
    return texture_init_common(t, w, h, c, data, 1);
}
extern "C" bool terra_texture_init_hdr(TerraTexture* t, size_t w, size_t h, size_t c, const float* data) {
    return texture_init_common(t, w, h, c, data, 4);
}
extern "C" TerraFloat3 terra_texture_read(TerraTexture* texture, size_t x, size_t y) {
    return abi3(tb_texture_read(texture_view(texture), x, y));
}
extern "C" TerraFloat3 terra_texture_sample(void* texture, const void* uv, const void*) {
    const TerraFloat2* t = (const TerraFloat2*)uv;
    return abi3(tb_texture_sample(texture_view((const TerraTexture*)texture), t->x, t->y));
}
extern "C" TerraFloat3 terra_texture_sample_latlong(void* texture, const void* dir, const void*) {
    return abi3(tb_texture_sample_latlong(texture_view((const TerraTexture*)texture), v3abi((const TerraFloat3*)dir)));
}
extern "C" void terra_texture_destroy(TerraTexture* texture) {                                     // :479-482
    terra_free(texture->pixels);
    texture->pixels = NULL;
}
extern "C" void terra_texture_finalize(void* p) {                                                  // :484-507 (sRGB -> linear)
    TerraTexture* t = (TerraTexture*)p;
    if (t == NULL || t->pixels == NULL) return;
    size_t n = (size_t)t->width * t->height * t->components;
    if (t->depth == 1) {
        uint8_t* px = (uint8_t*)t->pixels;
        for (size_t i = 0; i < n; ++i) px[i] = (uint8_t)(powf(px[i] / 255.f, 2.2f) * 255);
    } else if (t->depth == 4) {
        float* px = (float*)t->pixels;
        for (size_t i = 0; i < n; ++i) px[i] = powf(px[i], 2.2f);
    } else {
        assert(false);
    }
}

// =============================================================================== scene container
struct TbLightHost {                 // reference TerraLight, src/TerraPrivate.h:24-29
    TerraFloat3 power;
    float       area;
    size_t      object;
    std::vector<float> triangle_area;
};

// One issuing thread per device of a multi-GPU scene. A fanned-out terra_render launches ~45 kernels per device; issued
// from one host thread, device 7 of 8 started 1.3 ms after device 0 (r2m: e2e 13.8 ms against 12.3 ms of device work on
// C2 at N = 8). The threads live as long as the scene; run() hands every thread the same job and returns when all are done.
struct TbFanPool {
    std::vector<std::thread> threads;
    std::mutex m;
    std::condition_variable go, done;
    std::function<void(int)> job;
    uint64_t generation;
    int pending;
    bool quit;
    explicit TbFanPool(int n) : generation(0), pending(0), quit(false) {
        for (int g = 0; g < n; ++g) threads.emplace_back([this, g] { main(g); });
    }
    ~TbFanPool() {
        { std::lock_guard<std::mutex> l(m); quit = true; }
        go.notify_all();
        for (std::thread& t : threads) t.join();
    }
    void main(int g) {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> l(m);
        while (true) {
            go.wait(l, [&] { return quit || generation != seen; });
            if (quit) return;
            seen = generation;
            l.unlock();
            job(g);
            l.lock();
            if (--pending == 0) done.notify_one();
        }
    }
    void run(const std::function<void(int)>& f) {
        std::unique_lock<std::mutex> l(m);
        job = f;
        pending = (int)threads.size();
        ++generation;
        go.notify_all();
        done.wait(l, [&] { return pending == 0; });
    }
};

struct TbScene {
    TerraSceneOptions opts;          // committed
    TerraSceneOptions new_opts;      // what terra_scene_get_options hands out (Terra.c:24-43)
    TerraObject*      objects;
    size_t            objects_pop, objects_cap;
    std::vector<TbLightHost> lights;
    size_t            lights_triangles_count;
    bool              dirty_objects, dirty_lights, has_bvh;
    std::vector<TbHostNode> nodes;   // the reference-layout BVH
    int               stack_bound;
    int               bvh_mode;
    float             commit_ms[4];   // last commit: tree build, flatten, quantise, upload
    std::vector<uint32_t> rank_of_triangle;   // of the last flatten: [object_base[obj] + tri] -> rank (the light table needs it)
    std::vector<size_t>   object_base;
    std::vector<uint8_t>  presets;            // preset per object as tagged into the uploaded triangle records
    bool              host_boxes_stale;       // terra_b200_refit moved the boxes on the device only
    TbDevice*         dev;            // primary device: every entry point but a fanned-out terra_render runs here
    std::vector<int>  device_ids;     // devices of this scene (terra_b200_scene_set_devices / TERRA_B200_DEVICES), [0] = primary
    std::vector<TbDevice*> devs;      // one per device_ids entry once committed; devs[0] == dev
    bool              peers_ok;       // every pair of the scene's devices can map the other's memory
    size_t            fanout_min_paths;   // terra_render calls of at least this many paths are split over the devices
    TbFanPool*        fan;            // issuing threads of the fan-out, created by the first call that fans out
    size_t            fanout_chunk_paths; // paths per device in one row chunk of a fanned-out call (TERRA_B200_FANOUT_CHUNK)
    uint64_t          seed;
    bool              seed_fixed;
    uint64_t          render_calls;
    std::mutex        lock;
};

static int g_device_index = 0;
// Devices a new scene renders on: TERRA_B200_DEVICES="0,1,2,3" (or "all"), else the one terra_b200_set_device chose.
// The scene is replicated on each of them at commit; a large terra_render call is split over them by sample range.
static std::vector<int> default_device_ids() {
    std::vector<int> ids;
    const char* e = getenv("TERRA_B200_DEVICES");
    if (e && *e) {
        if (strcmp(e, "all") == 0) {
            const int n = tb_cuda_device_count();
            for (int i = 0; i < n; ++i) ids.push_back(i);
        } else {
            for (const char* p = e; *p;) {
                char* end = NULL;
                const long v = strtol(p, &end, 10);
                if (end == p) break;
                ids.push_back((int)v);
                p = *end == ',' ? end + 1 : end;
            }
        }
    }
    if (ids.empty()) ids.push_back(g_device_index);
    return ids;
}
// Acceleration structure the next commit builds: 0 = the reference's tree (x-only SAH, TerraBVH.c),
// 1 = binned SAH over three axes in the same node format (default; see tb_bvh.cpp).
// 2 = linear BVH built on the GPU (tb_lbvh.cu): the fast-build mode for large scenes.
enum { TB_BVH_REFERENCE = 0, TB_BVH_SAH3 = 1, TB_BVH_LBVH = 2 };
extern "C" void tb_lbvh_build(int device_index, const TerraObject* objects, int objects_count, std::vector<TbHostNode>* nodes_out, float* ms_out);
static float g_lbvh_ms[2] = { 0.f, 0.f };
static int g_bvh_mode = -1;
static int bvh_mode() {
    if (g_bvh_mode >= 0) return g_bvh_mode;
    const char* e = getenv("TERRA_B200_BVH");
    if (e && strcmp(e, "reference") == 0) return TB_BVH_REFERENCE;
    if (e && strcmp(e, "lbvh") == 0) return TB_BVH_LBVH;
    return TB_BVH_SAH3;
}
static void build_bvh(TbScene* s, int mode);
static void sync_host_nodes(TbScene* s);
enum { TB_PREALLOCATED_OBJECTS = 64 };

extern "C" HTerraScene terra_scene_create(void) {                                                   // Terra.c:130-138
    TbScene* s = new TbScene();
    memset(&s->opts, 0, sizeof s->opts);
    memset(&s->new_opts, 0, sizeof s->new_opts);
    s->objects = (TerraObject*)terra_malloc(sizeof(TerraObject) * TB_PREALLOCATED_OBJECTS);
    s->objects_pop = 0;
    s->objects_cap = TB_PREALLOCATED_OBJECTS;
    s->lights_triangles_count = 0;
    s->dirty_objects = s->dirty_lights = s->has_bvh = false;
    s->stack_bound = 0;
    s->bvh_mode = -1;
    memset(s->commit_ms, 0, sizeof s->commit_ms);
    s->host_boxes_stale = false;
    s->dev = NULL;
    s->device_ids = default_device_ids();
    s->peers_ok = false;
    s->fan = NULL;
    s->fanout_chunk_paths = (size_t)1 << 40;            // off: see terra_render
    if (const char* e = getenv("TERRA_B200_FANOUT_CHUNK")) s->fanout_chunk_paths = (size_t)strtoull(e, NULL, 10);
    if (s->fanout_chunk_paths < 1) s->fanout_chunk_paths = 1;
    s->fanout_min_paths = (size_t)1 << 24;
    if (const char* e = getenv("TERRA_B200_FANOUT_MIN")) s->fanout_min_paths = (size_t)strtoull(e, NULL, 10);
    s->seed = 0;
    s->seed_fixed = false;
    s->render_calls = 0;
    return s;
}

extern "C" TerraObject* terra_scene_add_object(HTerraScene h, size_t triangles_count) {            // Terra.c:140-155
    TbScene* s = (TbScene*)h;
    if (s->objects_pop == 256) tb_fatal("more than 256 objects: the 8-bit object index of the leaf packing overflows (TerraBVH.c:155)");
    if (triangles_count > ((size_t)1 << 23)) tb_fatal("more than 2^23 triangles in one object: leaf packing i << 8 overflows int32 (TerraBVH.c:155)");
    if (s->objects_pop == s->objects_cap) {
        s->objects_cap *= 2;      // element count, not bytes (reference bug)
        s->objects = (TerraObject*)terra_realloc(s->objects, sizeof(TerraObject) * s->objects_cap);
    }
    TerraObject* o = &s->objects[s->objects_pop];
    memset(o, 0, sizeof *o);
    o->triangles = (TerraTriangle*)terra_malloc(sizeof(TerraTriangle) * triangles_count);
    o->properties = (TerraTriangleProperties*)terra_malloc(sizeof(TerraTriangleProperties) * triangles_count);
    o->triangles_count = triangles_count;
    s->dirty_objects = true;
    s->dirty_lights = true;
    ++s->objects_pop;
    return o;
}

extern "C" size_t terra_scene_count_objects(HTerraScene h) { return ((TbScene*)h)->objects_pop; }
extern "C" TerraSceneOptions* terra_scene_get_options(HTerraScene h) { return &((TbScene*)h)->new_opts; }

extern "C" void terra_scene_clear(HTerraScene h) {                                                  // Terra.c:238-249
    TbScene* s = (TbScene*)h;
    for (size_t i = 0; i < s->objects_pop; ++i) {      // the reference leaks these ("TODO also free memory?")
        terra_free(s->objects[i].triangles);
        terra_free(s->objects[i].properties);
    }
    s->objects_pop = 0;
    s->lights.clear();
    s->dirty_objects = true;
}

extern "C" void terra_scene_destroy(HTerraScene h) {                                                // Terra.c:257-282
    TbScene* s = (TbScene*)h;
    if (s == NULL) { assert(false); return; }
    for (size_t i = 0; i < s->objects_pop; ++i) {
        terra_free(s->objects[i].triangles);
        terra_free(s->objects[i].properties);
    }
    terra_free(s->objects);
    delete s->fan;
    for (TbDevice* d : s->devs) tb_device_destroy(d);
    delete s;
}

// =============================================================================== commit
#include "tb_bvh.h"

static float triangle_area(const TerraTriangle* t) {                                                // Terra.c:1833-1838
    V3 ab = vsub(v3abi(&t->b), v3abi(&t->a));
    V3 ac = vsub(v3abi(&t->c), v3abi(&t->a));
    return vlen(vcross(ab, ac)) / 2;
}

struct TbTextureTable {
    std::vector<const TerraTexture*> list;
    int index_of(const TerraTexture* t) {
        for (size_t i = 0; i < list.size(); ++i) if (list[i] == t) return (int)i;
        list.push_back(t);
        return (int)list.size() - 1;
    }
};

// Function-pointer attributes cannot cross to the device: classify by pointer identity.
static TbDevAttr classify_attribute(const TerraAttribute* a, TbTextureTable* table, const char* what) {
    TbDevAttr d;
    d.vx = a->value.x; d.vy = a->value.y; d.vz = a->value.z;
    d.kind_tex = TB_ATTR_CONST;
    if (a->state == NULL) return d;                                                                 // Terra.c:1805
    int kind;
    if (a->eval == terra_texture_sample) kind = TB_ATTR_TEX2D;
    else if (a->eval == terra_texture_sample_latlong) kind = TB_ATTR_LATLONG;
    else {
        terra_log("terra_b200: %s attribute has an eval callback that is not one of this library's "
                  "(terra_texture_sample / terra_texture_sample_latlong)\n", what);
        tb_fatal("custom attribute callbacks cannot run on the GPU and there is no CPU fallback");
    }
    const TerraTexture* t = (const TerraTexture*)a->state;
    if (t->pixels == NULL) tb_fatal("attribute bound to a texture without pixels");
    d.kind_tex = kind | (table->index_of(t) << 2);
    return d;
}

static V3 host_attr_eval(const TerraAttribute* a, V2 uv) {                                          // Terra.c:1804-1810
    if (a->state == NULL) return v3abi(&a->value);
    TerraFloat2 t = { uv.x, uv.y };
    TerraFloat3 r = a->eval(a->state, &t, NULL);
    return v3abi(&r);
}

static void build_bvh(TbScene* s, int mode) {
    if (mode == TB_BVH_REFERENCE) tb_bvh_build(s->objects, (int)s->objects_pop, s->nodes);
    else if (mode == TB_BVH_LBVH) {
        if (tb_cuda_device_count() <= 0) tb_fatal("the lbvh build runs on the GPU: no CUDA device (there is no CPU fallback)");
        tb_lbvh_build(g_device_index, s->objects, (int)s->objects_pop, &s->nodes, g_lbvh_ms);
    }
    else tb_bvh_build_sah3(s->objects, (int)s->objects_pop, s->nodes);
    s->bvh_mode = mode;
    s->host_boxes_stale = false;
}

// After terra_b200_refit the host tree still holds the boxes of the last build. Anything that reads s->nodes
// (a commit that re-flattens without rebuilding, terra_b200_bvh_nodes) first brings the device's boxes back.
static void sync_host_nodes(TbScene* s) {
    if (!s->host_boxes_stale || !s->dev) return;
    std::vector<float> flat(s->nodes.size() * 16);
    tb_device_download_nodes(s->dev, flat.data(), s->nodes.size());
    for (size_t i = 0; i < s->nodes.size(); ++i) memcpy(&s->nodes[i].aabb[0], &flat[i * 16], 48);
    s->host_boxes_stale = false;
}

extern "C" void terra_scene_commit(HTerraScene h) {                                                 // Terra.c:162-236
    TbScene* s = (TbScene*)h;
    std::lock_guard<std::mutex> guard(s->lock);
    bool dirty_accelerator = s->dirty_objects || !s->has_bvh || s->bvh_mode != bvh_mode();
    if (s->opts.accelerator != s->new_opts.accelerator) dirty_accelerator = true;
    s->opts = s->new_opts;
    if (s->opts.accelerator != kTerraAcceleratorBVH) { assert(false); return; }
    size_t total = 0;
    for (size_t i = 0; i < s->objects_pop; ++i) total += s->objects[i].triangles_count;
    if (total < 2) tb_fatal("terra_scene_commit needs at least 2 triangles (the reference builder reads out of range below that)");
    if (!s->dev) {
        if (s->device_ids.size() == 1 && !getenv("TERRA_B200_DEVICES")) s->device_ids[0] = g_device_index;   // terra_b200_set_device after create
        for (int id : s->device_ids) s->devs.push_back(tb_device_create(id));
        s->dev = s->devs[0];
        s->peers_ok = s->devs.size() > 1 && tb_device_enable_peers(s->devs.data(), (int)s->devs.size()) != 0;
        if (s->devs.size() > 1 && !s->peers_ok)
            terra_log("terra_b200: the scene's devices cannot map each other's memory: terra_render stays on device %d\n", s->device_ids[0]);
    }

    if (s->dirty_lights) {                                                                          // Terra.c:194-231
        s->lights.clear();
        for (size_t i = 0; i < s->objects_pop; ++i) {
            V2 mid = { 0.5f, 0.5f };
            V3 e = host_attr_eval(&s->objects[i].material.emissive, mid);
            if (e.x == 0 && e.y == 0 && e.z == 0) continue;
            TbLightHost l;
            float area = 0;
            l.triangle_area.resize(s->objects[i].triangles_count);
            for (size_t j = 0; j < s->objects[i].triangles_count; ++j) {
                float a = triangle_area(&s->objects[i].triangles[j]);
                l.triangle_area[j] = a;
                area += a;
            }
            l.power = abi3(vmul(e, area * TB_PI));
            l.area = area;
            l.object = i;
            s->lights_triangles_count += s->objects[i].triangles_count;   // never reset, as in the reference
            s->lights.push_back(l);
        }
    }

    // ---- materials and textures are classified at EVERY commit: the reference reads object->material live at every
    //      hit (Terra.c:1745-1762), so an in-place material / emissive edit followed by a commit must show
    TbTextureTable table;
    TbDevAttr env = classify_attribute(&s->opts.environment_map, &table, "environment_map");
    std::vector<TbDevMaterial> mats(s->objects_pop);
    std::vector<uint8_t> presets(s->objects_pop);
    for (size_t i = 0; i < s->objects_pop; ++i) {
        const TerraMaterial* m = &s->objects[i].material;
        int preset = classify_bsdf(&m->bsdf);
        if (preset < 0) {
            terra_log("terra_b200: object %zu: material.bsdf is not one of terra_bsdf_{diffuse,phong,ggx,glass}_init\n", i);
            tb_fatal("custom BSDF callbacks cannot run on the GPU and there is no CPU fallback");
        }
        if (m->attributes_count > TERRA_MATERIAL_MAX_ATTRIBUTES) tb_fatal("attributes_count > TERRA_MATERIAL_MAX_ATTRIBUTES");
        memset(&mats[i], 0, sizeof mats[i]);
        mats[i].preset = preset;
        mats[i].attr_count = (int)m->attributes_count;
        mats[i].ior = m->ior;
        mats[i].light_index = -1;
        mats[i].emissive = classify_attribute(&m->emissive, &table, "emissive");
        for (size_t k = 0; k < m->attributes_count; ++k) mats[i].attr[k] = classify_attribute(&m->attributes[k], &table, "material");
        presets[i] = (uint8_t)preset;
    }
    // the triangle records carry object | preset << 8 (the trace kernels bin hits by material without touching the
    // material table): a preset swapped in place re-tags them, i.e. takes the geometry path without a rebuild
    const bool retag = presets != s->presets;
    typedef std::chrono::steady_clock Clock;
    auto ms_since = [](Clock::time_point t0) { return std::chrono::duration<float, std::milli>(Clock::now() - t0).count(); };
    if (dirty_accelerator || retag) {
        Clock::time_point t0 = Clock::now();
        if (dirty_accelerator) build_bvh(s, bvh_mode());
        else sync_host_nodes(s);                        // a refit left the up-to-date boxes on the device only
        s->host_boxes_stale = false;
        s->commit_ms[0] = ms_since(t0);
        t0 = Clock::now();
        TbFlatBvh flat;
        tb_bvh_flatten(s->nodes, s->objects, (int)s->objects_pop, &flat);
        s->commit_ms[1] = ms_since(t0);
        s->stack_bound = flat.stack_bound;
        for (size_t r = 0; r < total; ++r) {
            uint32_t obj;
            memcpy(&obj, &flat.tris[r * 12 + 7], 4);
            uint32_t tag = obj | ((uint32_t)mats[obj].preset << 8);
            memcpy(&flat.tris[r * 12 + 7], &tag, 4);
        }
        TbFlatScene up;
        memset(&up, 0, sizeof up);
        up.nodes = flat.nodes.data();   up.n_nodes = s->nodes.size();
        up.tris = flat.tris.data();     up.n_tris = total;
        up.props = flat.props.data();
        up.stack_bound = flat.stack_bound;
        up.exact_boxes = s->bvh_mode == TB_BVH_REFERENCE ? 1 : 0;
        t0 = Clock::now();
        TbQuantBvh quant;
        if (!up.exact_boxes) {
            tb_bvh_quantize(flat, &quant);
            up.qnodes = quant.nodes.data();
            memcpy(up.grid_lo, quant.grid_lo, sizeof up.grid_lo);
            memcpy(up.grid_cell, quant.grid_cell, sizeof up.grid_cell);
        }
        TbQuadBvh quad;
        if (!up.exact_boxes && s->nodes.size() > 0) {
            tb_bvh4_build(flat, quant, &quad);
            up.q4nodes = quad.nodes.data(); up.q4src = quad.src.data();
            up.n_q4 = quad.src.size() / 4; up.stack_bound4 = quad.stack_bound;
        }
        s->commit_ms[2] = ms_since(t0);
        t0 = Clock::now();
        tb_device_upload_geometry(s->dev, &up);
        // replicas: device to device (NVLink) when the devices are peers, else a second upload from the host
        for (size_t g = 1; g < s->devs.size(); ++g) {
            if (s->peers_ok) tb_device_clone_geometry(s->devs[g], s->dev);
            else tb_device_upload_geometry(s->devs[g], &up);
        }
        s->commit_ms[3] = ms_since(t0);
        s->rank_of_triangle.swap(flat.rank_of_triangle);
        s->object_base.swap(flat.object_base);
        s->presets = presets;
        s->has_bvh = true;
    }
    {
        std::vector<TbDevLight> lights(s->lights.size());
        std::vector<float> light_area;
        std::vector<uint32_t> light_rank;
        for (size_t l = 0; l < s->lights.size(); ++l) {
            size_t obj = s->lights[l].object;
            lights[l].object = (int32_t)obj;
            lights[l].tri_count = (uint32_t)s->objects[obj].triangles_count;
            lights[l].tri_base = (uint32_t)light_area.size();
            lights[l].area = s->lights[l].area;
            mats[obj].light_index = (int32_t)l;
            for (size_t j = 0; j < s->objects[obj].triangles_count; ++j) {
                light_area.push_back(s->lights[l].triangle_area[j]);
                light_rank.push_back(s->rank_of_triangle[s->object_base[obj] + j]);
            }
        }
        TbFlatScene up;
        memset(&up, 0, sizeof up);
        up.materials = mats.data();     up.n_objects = s->objects_pop;
        up.lights = lights.data();      up.n_lights = lights.size();
        up.light_tri_area = light_area.data();
        up.light_tri_rank = light_rank.data();
        up.n_light_tris = light_area.size();
        up.textures = table.list.data(); up.n_textures = table.list.size();
        up.lights_triangles_count = s->lights_triangles_count;
        up.env = env;
        Clock::time_point t0 = Clock::now();
        for (TbDevice* d : s->devs) tb_device_upload_tables(d, &up);
        if (!(dirty_accelerator || retag)) s->commit_ms[3] = ms_since(t0);
    }
    for (TbDevice* d : s->devs) tb_device_set_options(d, &s->opts, &env);
    s->dirty_objects = false;
    s->dirty_lights = false;
}

// =============================================================================== render
static TbCameraFrame camera_frame(const TerraCamera* c) {                                           // Terra.c:1770-1781, 1794
    V3 zaxis = vnorm(v3abi(&c->direction));
    V3 xaxis = vnorm(vcross(v3abi(&c->up), zaxis));
    V3 yaxis = vcross(zaxis, xaxis);
    TbCameraFrame f;
    f.rot[0][0] = xaxis.x; f.rot[0][1] = yaxis.x; f.rot[0][2] = zaxis.x;
    f.rot[1][0] = xaxis.y; f.rot[1][1] = yaxis.y; f.rot[1][2] = zaxis.y;
    f.rot[2][0] = xaxis.z; f.rot[2][1] = yaxis.z; f.rot[2][2] = zaxis.z;
    f.pos[0] = c->position.x; f.pos[1] = c->position.y; f.pos[2] = c->position.z;
    f.tan_half_fov = (float)tan((c->fov * 0.0174533f) / 2);
    return f;
}

static size_t effective_spp(const TerraSceneOptions* o) {                                           // Terra.c:516-527
    size_t spp = o->samples_per_pixel;
    if (o->sampling_method == kTerraSamplingMethodStratified) {
        size_t cur = o->strata * o->strata;
        if (cur < 2) return spp > cur ? spp : cur;     // the reference loops forever on strata <= 1; do not
        while (spp > cur) cur *= cur;
        spp = cur;
    }
    return spp;
}

static uint64_t next_seed(TbScene* s) {
    uint64_t call = s->render_calls++;
    uint64_t base = s->seed_fixed ? s->seed : ((uint64_t)time(NULL) ^ (uint64_t)(uintptr_t)&exit);  // Terra.c:679
    return base * 0x9E3779B97F4A7C15ull + call * 0xD1B54A32D192ED03ull + 1;
}

static TbScene* committed(HTerraScene h, const char* who) {
    TbScene* s = (TbScene*)h;
    if (!s || !s->dev || !s->has_bvh) {
        terra_log("terra_b200: %s called on a scene that was never committed\n", who);
        tb_fatal("scene not committed");
    }
    return s;
}

extern "C" void terra_render(const TerraCamera* camera, HTerraScene h, const TerraFramebuffer* fb,
                             size_t x, size_t y, size_t width, size_t height) {                     // Terra.c:512-635
    TbScene* s = committed(h, "terra_render");
    if (width == 0 || height == 0) return;
    if (x + width > fb->width || y + height > fb->height) tb_fatal("terra_render region outside the framebuffer");
    std::lock_guard<std::mutex> guard(s->lock);
    TbCameraFrame cam = camera_frame(camera);
    const size_t spp = effective_spp(&s->opts);
    const uint64_t seed = next_seed(s);
    const size_t n_dev = s->devs.size();
    if (n_dev > 1 && s->peers_ok && width * height * spp >= s->fanout_min_paths && n_dev <= 16) {
        // ---- the multi-GPU splitter behind the reference's call: the scene is replicated, device g
        //      renders sample range g of the whole region, then finalises and returns row band g — a reduce-scatter
        //      over NVLink fused into the accumulate / tonemap kernel, and every PCIe link carries 1/n of the
        //      framebuffer in each direction (tb_render.cu, tb_device_partial_*). One host thread per device issues
        //      that device's launches (TbFanPool); the second phase starts when every device has RECORDED the event
        //      the others wait on (a wait on an event not yet recorded would be a no-op).
        //      A large region goes through as TWO row chunks, the first seven eighths and the rest: the framebuffer
        //      traffic of the first (PCIe in and out, NVLink reduce) runs beside the rendering of the second, and only
        //      an eighth of it is left exposed at the end. Device g finalises and returns slice g of each chunk.
        //      OFF by default (terra_b200_set_fanout_chunk_paths / TERRA_B200_FANOUT_CHUNK turn it on). Measured on C2,
        //      r2o: a chunk costs ~0.2 ms of launch tails and short last-bounce waves, which is what the hidden
        //      transfers are worth — two chunks 49.02 against 48.91 ms at N = 2, 13.36 against 13.20 ms at N = 8
        //      (equal chunks of 16M paths: 50.5 ms at N = 2).
        std::vector<size_t> base(n_dev + 1);
        for (size_t g = 0; g <= n_dev; ++g) base[g] = spp * g / n_dev;
        std::vector<size_t> crow;
        crow.push_back(0);
        if (width * height * spp / n_dev >= 2 * s->fanout_chunk_paths && height >= 16) {
            size_t tail = (height / 8 + 3) & ~(size_t)3;                                  // 8x4 pixel tiles stay whole
            crow.push_back(height - tail);
        }
        crow.push_back(height);
        const size_t n_chunks = crow.size() - 1;
        size_t stage_px = 0;                                                              // slices of one device, all chunks
        for (size_t c = 0; c < n_chunks; ++c) stage_px += ((crow[c + 1] - crow[c]) / n_dev + 1) * width;
        if (!s->fan || s->fan->threads.size() != n_dev) { delete s->fan; s->fan = new TbFanPool((int)n_dev); }
        for (size_t c = 0; c < n_chunks; ++c) {
            const size_t rows = crow[c + 1] - crow[c];
            auto slice0 = [&](size_t g) { return crow[c] + rows * g / n_dev; };
            s->fan->run([&](int g) {
                tb_device_partial_begin(s->devs[g], &cam, fb, x, y, width, height, c, crow[c], rows, slice0(g), slice0(g + 1) - slice0(g),
                                        stage_px, base[g], base[g + 1] - base[g], seed);
            });
            s->fan->run([&](int g) {
                tb_device_partial_finish(s->devs[g], s->devs.data(), (int)n_dev, fb, x, y, width, height, slice0(g), slice0(g + 1) - slice0(g),
                                         spp, c + 1 == n_chunks);
            });
        }
        for (size_t g = 0; g < n_dev; ++g) tb_device_partial_wait(s->devs[g], s->devs.data(), (int)n_dev);
        return;
    }
    tb_device_render(s->dev, &cam, fb, x, y, width, height, spp, seed);
}

// =============================================================================== async tile jobs
// The satellite front-end renders a frame as square tiles pushed to worker threads that each call
// terra_render and report tile begin / end to the UI thread, with an atomic count of tiles left
// (satellite/src/Renderer.cpp:70-98, 316-349). The same contract as a library call: the region is cut into
// tiles in the same row-major order, ONE worker thread per job walks them (a tile already fills the GPU;
// the scene lock serialises jobs that share a scene), and the caller polls, waits or cancels.
struct TerraB200Job {
    TerraCamera camera;
    HTerraScene scene;
    const TerraFramebuffer* fb;
    size_t x, y, w, h, tile;
    TerraB200TileEvent on_begin, on_end;
    void* user;
    std::atomic<int> tiles_left;
    std::atomic<bool> cancel;
    std::thread worker;
};

static void job_main(TerraB200Job* j) {
    const size_t t = j->tile ? j->tile : (j->w > j->h ? j->w : j->h);
    for (size_t ty = 0; ty < j->h; ty += t)
        for (size_t tx = 0; tx < j->w; tx += t) {
            const size_t tw = tx + t < j->w ? t : j->w - tx, th = ty + t < j->h ? t : j->h - ty;
            if (!j->cancel.load()) {
                if (j->on_begin) j->on_begin(j->x + tx, j->y + ty, tw, th, j->user);
                terra_render(&j->camera, j->scene, j->fb, j->x + tx, j->y + ty, tw, th);
                if (j->on_end) j->on_end(j->x + tx, j->y + ty, tw, th, j->user);
            }
            j->tiles_left.fetch_sub(1);
        }
}

extern "C" HTerraB200Job terra_b200_render_async(const TerraCamera* camera, HTerraScene h, const TerraFramebuffer* fb,
                                                 size_t x, size_t y, size_t width, size_t height, size_t tile_size,
                                                 TerraB200TileEvent on_tile_begin, TerraB200TileEvent on_tile_end, void* user) {
    committed(h, "terra_b200_render_async");
    if (x + width > fb->width || y + height > fb->height) tb_fatal("terra_b200_render_async region outside the framebuffer");
    TerraB200Job* j = new TerraB200Job();
    j->camera = *camera; j->scene = h; j->fb = fb;
    j->x = x; j->y = y; j->w = width; j->h = height; j->tile = tile_size;
    j->on_begin = on_tile_begin; j->on_end = on_tile_end; j->user = user;
    const size_t t = tile_size ? tile_size : (width > height ? width : height);
    const size_t nx = t ? (width + t - 1) / t : 0, ny = t ? (height + t - 1) / t : 0;
    j->tiles_left.store((int)(nx * ny));
    j->cancel.store(false);
    j->worker = std::thread(job_main, j);
    return j;
}
extern "C" int terra_b200_job_poll(HTerraB200Job j) { return j->tiles_left.load(); }
extern "C" void terra_b200_job_wait(HTerraB200Job j) { if (j->worker.joinable()) j->worker.join(); }
extern "C" void terra_b200_job_cancel(HTerraB200Job j) { j->cancel.store(true); terra_b200_job_wait(j); }
extern "C" void terra_b200_job_release(HTerraB200Job j) { terra_b200_job_wait(j); delete j; }

// =============================================================================== extensions
extern "C" int  terra_b200_device_count(void) { return tb_cuda_device_count(); }
extern "C" void terra_b200_set_device(int device) { g_device_index = device; }
extern "C" double terra_b200_probe_bandwidth(size_t bytes, int mode) { return tb_device_probe_bandwidth(g_device_index, bytes, mode); }
extern "C" void terra_b200_set_seed(HTerraScene h, uint64_t seed) {
    TbScene* s = (TbScene*)h;
    s->seed = seed; s->seed_fixed = true; s->render_calls = 0;
}
extern "C" void terra_b200_get_stats(HTerraScene h, TerraB200Stats* out) {      // summed over the scene's devices
    TbScene* s = committed(h, "terra_b200_get_stats");
    memset(out, 0, sizeof *out);
    for (TbDevice* d : s->devs) {
        TerraB200Stats t;
        tb_device_get_stats(d, &t);
        out->kernel_launches += t.kernel_launches; out->rays_traced += t.rays_traced; out->paths_traced += t.paths_traced;
        out->trace_kernel_ms += t.trace_kernel_ms; out->trace_kernel_launches += t.trace_kernel_launches;
        out->render_ms = t.render_ms > out->render_ms ? t.render_ms : out->render_ms;     // the devices of one call run side by side
        out->h2d_bytes += t.h2d_bytes; out->d2h_bytes += t.d2h_bytes; out->nodes_visited += t.nodes_visited; out->tris_tested += t.tris_tested;
        out->graph_replays += t.graph_replays; out->coherent_rays += t.coherent_rays; out->coherent_kernel_ms += t.coherent_kernel_ms;
        out->coherent_kernel_launches += t.coherent_kernel_launches; out->coherent_nodes_visited += t.coherent_nodes_visited;
        out->coherent_tris_tested += t.coherent_tris_tested; out->p2p_bytes += t.p2p_bytes;
    }
}
extern "C" void terra_b200_round_stats(HTerraScene h, uint64_t* out10) { tb_device_round_stats(committed(h, "terra_b200_round_stats")->dev, out10); }
extern "C" const char* terra_b200_incoherent_kernel(HTerraScene h) { return tb_device_incoherent_kernel(committed(h, "terra_b200_incoherent_kernel")->dev); }
extern "C" void terra_b200_reset_stats(HTerraScene h) { for (TbDevice* d : committed(h, "terra_b200_reset_stats")->devs) tb_device_reset_stats(d); }
extern "C" void terra_b200_set_env_lit(HTerraScene h, int on) { for (TbDevice* d : committed(h, "terra_b200_set_env_lit")->devs) tb_device_set_env_lit(d, on); }
// Devices of a scene, before its first commit (default: TERRA_B200_DEVICES, else the terra_b200_set_device one).
extern "C" void terra_b200_scene_set_devices(HTerraScene h, const int* device_ids, int count) {
    TbScene* s = (TbScene*)h;
    if (s->dev) tb_fatal("terra_b200_scene_set_devices after the first terra_scene_commit");
    if (count < 1 || count > 16) tb_fatal("terra_b200_scene_set_devices: 1 to 16 devices");
    s->device_ids.assign(device_ids, device_ids + count);
}
extern "C" int terra_b200_scene_device_count(HTerraScene h) { return (int)((TbScene*)h)->device_ids.size(); }
extern "C" void terra_b200_set_fanout_min_paths(HTerraScene h, size_t min_paths) { ((TbScene*)h)->fanout_min_paths = min_paths; }
extern "C" void terra_b200_set_fanout_chunk_paths(HTerraScene h, size_t paths) { ((TbScene*)h)->fanout_chunk_paths = paths ? paths : 1; }
extern "C" void terra_b200_distribution_2d_sample(HTerraScene h, const float* f, size_t nx, size_t ny, size_t n,
                                                  const float* e1, const float* e2, float* out4) {
    TbScene* s = committed(h, "terra_b200_distribution_2d_sample");
    if (nx == 0 || ny == 0) tb_fatal("terra_b200_distribution_2d_sample: empty function");
    std::lock_guard<std::mutex> guard(s->lock);
    tb_device_distribution_2d_sample(s->dev, f, nx, ny, n, e1, e2, out4);
}
// Re-fit the committed tree to vertex data edited in place (SURVEY §8f N1 "build / refit"): same objects, same
// triangle counts, any new positions / normals / texcoords. Light areas are refreshed the way commit computes them
// (Terra.c:208-215). The tree keeps its topology, so quality degrades with large motion; results do not.
extern "C" float terra_b200_refit(HTerraScene h) {
    TbScene* s = committed(h, "terra_b200_refit");
    std::lock_guard<std::mutex> guard(s->lock);
    if (s->dirty_objects) tb_fatal("terra_b200_refit: objects were added or cleared since the last commit (commit instead)");
    if (s->bvh_mode == TB_BVH_REFERENCE) tb_fatal("terra_b200_refit: the reference tree's boxes carry its builder's +1e-4 merges and cannot be re-fitted (use the sah3 or lbvh tree)");
    size_t total = 0;
    std::vector<uint32_t> base(s->objects_pop);
    for (size_t i = 0; i < s->objects_pop; ++i) { base[i] = (uint32_t)total; total += s->objects[i].triangles_count; }
    static_assert(sizeof(TerraTriangle) == 36 && sizeof(TerraTriangleProperties) == 60, "record sizes the refit kernels assume");
    std::vector<const float*> tris(s->objects_pop), props(s->objects_pop);
    for (size_t i = 0; i < s->objects_pop; ++i) { tris[i] = (const float*)s->objects[i].triangles; props[i] = (const float*)s->objects[i].properties; }
    std::vector<TbDevLight> lights(s->lights.size());
    std::vector<float> light_area;
    for (size_t l = 0; l < s->lights.size(); ++l) {
        const size_t obj = s->lights[l].object;
        float area = 0;
        for (size_t j = 0; j < s->objects[obj].triangles_count; ++j) {
            const float a = triangle_area(&s->objects[obj].triangles[j]);
            s->lights[l].triangle_area[j] = a;
            area += a;
        }
        V2 mid = { 0.5f, 0.5f };
        s->lights[l].power = abi3(vmul(host_attr_eval(&s->objects[obj].material.emissive, mid), area * TB_PI));
        s->lights[l].area = area;
        lights[l].object = (int32_t)obj;
        lights[l].tri_count = (uint32_t)s->objects[obj].triangles_count;
        lights[l].tri_base = (uint32_t)light_area.size();
        lights[l].area = area;
        light_area.insert(light_area.end(), s->lights[l].triangle_area.begin(), s->lights[l].triangle_area.end());
    }
    float ms = 0.f;
    tb_device_refit(s->dev, tris.data(), props.data(), base.data(), s->objects_pop, lights.data(), lights.size(),
                    light_area.data(), light_area.size(), &ms);
    s->host_boxes_stale = true;
    return ms;
}
extern "C" void terra_b200_set_graph_paths(HTerraScene h, size_t max_paths) { for (TbDevice* d : committed(h, "terra_b200_set_graph_paths")->devs) tb_device_set_graph_paths(d, max_paths); }
extern "C" void terra_b200_set_counting(HTerraScene h, int on) { for (TbDevice* d : committed(h, "terra_b200_set_counting")->devs) tb_device_set_counting(d, on); }
extern "C" const void* terra_b200_bvh_nodes(HTerraScene h, int* nodes_count) {
    TbScene* s = (TbScene*)h;
    std::lock_guard<std::mutex> guard(s->lock);
    sync_host_nodes(s);
    *nodes_count = (int)s->nodes.size();
    return s->nodes.data();
}
extern "C" void terra_b200_commit_ms(HTerraScene h, float* out4) { memcpy(out4, ((TbScene*)h)->commit_ms, 4 * sizeof(float)); }
extern "C" void terra_b200_lbvh_build_ms(float* device_ms, float* with_upload_ms) {
    if (device_ms) *device_ms = g_lbvh_ms[0];
    if (with_upload_ms) *with_upload_ms = g_lbvh_ms[1];
}
extern "C" int terra_b200_bvh_stack_bound(HTerraScene h) { return ((TbScene*)h)->stack_bound; }
// Host-only half of terra_scene_commit (no CUDA device needed): build the BVH and flatten it.
// Lets the CPU test-suite check the builder and the flatten against the reference's node array.
extern "C" void terra_b200_set_bvh_mode(int mode) { g_bvh_mode = mode; }
extern "C" int terra_b200_build_host_bvh(HTerraScene h, int mode, float* flat_nodes_out, float* flat_tris_out) {
    TbScene* s = (TbScene*)h;
    std::lock_guard<std::mutex> guard(s->lock);
    build_bvh(s, mode);
    s->has_bvh = false;             // a later commit rebuilds for the device
    TbFlatBvh flat;
    tb_bvh_flatten(s->nodes, s->objects, (int)s->objects_pop, &flat);
    s->stack_bound = flat.stack_bound;
    if (flat_nodes_out) memcpy(flat_nodes_out, flat.nodes.data(), flat.nodes.size() * sizeof(float));
    if (flat_tris_out) memcpy(flat_tris_out, flat.tris.data(), flat.tris.size() * sizeof(float));
    return (int)s->nodes.size();
}

// Host-only: the 4-wide tree terra_scene_commit derives for k_trace_v7 (tb_bvh4_build) from the tree built with `mode`,
// with the quantised binary nodes it copies its boxes from. Returns the number of wide nodes; fills at most `cap` of them.
extern "C" int terra_b200_build_host_bvh4(HTerraScene h, int mode, uint32_t* q4nodes_out, int32_t* src_out, uint32_t* qnodes_out, int cap, int* stack_bound_out) {
    TbScene* s = (TbScene*)h;
    std::lock_guard<std::mutex> guard(s->lock);
    build_bvh(s, mode);
    s->has_bvh = false;
    TbFlatBvh flat;
    tb_bvh_flatten(s->nodes, s->objects, (int)s->objects_pop, &flat);
    TbQuantBvh quant;
    tb_bvh_quantize(flat, &quant);
    TbQuadBvh quad;
    tb_bvh4_build(flat, quant, &quad);
    const int n = (int)(quad.src.size() / 4);
    const int m = n < cap ? n : cap;
    if (q4nodes_out) memcpy(q4nodes_out, quad.nodes.data(), (size_t)m * 64);
    if (src_out) memcpy(src_out, quad.src.data(), (size_t)m * 16);
    if (qnodes_out) memcpy(qnodes_out, quant.nodes.data(), quant.nodes.size() * 4);
    if (stack_bound_out) *stack_bound_out = quad.stack_bound;
    return n;
}

extern "C" void* terra_b200_host_alloc(size_t bytes) { return tb_host_alloc_pinned(bytes ? bytes : 1); }
extern "C" void terra_b200_host_free(void* p) { if (p) tb_host_free_pinned(p); }
extern "C" void terra_b200_trace_rays(HTerraScene h, size_t n, const float* origins, const float* directions, TerraB200Hit* hits_out) {
    TbScene* s = committed(h, "terra_b200_trace_rays");
    std::lock_guard<std::mutex> guard(s->lock);
    tb_device_trace_host(s->dev, n, origins, directions, hits_out);
}
extern "C" void terra_b200_trace_rays_device(HTerraScene h, size_t n, const void* d_o, const void* d_d, void* d_hits, void* stream) {
    TbScene* s = committed(h, "terra_b200_trace_rays_device");
    std::lock_guard<std::mutex> guard(s->lock);
    tb_device_trace_resident(s->dev, n, d_o, d_d, d_hits, stream);
}
extern "C" int terra_b200_trace_rays_packed_noalloc(HTerraScene h, size_t n, const float* d_o3, const float* d_d3, void* d_hits, void* stream) {
    TbScene* s = committed(h, "terra_b200_trace_rays_packed_noalloc");
    std::lock_guard<std::mutex> guard(s->lock);
    return tb_device_trace_packed_noalloc(s->dev, n, d_o3, d_d3, d_hits, stream);
}
extern "C" void terra_b200_primary_hits(const TerraCamera* camera, HTerraScene h, size_t width, size_t height, TerraB200Hit* hits_out) {
    TbScene* s = committed(h, "terra_b200_primary_hits");
    std::lock_guard<std::mutex> guard(s->lock);
    TbCameraFrame cam = camera_frame(camera);
    tb_device_primary_hits(s->dev, &cam, width, height, hits_out);
}
extern "C" void terra_b200_render_device(const TerraCamera* camera, HTerraScene h, size_t fbw, size_t fbh,
                                         size_t x, size_t y, size_t width, size_t height,
                                         size_t sample_base, size_t spp, void* d_results, void* stream) {
    TbScene* s = committed(h, "terra_b200_render_device");
    std::lock_guard<std::mutex> guard(s->lock);
    TbCameraFrame cam = camera_frame(camera);
    uint64_t seed = s->seed_fixed ? s->seed * 0x9E3779B97F4A7C15ull + 1 : next_seed(s);
    tb_device_render_resident(s->dev, &cam, fbw, fbh, x, y, width, height, sample_base, spp, seed, d_results, stream);
}
extern "C" void terra_b200_resolve_device(HTerraScene h, size_t fbw, size_t fbh, const void* d_results, void* d_pixels, void* stream) {
    TbScene* s = committed(h, "terra_b200_resolve_device");
    tb_device_resolve(s->dev, fbw, fbh, d_results, d_pixels, stream);
}
extern "C" void terra_b200_bsdf_kat(HTerraScene h, size_t object_idx, size_t n, const float* e123, const float* wo,
                                    const float* normal, float* out) {
    TbScene* s = committed(h, "terra_b200_bsdf_kat");
    if (object_idx >= s->objects_pop) tb_fatal("terra_b200_bsdf_kat: object index out of range");
    tb_device_bsdf_kat(s->dev, object_idx, n, e123, wo, normal, out);
}
