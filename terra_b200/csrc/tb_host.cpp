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
#include <mutex>
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
extern "C" bool terra_framebuffer_create(TerraFramebuffer* fb, size_t width, size_t height) {     // :309-326
    if (width == 0 || height == 0) return false;
    fb->width = width;
    fb->height = height;
    fb->pixels = (TerraFloat3*)terra_malloc(sizeof(TerraFloat3) * width * height);
    fb->results = (TerraRawIntegrationResult*)terra_malloc(sizeof(TerraRawIntegrationResult) * width * height);
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
    terra_free(fb->results);
    terra_free(fb->pixels);
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
    TbDevice*         dev;
    uint64_t          seed;
    bool              seed_fixed;
    uint64_t          render_calls;
    std::mutex        lock;
};

static int g_device_index = 0;
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
    s->dev = NULL;
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
    if (s->dev) tb_device_destroy(s->dev);
    delete s;
}
