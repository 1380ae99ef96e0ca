// tb_render.cu — the CUDA layer of libterra_b200.so (sm_100a): wavefront path tracer that replaces
// the per-pixel loop of the reference's terra_render (src/Terra.c:512-635).
//
// Kernels (SURVEY §2.2 names in brackets):
//   k_raygen      [K1]  camera rays            Terra.c:553-560, 1783-1799, 1702-1710
//   k_trace       [K2]  persistent closest hit Terra.c:1623-1657, TerraBVH.c:250-310,
//                        + material binning     Terra.c:851-878, TerraGeometry.c:98-259
//   k_shade       [K3+K4] surface, integrator, BSDF sample/pdf/eval, Russian roulette, next ray
//                                               Terra.c:1726-1764, 1039-1097, 1339-1587; TerraPresets.c
//   k_accumulate / k_finalize [K6]              Terra.c:569-630, 1815-1828
// Compiled with -fmad=false: the reference's x86-64 build has no FMA contraction and the primary
// closest-hit parity is bit-exact only if every mul/add rounds separately.
// No tensor cores: there is no dense contraction anywhere on this path.
#include <cuda_runtime.h>

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "tb_internal.h"
#include "tb_bsdf.h"
#include "tb_texture.h"

extern "C" void terra_log(const char* fmt, ...);

#define TB_CUDA(call)                                                                          \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            terra_log("terra_b200: CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_),        \
                      __FILE__, __LINE__, cudaGetErrorString(e_));                             \
            abort();                                                                           \
        }                                                                                      \
    } while (0)

// ------------------------------------------------------------------------------------ tunables
#define TB_TRACE_THREADS      128          // threads per trace CTA
#define TB_TRACE_CTAS_PER_SM  8            // persistent grid = SMs x this
#define TB_STACK_SMEM         16           // traversal stack entries held in shared memory ...
#define TB_STACK_LOCAL        112          // ... deeper entries spill to local memory
#define TB_REFETCH_BELOW      20           // dynamic fetch: refill the warp when fewer lanes are busy
#define TB_SHADE_THREADS      256
#define TB_QUEUE_SENTINEL     0x7fffffff

// ------------------------------------------------------------------------------------ RNG
// Counter-based: splitmix64 over (path key, dimension). The reference draws BSDF/RR numbers from
// libc rand() and pixel jitter from a time-seeded PCG32, so CPU<->GPU parity beyond
// the first hit is statistical by construction; what IS reproduced is the float mapping:
// (float)rand() / RAND_MAX can return exactly 1.0, and PCG's u32 * 2^-32 can round to 1.0.
__host__ __device__ __forceinline__ uint64_t tb_mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ uint64_t tb_path_key(uint64_t seed, uint32_t pixel, uint32_t sample) {
    return tb_mix64(seed ^ (((uint64_t)pixel << 32) | sample) * 0x9E3779B97F4A7C15ull);
}
__device__ __forceinline__ uint32_t tb_rand_u32(uint64_t key, uint32_t dim) {
    return (uint32_t)(tb_mix64(key + (uint64_t)(dim + 1) * 0x9E3779B97F4A7C15ull) >> 32);
}
__device__ __forceinline__ float tb_randf_libc(uint64_t key, uint32_t dim) {     // ~ (float)rand() / RAND_MAX, Terra.c:115
    return __uint2float_rn(tb_rand_u32(key, dim) >> 1) * 4.6566128730773926e-10f;    // / 2^31 == / (float)RAND_MAX
}
__device__ __forceinline__ float tb_randf_pcg(uint64_t key, uint32_t dim) {      // ~ rndi * 2^-32, Terra.c:698-700
    return __uint2float_rn(tb_rand_u32(key, dim)) * 2.3283064365386963e-10f;
}
enum {  // dimensions per bounce (16 reserved)
    TB_DIM_JITTER_X = 0, TB_DIM_JITTER_Y = 1,
    TB_DIM_BSDF0 = 0, TB_DIM_BSDF1 = 1, TB_DIM_BSDF2 = 2, TB_DIM_RR = 3,
    TB_DIM_LIGHT_PICK = 4, TB_DIM_TRI_PICK = 5, TB_DIM_LIGHT_U = 6, TB_DIM_LIGHT_V = 7,
    TB_DIM_MIS0 = 8, TB_DIM_MIS1 = 9, TB_DIM_MIS2 = 10,
    TB_DIMS_PER_BOUNCE = 16
};

// ------------------------------------------------------------------------------------ device state
struct TbWave {                // wavefront buffers, `capacity` path slots
    size_t   capacity;
    float4*  ray_o;            // xyz origin (before the 0.001*dir raycast offset)
    float4*  ray_d;            // xyz direction
    float4*  thr;              // throughput
    float4*  lo;               // radiance gathered so far by this path
    float4*  hit;              // t, rank (int bits), u, v
    float4*  aux_o;            // light / MIS visibility rays (2 per slot)
    float4*  aux_d;
    float4*  aux_hit;
    int*     queue[2];         // live path slots, ping-pong
    int*     matq;             // TB_PRESET_COUNT x capacity: slots binned by preset after the trace
    int*     auxq;             // aux ray ids to trace
    int*     counters;         // [0] fetch cursor, [1] queue count cur, [2] next, [3] aux count, [4..7] matq counts
};

struct TbEventPair { cudaEvent_t a, b; };

struct TbDevice {
    int          index;
    int          sm_count;
    cudaStream_t stream;
    TbDevScene   sc;
    // owned device memory
    float4 *d_nodes, *d_tris, *d_props;
    TbDevMaterial* d_mats;
    TbDevTexture*  d_texs;
    TbDevLight*    d_lights;
    float*         d_light_area;
    uint32_t*      d_light_rank;
    std::vector<void*> tex_pixels;
    int          stack_bound;
    TbWave       wave;
    // per-call staging for terra_render
    size_t       fb_capacity;
    TerraRawIntegrationResult* d_results;
    float*       d_pixels;
    float4*      d_acc;
    size_t       acc_capacity;
    // scratch for the host-buffer trace entry points
    size_t       io_capacity;
    float4      *d_io_o, *d_io_d;
    TerraB200Hit* d_io_hit;
    // stats
    TerraB200Stats stats;
    std::vector<TbEventPair> ev_pool;
    size_t       ev_used;
    cudaEvent_t  ev_call_a, ev_call_b;
};

static void tb_free(void* p) { if (p) TB_CUDA(cudaFree(p)); }
template <typename T> static T* tb_alloc(size_t n) {
    void* p = NULL;
    TB_CUDA(cudaMalloc(&p, (n ? n : 1) * sizeof(T)));
    return (T*)p;
}

extern "C" int tb_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

static void tb_wave_free(TbWave* w) {
    tb_free(w->ray_o); tb_free(w->ray_d); tb_free(w->thr); tb_free(w->lo); tb_free(w->hit);
    tb_free(w->aux_o); tb_free(w->aux_d); tb_free(w->aux_hit);
    tb_free(w->queue[0]); tb_free(w->queue[1]); tb_free(w->matq); tb_free(w->auxq); tb_free(w->counters);
    memset(w, 0, sizeof *w);
}
static void tb_wave_reserve(TbWave* w, size_t capacity) {
    if (w->capacity >= capacity) return;
    tb_wave_free(w);
    w->capacity = capacity;
    w->ray_o = tb_alloc<float4>(capacity);
    w->ray_d = tb_alloc<float4>(capacity);
    w->thr = tb_alloc<float4>(capacity);
    w->lo = tb_alloc<float4>(capacity);
    w->hit = tb_alloc<float4>(capacity);
    w->aux_o = tb_alloc<float4>(capacity * 2);
    w->aux_d = tb_alloc<float4>(capacity * 2);
    w->aux_hit = tb_alloc<float4>(capacity * 2);
    w->queue[0] = tb_alloc<int>(capacity);
    w->queue[1] = tb_alloc<int>(capacity);
    w->matq = tb_alloc<int>(capacity * TB_PRESET_COUNT);
    w->auxq = tb_alloc<int>(capacity * 2);
    w->counters = tb_alloc<int>(16);
}

extern "C" TbDevice* tb_device_create(int device_index) {
    int n = tb_cuda_device_count();
    if (n <= 0) {
        terra_log("terra_b200: no CUDA device visible. This library has no CPU fallback: "
                  "terra_scene_commit / terra_render need a B200 (sm_100a).\n");
        abort();
    }
    if (device_index < 0 || device_index >= n) { terra_log("terra_b200: device %d out of range (%d visible)\n", device_index, n); abort(); }
    TbDevice* d = new TbDevice();
    d->index = device_index;
    TB_CUDA(cudaSetDevice(device_index));
    cudaDeviceProp prop;
    TB_CUDA(cudaGetDeviceProperties(&prop, device_index));
    d->sm_count = prop.multiProcessorCount;
    TB_CUDA(cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking));
    memset(&d->sc, 0, sizeof d->sc);
    d->d_nodes = d->d_tris = d->d_props = NULL;
    d->d_mats = NULL; d->d_texs = NULL; d->d_lights = NULL; d->d_light_area = NULL; d->d_light_rank = NULL;
    d->stack_bound = 0;
    memset(&d->wave, 0, sizeof d->wave);
    d->fb_capacity = 0; d->d_results = NULL; d->d_pixels = NULL; d->d_acc = NULL; d->acc_capacity = 0;
    d->io_capacity = 0; d->d_io_o = d->d_io_d = NULL; d->d_io_hit = NULL;
    memset(&d->stats, 0, sizeof d->stats);
    d->ev_used = 0;
    TB_CUDA(cudaEventCreate(&d->ev_call_a));
    TB_CUDA(cudaEventCreate(&d->ev_call_b));
    return d;
}

static void tb_scene_free(TbDevice* d) {
    tb_free(d->d_nodes); tb_free(d->d_tris); tb_free(d->d_props); tb_free(d->d_mats); tb_free(d->d_texs);
    tb_free(d->d_lights); tb_free(d->d_light_area); tb_free(d->d_light_rank);
    for (void* p : d->tex_pixels) tb_free(p);
    d->tex_pixels.clear();
    d->d_nodes = d->d_tris = d->d_props = NULL;
    d->d_mats = NULL; d->d_texs = NULL; d->d_lights = NULL; d->d_light_area = NULL; d->d_light_rank = NULL;
}

extern "C" void tb_device_destroy(TbDevice* d) {
    TB_CUDA(cudaSetDevice(d->index));
    TB_CUDA(cudaStreamSynchronize(d->stream));
    tb_scene_free(d);
    tb_wave_free(&d->wave);
    tb_free(d->d_results); tb_free(d->d_pixels); tb_free(d->d_acc);
    tb_free(d->d_io_o); tb_free(d->d_io_d); tb_free(d->d_io_hit);
    for (TbEventPair& e : d->ev_pool) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    cudaEventDestroy(d->ev_call_a);
    cudaEventDestroy(d->ev_call_b);
    cudaStreamDestroy(d->stream);
    delete d;
}

template <typename T> static T* tb_upload(const T* host, size_t n, TbDevice* d) {
    T* p = tb_alloc<T>(n);
    if (n) TB_CUDA(cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice));
    d->stats.h2d_bytes += n * sizeof(T);
    return p;
}

extern "C" void tb_device_upload(TbDevice* d, const TbFlatScene* f) {
    TB_CUDA(cudaSetDevice(d->index));
    TB_CUDA(cudaStreamSynchronize(d->stream));
    if (f->stack_bound > TB_STACK_SMEM + TB_STACK_LOCAL) {
        terra_log("terra_b200: BVH depth %d exceeds the traversal stack (%d entries)\n", f->stack_bound, TB_STACK_SMEM + TB_STACK_LOCAL);
        abort();
    }
    tb_scene_free(d);
    d->d_nodes = (float4*)tb_upload(f->nodes, f->n_nodes * 16, d);
    d->d_tris = (float4*)tb_upload(f->tris, f->n_tris * 12, d);
    d->d_props = (float4*)tb_upload(f->props, f->n_tris * 16, d);
    std::vector<TbDevMaterial> mats(f->materials, f->materials + f->n_objects);
    d->d_mats = tb_upload(mats.data(), mats.size(), d);
    std::vector<TbDevTexture> texs(f->n_textures);
    for (size_t i = 0; i < f->n_textures; ++i) {
        const TerraTexture* t = f->textures[i];
        size_t bytes = (size_t)t->width * t->height * t->components * t->depth;
        void* px = tb_upload((const uint8_t*)t->pixels, bytes, d);
        d->tex_pixels.push_back(px);
        texs[i].pixels = px;
        texs[i].width = t->width; texs[i].height = t->height;
        texs[i].components = t->components; texs[i].depth = t->depth;
        texs[i].filter = t->filter; texs[i].address_mode = t->address_mode;
        texs[i].pad_ = 0;
    }
    d->d_texs = tb_upload(texs.data(), texs.size(), d);
    d->d_lights = tb_upload(f->lights, f->n_lights, d);
    d->d_light_area = tb_upload(f->light_tri_area, f->n_light_tris, d);
    d->d_light_rank = tb_upload(f->light_tri_rank, f->n_light_tris, d);
    d->stack_bound = f->stack_bound;
    TbDevScene& sc = d->sc;
    sc.nodes = d->d_nodes; sc.tris = d->d_tris; sc.props = d->d_props;
    sc.materials = d->d_mats; sc.textures = d->d_texs; sc.lights = d->d_lights;
    sc.light_tri_area = d->d_light_area; sc.light_tri_rank = d->d_light_rank;
    sc.n_nodes = (int)f->n_nodes; sc.n_tris = (int)f->n_tris;
    sc.n_objects = (int)f->n_objects; sc.n_lights = (int)f->n_lights;
    sc.lights_triangles_count = f->lights_triangles_count;
    sc.env = f->env;
    sc.preset_mask = 0;
    for (size_t i = 0; i < f->n_objects; ++i) sc.preset_mask |= 1 << f->materials[i].preset;
}

extern "C" void tb_device_set_options(TbDevice* d, const TerraSceneOptions* o, const TbDevAttr* env) {
    TbDevScene& sc = d->sc;
    sc.env = *env;
    sc.integrator = (int)o->integrator;
    sc.bounces = (int)o->bounces;
    sc.tonemap = (int)o->tonemapping_operator;
    sc.jitter = o->subpixel_jitter;
    sc.exposure = o->manual_exposure;
    sc.gamma = o->gamma;
    sc.inv_gamma = 1.f / o->gamma;                      // `1.f / scene->opts.gamma`, Terra.c:582
}
