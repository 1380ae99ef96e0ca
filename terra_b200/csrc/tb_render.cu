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
//   k_miss        env-lit extension             Terra.c:1056 restored; tb_distribution.cuh: Terra.c:760-846 on the device
// Host side of this file: wave loop (tb_render_region), CUDA-graph replay of small calls (tb_graph_for),
// pipelined host ray batches (tb_device_trace_host), uploads, statistics.
// Compiled with -fmad=false: the reference's x86-64 build has no FMA contraction and the primary
// closest-hit parity is bit-exact only if every mul/add rounds separately.
// No tensor cores: there is no dense contraction anywhere on this path.
#include <cuda_runtime.h>

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <thread>
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
#ifndef TB_TRACE_CTAS_PER_SM
#define TB_TRACE_CTAS_PER_SM  8            // persistent grid = SMs x this
#endif
#define TB_STACK_SMEM         16           // traversal stack entries held in shared memory ...
#define TB_STACK_LOCAL        112          // ... deeper entries spill to local memory
#define TB_TRI_MIN            6            // rounds kernels: parked leaves needed before a triangle round (grid of profiles/r3c_grid.jsonl)
#define TB_REFETCH_ROUNDS     24           // rounds kernels: refill the warp when fewer lanes own a ray (grid of profiles/r3c_grid.jsonl: 20 / 22 / 24 / 26)
#define TB_REFETCH_BELOW      20           // dynamic fetch: refill the warp when fewer lanes are busy
#define TB_SHADE_THREADS      256
#ifndef TB_SHADE_MIN_BLOCKS
#define TB_SHADE_MIN_BLOCKS    1            // resident CTAs per SM the direct / debug shade kernels are compiled for (register cap)
#endif
#ifndef TB_SHADE_MIN_BLOCKS_SIMPLE
#define TB_SHADE_MIN_BLOCKS_SIMPLE 3        // ... and the simple-integrator ones (<= 85 registers)
#endif
#define TB_QUEUE_SENTINEL     0x7fffffff
#define TB_K_V6_ID            2            // = TB_K_V6 / TB_K_PACKET below (the device is created before the enum is seen)
#define TB_K_PACKET_ID        3
#define TB_K_V7_ID            4
#define TB_V7_MIN_MB          400          // k_trace_v7 by default for scenes whose quantised nodes + triangles exceed this (L2: 126 MB; C3 at
                                           // 3M triangles = 240 MB: v8 4931 / v7 4689 Mrays/s, at 10M = 800 MB: 3324 / 4290, profiles/r3f, r3e)
#define TB_K_V8_ID            5
#define TB_V8_SMEM_DEPTH      40           // k_trace_v8: rows of the shared stack, 4 bytes per entry (deeper trees: + local memory)
#define TB_LEAN_WORDS         10           // k_trace_v7 / v8: words of per-ray state kept in shared memory behind the stack rows
#define TB_V7_SMEM_DEPTH      16           // k_trace_v7: rows of the shared stack ...
#define TB_V7_LOCAL           48           // ... and entries beyond them in local memory; deeper 4-wide trees are not uploaded (the scene runs k_trace_v6)

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
    TB_DIM_ENV0 = 11, TB_DIM_ENV1 = 12,        // explicit environment sample (env_lit mode 2)
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
    float4*  aux_o;            // light / MIS visibility rays (2 per slot: [2*slot + k]) + explicit environment rays ([2*capacity + slot])
    float4*  aux_d;
    float4*  aux_hit;
    int*     queue[2];         // live path slots, ping-pong
    int*     matq;             // TB_PRESET_COUNT x capacity: slots binned by preset after the trace
    int*     auxq;             // aux ray ids to trace
    int*     counters;         // [0] fetch cursor, [1] queue count cur, [2] next, [3] aux count, [4..7] matq counts
};

// What changes between two terra_render calls of the same shape. A call replayed from a CUDA graph reads
// these from device memory (k_raygen is the only kernel that needs them) so that the graph's kernel arguments
// stay constant.
struct TbCallParams {
    TbCameraFrame cam;
    int      x, y, fbw, fbh;
    uint32_t sample_base, pad_;
    uint64_t seed;
    void*    results;           // device-resident calls: the caller's full-framebuffer results array (k_finalize)
};

struct TbTextureSig { const void* host; const void* pixels; uint32_t width, height, components, depth, filter, address_mode; };
struct TbEventPair { cudaEvent_t a, b; int coherent; };

struct TbDevice {
    int          index;
    int          sm_count;
    cudaStream_t stream;
    TbDevScene   sc;
    // owned device memory
    float4 *d_nodes, *d_tris, *d_props;
    uint4*  d_qnodes;
    uint4*  d_q4nodes;         // 4-wide tree of k_trace_v7 (4 x uint4 per node) and, for refits, where each slot's box comes from
    int*    d_q4src;
    size_t  n_q4;
    int     stack_bound4;
    float   grid_lo[3], grid_cell[3];
    TbDevMaterial* d_mats;
    TbDevTexture*  d_texs;
    TbDevLight*    d_lights;
    float*         d_light_area;
    uint32_t*      d_light_rank;
    std::vector<void*> tex_pixels;
    std::vector<TbDevTexture> h_texs;    // host copy of the texture table (dims for the environment importance map)
    std::vector<TbTextureSig> tex_sig;   // what the device copies of the textures were made from
    float       *d_env_f, *d_env_cdf, *d_env_rows, *d_env_marg;   // TbDevDist2D storage of sc.env_dist
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
    // terra_b200_trace_rays pipeline: copy-in / compute / copy-out streams, two slots of raw xyz input and hits
    cudaStream_t st_in, st_out;
    size_t       pipe_capacity;
    float       *d_pipe_o3[2], *d_pipe_d3[2];
    TerraB200Hit* d_pipe_hit[2];
    cudaEvent_t  ev_in[2], ev_pack[2], ev_trace[2], ev_out[2];
    // pageable caller buffers: page-locked bounce buffers, filled / drained by host threads while the GPU works
    size_t       stage_capacity;
    float       *h_stage_o3[2], *h_stage_d3[2];
    TerraB200Hit* h_stage_hit[2];
    // stats
    TerraB200Stats stats;
    std::vector<TbEventPair> ev_pool;
    size_t       ev_used;
    cudaEvent_t  ev_call_a, ev_call_b, ev_legacy;
    cudaEvent_t  ev_partial, ev_band;      // multi-device terra_render: this device's accumulator is complete / its band is finalised
    cudaEvent_t  ev_slice_in;              // ... its slice of the caller's results[] has arrived
    size_t       stage_off;                // ... pixels of d_results / d_pixels the slices of this call have taken so far
    int          counting;     // use the counting variant of k_trace (work counters for the roofline)
    int          exact_boxes;  // reference tree: reference box arithmetic
    int          trace_kind;   // TB_K_* forced by TERRA_B200_TRACE, or TB_K_AUTO
    int          fused_raygen;      // 1: camera rays are generated inside the packet kernel (no k_raygen, no ray arrays for bounce 0)
    int          coherent_kind;     // TB_K_* used for camera-ray launches when trace_kind is auto
    int          incoherent_kind;   // TB_K_* used for incoherent launches when trace_kind is auto
    size_t       v7_min_bytes;      // auto: k_trace_v7 when quantised nodes + triangle records are at least this large
    int          warp_samples; // at most this many samples of one pixel share a warp (TERRA_B200_WARP_SAMPLES, default 32)
    int          tile_pixels;  // 1: waves walk the region in 8x4 pixel tiles (TERRA_B200_TILE=0 turns it off)
    int          tri_min, refetch_below;
    // CUDA-graph replay of small terra_render calls (launch-bound: ~45 launches for a 128x128 tile)
    struct Graph { size_t w, h, spp; int kind; uint64_t epoch; cudaGraphExec_t exec; uint64_t launches, paths, last_use; };
    std::vector<Graph> graphs;
    std::vector<uint64_t> graph_seen;      // shapes rendered once without a graph (w | h << 20 | spp << 40, xor epoch)
    uint64_t     graph_epoch, graph_clock;  // epoch: bumped whenever a pointer or scene constant a graph baked in changes
    size_t       graph_max_paths;           // TERRA_B200_GRAPH_PATHS (0 = never use graphs)
    bool         capturing;
    int          env_mode;                  // what terra_b200_set_env_lit asked for (0 / 1 / 2), re-applied at every commit
    TbCallParams *d_call, *h_call;
    unsigned char* noalloc_scratch;         // device address of g_noalloc_scratch ON THIS DEVICE (cudaGetSymbolAddress in tb_device_create)
};

static void tb_free(void* p) { if (p) TB_CUDA(cudaFree(p)); }
template <typename T> static T* tb_alloc(size_t n) {
    void* p = NULL;
    TB_CUDA(cudaMalloc(&p, (n ? n : 1) * sizeof(T)));
    return (T*)p;
}

// Page-locked host memory for framebuffers the library allocates itself (terra_framebuffer_create): the
// H2D of results[] and the D2H of results[] + pixels[] inside terra_render then run at PCIe speed instead
// of being staged through the driver's bounce buffers. NULL when there is no device or no memory to pin.
extern "C" void* tb_host_alloc_pinned(size_t bytes) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { cudaGetLastError(); return NULL; }
    void* p = NULL;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return NULL; }
    return p;
}
extern "C" void tb_host_free_pinned(void* p) { cudaFreeHost(p); }

extern "C" int tb_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

static void tb_wave_free(TbWave* w) {
    tb_free(w->ray_o); tb_free(w->ray_d); tb_free(w->thr); tb_free(w->lo); tb_free(w->hit);
    tb_free(w->aux_o); tb_free(w->aux_d); tb_free(w->aux_hit);
    tb_free(w->queue[0]); tb_free(w->queue[1]); tb_free(w->matq); tb_free(w->auxq);
    int* counters = w->counters;                 // lives as long as the device (holds the ray statistics)
    memset(w, 0, sizeof *w);
    w->counters = counters;
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
    w->aux_o = tb_alloc<float4>(capacity * 3);
    w->aux_d = tb_alloc<float4>(capacity * 3);
    w->aux_hit = tb_alloc<float4>(capacity * 3);
    w->queue[0] = tb_alloc<int>(capacity);
    w->queue[1] = tb_alloc<int>(capacity);
    w->matq = tb_alloc<int>(capacity * TB_PRESET_COUNT);
    w->auxq = tb_alloc<int>(capacity * 3);
}

// Scratch of terra_b200_trace_rays_packed_noalloc: a fetch cursor (first 256 bytes) and the float4 origins / directions of up
// to TB_NOALLOC_MAX_RAYS rays. File-scope device memory: every GPU that loads the library holds these 8 MB until the
// process exits. ONE array per device shared by every call: it keeps the last call's contents (the cursor is cleared by
// every call, the ray slots are overwritten), and two calls that overlap on one GPU race.
#define TB_NOALLOC_MAX_RAYS 262144          /* = TERRA_B200_NOALLOC_MAX_RAYS, include/terra_b200.h */
#define TB_NOALLOC_HEADER   256
static __device__ __align__(256) unsigned char g_noalloc_scratch[TB_NOALLOC_HEADER + (size_t)TB_NOALLOC_MAX_RAYS * 32];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes
// from cudaGetSymbolAddress, per device, with that device current.
static unsigned char* tb_noalloc_scratch_address() {
    unsigned char* p = NULL;
    TB_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    return p;
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
    d->noalloc_scratch = tb_noalloc_scratch_address();      // this device is current: the symbol resolves to ITS copy of the array
    memset(&d->sc, 0, sizeof d->sc);
    d->d_nodes = d->d_tris = d->d_props = NULL;
    d->d_qnodes = NULL;
    d->d_q4nodes = NULL; d->d_q4src = NULL; d->n_q4 = 0; d->stack_bound4 = 0;
    d->d_mats = NULL; d->d_texs = NULL; d->d_lights = NULL; d->d_light_area = NULL; d->d_light_rank = NULL;
    d->stack_bound = 0;
    memset(&d->wave, 0, sizeof d->wave);
    d->wave.counters = tb_alloc<int>(64);
    TB_CUDA(cudaMemset(d->wave.counters, 0, 64 * sizeof(int)));
    d->fb_capacity = 0; d->d_results = NULL; d->d_pixels = NULL; d->d_acc = NULL; d->acc_capacity = 0;
    d->io_capacity = 0; d->d_io_o = d->d_io_d = NULL; d->d_io_hit = NULL;
    d->st_in = d->st_out = NULL; d->pipe_capacity = 0;
    d->d_env_f = d->d_env_cdf = d->d_env_rows = d->d_env_marg = NULL;
    d->graph_epoch = 1; d->graph_clock = 0; d->capturing = false; d->env_mode = 0;
    d->d_call = tb_alloc<TbCallParams>(1);
    TB_CUDA(cudaHostAlloc((void**)&d->h_call, 8 * sizeof(TbCallParams), cudaHostAllocDefault));      // TB_CALL_SLOTS
    for (int k = 0; k < 2; ++k) { d->d_pipe_o3[k] = d->d_pipe_d3[k] = NULL; d->d_pipe_hit[k] = NULL; }
    d->stage_capacity = 0;
    for (int k = 0; k < 2; ++k) { d->h_stage_o3[k] = d->h_stage_d3[k] = NULL; d->h_stage_hit[k] = NULL; }
    memset(&d->stats, 0, sizeof d->stats);
    d->ev_used = 0;
    d->counting = 0;
    d->exact_boxes = 0;
    {
        // A/B switches: TERRA_B200_TRACE forces one kernel for every launch; TERRA_B200_INCOHERENT / TERRA_B200_COHERENT
        // choose it for bounce / visibility / API rays and for camera rays. walk = per-ray while-while (the work
        // counters' kernel), v6 = rounds kernel (tb_trace_v6.cuh), packet = one warp per 32-ray packet.
        // v7 = the rounds kernel on the 4-wide tree (tb_trace_v7.cuh); a scene without that tree (reference mode) runs v6.
        // v8 = the rounds kernel on the binary tree with 4-byte stack entries and the triangle-round state in shared memory
        // (tb_trace_v8.cuh): the default for scenes that fit L2.
        static const char* names[] = { "auto", "walk", "v6", "packet", "v7", "v8" };
        auto kind_of = [&](const char* env, int dflt) {
            const char* e = getenv(env);
            for (int k = 0; e && k < 6; ++k) if (strcmp(e, names[k]) == 0) return k;
            return dflt;
        };
        d->trace_kind = kind_of("TERRA_B200_TRACE", 0);
        d->incoherent_kind = kind_of("TERRA_B200_INCOHERENT", 0);
        const char* wm = getenv("TERRA_B200_V7_MIN_MB"); d->v7_min_bytes = (size_t)(wm ? atof(wm) : TB_V7_MIN_MB) << 20;
        d->coherent_kind = kind_of("TERRA_B200_COHERENT", TB_K_PACKET_ID);
        const char* t = getenv("TERRA_B200_TRI_MIN"); d->tri_min = t ? atoi(t) : TB_TRI_MIN;
        const char* tp = getenv("TERRA_B200_TILE"); d->tile_pixels = tp ? atoi(tp) : 1;
        const char* ws = getenv("TERRA_B200_WARP_SAMPLES"); d->warp_samples = ws ? atoi(ws) : 32;
        const char* gp = getenv("TERRA_B200_GRAPH_PATHS"); d->graph_max_paths = gp ? (size_t)strtoull(gp, NULL, 10) : ((size_t)1 << 22);
        const char* f = getenv("TERRA_B200_REFETCH"); d->refetch_below = f ? atoi(f) : TB_REFETCH_ROUNDS;
        const char* fr = getenv("TERRA_B200_FUSED_RAYGEN"); d->fused_raygen = fr ? atoi(fr) : 1;
    }
    TB_CUDA(cudaEventCreate(&d->ev_call_a));
    TB_CUDA(cudaEventCreate(&d->ev_call_b));
    TB_CUDA(cudaEventCreateWithFlags(&d->ev_legacy, cudaEventDisableTiming));
    TB_CUDA(cudaEventCreateWithFlags(&d->ev_partial, cudaEventDisableTiming));
    TB_CUDA(cudaEventCreateWithFlags(&d->ev_band, cudaEventDisableTiming));
    TB_CUDA(cudaEventCreateWithFlags(&d->ev_slice_in, cudaEventDisableTiming));
    d->stage_off = 0;
    return d;
}

static void tb_scene_free(TbDevice* d) {
    tb_free(d->d_nodes); tb_free(d->d_tris); tb_free(d->d_props); tb_free(d->d_mats); tb_free(d->d_texs);
    tb_free(d->d_qnodes); d->d_qnodes = NULL;
    tb_free(d->d_q4nodes); tb_free(d->d_q4src); d->d_q4nodes = NULL; d->d_q4src = NULL; d->n_q4 = 0;
    tb_free(d->d_lights); tb_free(d->d_light_area); tb_free(d->d_light_rank);
    for (void* p : d->tex_pixels) tb_free(p);
    d->tex_pixels.clear();
    d->tex_sig.clear();
    tb_free(d->d_env_f); tb_free(d->d_env_cdf); tb_free(d->d_env_rows); tb_free(d->d_env_marg);
    d->d_env_f = d->d_env_cdf = d->d_env_rows = d->d_env_marg = NULL;
    d->sc.env_nee = 0;
    d->d_nodes = d->d_tris = d->d_props = NULL;
    d->d_mats = NULL; d->d_texs = NULL; d->d_lights = NULL; d->d_light_area = NULL; d->d_light_rank = NULL;
}

extern "C" void tb_device_destroy(TbDevice* d) {
    TB_CUDA(cudaSetDevice(d->index));
    TB_CUDA(cudaStreamSynchronize(d->stream));
    tb_scene_free(d);
    tb_wave_free(&d->wave);
    tb_free(d->wave.counters);
    tb_free(d->d_results); tb_free(d->d_pixels); tb_free(d->d_acc);
    tb_free(d->d_io_o); tb_free(d->d_io_d); tb_free(d->d_io_hit);
    for (int k = 0; k < 2; ++k) { tb_free(d->d_pipe_o3[k]); tb_free(d->d_pipe_d3[k]); tb_free(d->d_pipe_hit[k]); }
    for (int k = 0; k < 2; ++k) { if (d->h_stage_o3[k]) cudaFreeHost(d->h_stage_o3[k]); if (d->h_stage_d3[k]) cudaFreeHost(d->h_stage_d3[k]); if (d->h_stage_hit[k]) cudaFreeHost(d->h_stage_hit[k]); }
    if (d->st_in) {
        for (int k = 0; k < 2; ++k) { cudaEventDestroy(d->ev_in[k]); cudaEventDestroy(d->ev_pack[k]); cudaEventDestroy(d->ev_trace[k]); cudaEventDestroy(d->ev_out[k]); }
        cudaStreamDestroy(d->st_in); cudaStreamDestroy(d->st_out);
    }
    for (TbDevice::Graph& g : d->graphs) cudaGraphExecDestroy(g.exec);
    tb_free(d->d_call); cudaFreeHost(d->h_call);
    for (TbEventPair& e : d->ev_pool) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    cudaEventDestroy(d->ev_call_a);
    cudaEventDestroy(d->ev_call_b);
    cudaEventDestroy(d->ev_legacy);
    cudaEventDestroy(d->ev_partial);
    cudaEventDestroy(d->ev_band);
    cudaEventDestroy(d->ev_slice_in);
    cudaStreamDestroy(d->stream);
    delete d;
}

template <typename T> static T* tb_upload(const T* host, size_t n, TbDevice* d) {
    T* p = tb_alloc<T>(n);
    if (n) TB_CUDA(cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice));
    d->stats.h2d_bytes += n * sizeof(T);
    return p;
}

// Geometry half of a commit: nodes, rank-ordered triangle / property records, quantised nodes.
extern "C" void tb_device_upload_geometry(TbDevice* d, const TbFlatScene* f) {
    TB_CUDA(cudaSetDevice(d->index));
    TB_CUDA(cudaStreamSynchronize(d->stream));
    if (f->stack_bound > TB_STACK_SMEM + TB_STACK_LOCAL) {
        terra_log("terra_b200: BVH depth %d exceeds the traversal stack (%d entries)\n", f->stack_bound, TB_STACK_SMEM + TB_STACK_LOCAL);
        abort();
    }
    d->graph_epoch++;
    tb_free(d->d_nodes); tb_free(d->d_tris); tb_free(d->d_props); tb_free(d->d_qnodes);
    d->d_qnodes = NULL;
    d->d_nodes = (float4*)tb_upload(f->nodes, f->n_nodes * 16, d);
    d->d_tris = (float4*)tb_upload(f->tris, f->n_tris * 12, d);
    if (f->qnodes) {
        d->d_qnodes = (uint4*)tb_upload(f->qnodes, f->n_nodes * 8, d);
        memcpy(d->grid_lo, f->grid_lo, sizeof d->grid_lo);
        memcpy(d->grid_cell, f->grid_cell, sizeof d->grid_cell);
    }
    d->d_props = (float4*)tb_upload(f->props, f->n_tris * 16, d);
    tb_free(d->d_q4nodes); tb_free(d->d_q4src);
    d->d_q4nodes = NULL; d->d_q4src = NULL; d->n_q4 = 0; d->stack_bound4 = 0;
    if (f->q4nodes && f->n_q4 && f->stack_bound4 <= TB_V7_SMEM_DEPTH + TB_V7_LOCAL) {
        d->d_q4nodes = (uint4*)tb_upload(f->q4nodes, f->n_q4 * 16, d);
        d->d_q4src = (int*)tb_upload(f->q4src, f->n_q4 * 4, d);
        d->n_q4 = f->n_q4; d->stack_bound4 = f->stack_bound4;
    }
    d->stack_bound = f->stack_bound;
    d->exact_boxes = f->exact_boxes;
    TbDevScene& sc = d->sc;
    sc.nodes = d->d_nodes; sc.tris = d->d_tris; sc.props = d->d_props;
    sc.n_nodes = (int)f->n_nodes; sc.n_tris = (int)f->n_tris;
}

// Table half of a commit, run at EVERY commit (the reference reads object->material live at every hit, so an
// in-place material edit followed by a commit must show): materials, light table, textures. Texel data goes up
// again only when the set of textures (address, pixel pointer, shape, modes) changed.
extern "C" void tb_device_upload_tables(TbDevice* d, const TbFlatScene* f) {
    TB_CUDA(cudaSetDevice(d->index));
    TB_CUDA(cudaStreamSynchronize(d->stream));
    d->graph_epoch++;
    tb_free(d->d_mats); tb_free(d->d_lights); tb_free(d->d_light_area); tb_free(d->d_light_rank);
    std::vector<TbDevMaterial> mats(f->materials, f->materials + f->n_objects);
    d->d_mats = tb_upload(mats.data(), mats.size(), d);
    d->d_lights = tb_upload(f->lights, f->n_lights, d);
    d->d_light_area = tb_upload(f->light_tri_area, f->n_light_tris, d);
    d->d_light_rank = tb_upload(f->light_tri_rank, f->n_light_tris, d);
    std::vector<TbTextureSig> sig(f->n_textures);
    for (size_t i = 0; i < f->n_textures; ++i) {
        const TerraTexture* t = f->textures[i];
        TbTextureSig g;
        memset(&g, 0, sizeof g);
        g.host = t; g.pixels = t->pixels; g.width = t->width; g.height = t->height;
        g.components = t->components; g.depth = t->depth; g.filter = t->filter; g.address_mode = t->address_mode;
        sig[i] = g;
    }
    const bool same = sig.size() == d->tex_sig.size() &&
                      (sig.empty() || memcmp(sig.data(), d->tex_sig.data(), sig.size() * sizeof(TbTextureSig)) == 0);
    if (!same) {
        for (void* p : d->tex_pixels) tb_free(p);
        d->tex_pixels.clear();
        tb_free(d->d_texs);
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
        d->h_texs = texs;
        d->tex_sig = sig;
    }
    TbDevScene& sc = d->sc;
    sc.materials = d->d_mats; sc.textures = d->d_texs; sc.lights = d->d_lights;
    sc.light_tri_area = d->d_light_area; sc.light_tri_rank = d->d_light_rank;
    sc.n_objects = (int)f->n_objects; sc.n_lights = (int)f->n_lights;
    sc.lights_triangles_count = f->lights_triangles_count;
    sc.env = f->env;
    sc.preset_mask = 0;
    for (size_t i = 0; i < f->n_objects; ++i) sc.preset_mask |= 1 << f->materials[i].preset;
}

// The device's float nodes (16 floats each, the flatten layout) back to the host: after a refit they are the
// only up-to-date boxes.
extern "C" void tb_device_download_nodes(TbDevice* d, float* out, size_t n_nodes) {
    TB_CUDA(cudaSetDevice(d->index));
    TB_CUDA(cudaStreamSynchronize(d->stream));
    TB_CUDA(cudaMemcpy(out, d->d_nodes, n_nodes * 16 * sizeof(float), cudaMemcpyDeviceToHost));
    d->stats.d2h_bytes += n_nodes * 64;
}

extern "C" void tb_device_set_env_lit(TbDevice* d, int on);
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
    d->graph_epoch++;                                   // graphs bake the scene constants in
    if (d->env_mode == 2) tb_device_set_env_lit(d, 2);  // the importance map belongs to the environment of THIS commit
}

// ==================================================================================== K2: closest hit
// Ray in the frame the reference intersects in: origin already moved 0.001 along the direction
// (terra_scene_raycast, Terra.c:1628-1630), inverse direction (terra_ray, :1706-1708) and the
// Wald-2013 shear/scale + axis permutation (terra_ray_triangle_intersection_init,
// TerraGeometry.c:98-134).
struct TbRay {
    V3    o, inv;
    float sx, sy, sz;
    int   kx, ky, kz;
};

// Axis select as two SELPs: written as a ternary the compiler sank the nine vertex subtractions of
// tb_triangle into per-axis branches (BSSY/BRA/BSYNC per select, divergent across lanes).
__device__ __forceinline__ float tb_sel(V3 v, int k) {
    float r;
    asm("{\n\t.reg .pred p0, p1;\n\tsetp.eq.s32 p0, %4, 0;\n\tsetp.eq.s32 p1, %4, 1;\n\t"
        "selp.f32 %0, %2, %3, p1;\n\tselp.f32 %0, %1, %0, p0;\n\t}" : "=&f"(r) : "f"(v.x), "f"(v.y), "f"(v.z), "r"(k));
    return r;
}

__device__ __forceinline__ V3 tb_offset_origin(V3 o, V3 d) { return vadd(o, vmul(d, 0.001f)); }

__device__ __forceinline__ TbRay tb_ray_setup(V3 origin, V3 d) {
    TbRay r;
    r.o = tb_offset_origin(origin, d);
    r.inv = v3(1.f / d.x, 1.f / d.y, 1.f / d.z);
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz;                                             // terra_max_coefff3, TerraMath.inl:155-169
    if (ax > ay) kz = ax > az ? 0 : 2;
    else         kz = ay > az ? 1 : 2;
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    float dz = tb_sel(d, kz);
    if (dz < 0.f) { int t = kx; kx = ky; ky = t; }
    r.sz = 1.f / dz;
    r.sx = tb_sel(d, kx) * r.sz;
    r.sy = tb_sel(d, ky) * r.sz;
    r.kx = kx; r.ky = ky; r.kz = kz;
    return r;
}

// terra_ray_aabb_intersection, Terra.c:851-878. Operand order of every min/max kept (NaN behaviour).
__device__ __forceinline__ bool tb_box(const TbRay& r, float mnx, float mny, float mnz, float mxx, float mxy, float mxz, float* tmin_out) {
    float t1 = (mnx - r.o.x) * r.inv.x;
    float t2 = (mxx - r.o.x) * r.inv.x;
    float tmin = tminf(t1, t2);
    float tmax = tmaxf(t1, t2);
    t1 = (mny - r.o.y) * r.inv.y;
    t2 = (mxy - r.o.y) * r.inv.y;
    tmin = tmaxf(tmin, tminf(t1, t2));
    tmax = tminf(tmax, tmaxf(t1, t2));
    t1 = (mnz - r.o.z) * r.inv.z;
    t2 = (mxz - r.o.z) * r.inv.z;
    tmin = tmaxf(tmin, tminf(t1, t2));
    tmax = tminf(tmax, tmaxf(t1, t2));
    *tmin_out = tmin;
    return tmax > tmaxf(tmin, 0.f);
}

// terra_ray_triangle_intersection_query, TerraGeometry.c:159-259: watertight, no culling, fp64 retry.
__device__ __forceinline__ bool tb_triangle(const TbRay& r, V3 a, V3 b, V3 c, float* t_out, float* u_out, float* v_out) {
    const V3 A = vsub(a, r.o), B = vsub(b, r.o), C = vsub(c, r.o);
    const float Akz = tb_sel(A, r.kz), Bkz = tb_sel(B, r.kz), Ckz = tb_sel(C, r.kz);
    const float Ax = tb_sel(A, r.kx) - r.sx * Akz;
    const float Ay = tb_sel(A, r.ky) - r.sy * Akz;
    const float Bx = tb_sel(B, r.kx) - r.sx * Bkz;
    const float By = tb_sel(B, r.ky) - r.sy * Bkz;
    const float Cx = tb_sel(C, r.kx) - r.sx * Ckz;
    const float Cy = tb_sel(C, r.ky) - r.sy * Ckz;
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.f || V == 0.f || W == 0.f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    const uint32_t sign_mask = fbits(U) & 0x80000000u;
    if (sign_mask != (fbits(V) & 0x80000000u)) return false;
    if (sign_mask != (fbits(W) & 0x80000000u)) return false;
    const float det = U + V + W;
    if (det == 0.f) return false;
    const float Az = r.sz * Akz, Bz = r.sz * Bkz, Cz = r.sz * Ckz;
    const float ray_depth = U * Az + V * Bz + W * Cz;
    if (bitsf(fbits(ray_depth) ^ sign_mask) < 0.f) return false;
    const float inv_det = 1.f / det;
    *u_out = U * inv_det;
    *v_out = V * inv_det;
    *t_out = ray_depth * inv_det;
    return true;
}

struct TbTraceArgs {
    const float4* nodes;
    const float4* tris;
    const float4* ray_o;        // indexed by slot
    const float4* ray_d;
    const int*    queue;        // NULL: slot = ray index
    const int*    count_ptr;    // device count (wavefront) or NULL -> count
    int           count;
    int*          cursor;       // global fetch cursor (zeroed before launch)
    float4*       hit_out;      // by slot: (t, rank | prim, u, v)
    int*          matq;         // NULL or TB_PRESET_COUNT queues of `matq_stride`
    int*          matq_count;
    int           matq_stride;
    int           emit_prim;    // 1: write object|tri<<8 (API), 0: write rank (wavefront)
    unsigned long long* visit_counters;   // COUNT variant only: [0] nodes fetched, [1] triangles tested
    const uint4*  qnodes;       // quantised 32-byte nodes (tb_trace.cuh), or NULL
    float         grid_lo[3], grid_cell[3];
    int           tri_min, refetch_below;   // rounds kernels: parked leaves needed for a triangle round; refill below this many live lanes
    int           stack_smem;               // k_trace_v6: stack entries per lane held in (dynamic) shared memory
    float4*       lo;                       // fused ray generation: zeroed for the paths that leave the scene (k_trace_packet<.., GEN>)
    const uint4*  q4nodes;                  // k_trace_v7: the 4-wide tree, 4 x uint4 per node
    unsigned long long* round_stats;        // COUNT: lane-rounds by state, see terra_b200_round_stats
};

// ==================================================================================== K1: ray generation
struct TbBatch {               // which (pixel, sample) pairs a wave holds: slot = s_local * n_pixels + p_local
    int      x, y, w, h;       // render region inside the framebuffer
    int      fbw, fbh;
    int      p0, n_pixels;     // region-linear pixel range of this batch
    int      s0, n_samples;    // sample range of this batch (relative to sample_base)
    uint32_t sample_base;
    uint64_t seed;
    int      tiled;            // 1: region-linear index q walks 8x4 pixel tiles (w % 8 == 0 && h % 4 == 0)
    int      sub;              // tiled: samples of one pixel that share a warp (power of two dividing n_samples, <= 32)
};

// Everything a camera ray is a function of. The packet kernel and the first shade pass derive the ray from the
// slot index with this instead of reading it back from HBM (`on`), k_raygen writes it out for the kernels that
// cannot (work-counting walk, A/B kernels).
struct TbGenArgs {
    TbBatch batch;              // wave shape; x, y, fbw, fbh, sample_base and seed are NOT read on the device (see call)
    float   jitter;
    int     on;                 // 0: rays come from the ray arrays
    const TbCallParams* call;   // what differs between two calls of one shape — region origin, framebuffer size, seed, camera —
                                // in device memory, so that the arguments of a kernel captured in a CUDA graph never change
};

struct TbRaygenArgs {
    TbGenArgs gen;
    float4 *ray_o, *ray_d, *thr, *lo;
    int*    queue;
};

// Region-linear index -> pixel offset inside the region. Tiled: consecutive indices fill an 8x4 pixel
// tile, so the 32 camera rays of a warp span 8x4 pixels instead of 32x1 (more coherent traversal).
__device__ __forceinline__ void tb_region_pixel(const TbBatch& b, int pr, int* rx, int* ry) {
    if (b.tiled) {
        const int tile = pr >> 5, in = pr & 31, tiles_x = b.w >> 3;
        const int ty = tile / tiles_x, tx = tile - ty * tiles_x;
        *rx = tx * 8 + (in & 7);
        *ry = ty * 4 + (in >> 3);
    } else {
        *ry = pr / b.w;
        *rx = pr - *ry * b.w;
    }
}

// Wave slot <-> (pixel of the batch, sample of the batch). Linear: sample-major. Tiled: a warp's 32 slots are
// `sub` samples of each of 32/sub neighbouring pixels (sub = gcd(n_samples, 32) unless capped), and
// consecutive warps are the next samples of the same pixels, then the next pixels of the 8x4 tile order:
//   slot = ((p / P) * (n_samples / sub) + s / sub) * 32 + (p % P) * sub + s % sub,   P = 32 / sub.
// sub = 32: a warp is 32 samples of ONE pixel — its camera rays differ by the sub-pixel jitter only, so the
// packet kernel walks barely more nodes than a single ray would, and the bounce rays of a pixel sit together.
__device__ __forceinline__ int tb_slot_of(const TbBatch& b, int p_local, int s_local) {
    if (b.tiled) {
        const int sub = b.sub, P = 32 / sub;
        return ((p_local / P) * (b.n_samples / sub) + s_local / sub) * 32 + (p_local % P) * sub + s_local % sub;
    }
    return s_local * b.n_pixels + p_local;
}

// -> pixel offset inside the region and sample index relative to the call's sample_base
__device__ __forceinline__ void tb_slot_to_pixel(const TbBatch& b, int slot, int* rx_out, int* ry_out, uint32_t* sample) {
    int s_local, p_local;
    if (b.tiled) {
        const int sub = b.sub, P = 32 / sub, sg = b.n_samples / sub;
        const int group = slot >> 5, in = slot & 31;
        const int pgroup = group / sg, sgroup = group - pgroup * sg;
        p_local = pgroup * P + in / sub;
        s_local = sgroup * sub + in % sub;
    } else {
        s_local = slot / b.n_pixels;
        p_local = slot - s_local * b.n_pixels;
    }
    tb_region_pixel(b, b.p0 + p_local, rx_out, ry_out);
    *sample = (uint32_t)(b.s0 + s_local);
}

// terra_camera_perspective_sample + terra_transformf3 + terra_ray (Terra.c:555-560, 1783-1799)
__device__ __forceinline__ V3 tb_camera_dir(const TbCameraFrame& cam, int px, int py, int fbw, int fbh, float jitter, float r1, float r2) {
    const float dx = -jitter + 2 * r1 * jitter;
    const float dy = -jitter + 2 * r2 * jitter;
    const float ndc_x = ((float)px + 0.5f + dx) / (float)fbw;
    const float ndc_y = ((float)py + 0.5f + dy) / (float)fbh;
    const float screen_x = 2 * ndc_x - 1;
    const float screen_y = 1 - 2 * ndc_y;
    const float aspect_ratio = (float)fbw / (float)fbh;
    const float frustum_x = screen_x * aspect_ratio * cam.tan_half_fov;
    const float frustum_y = screen_y * cam.tan_half_fov;
    V3 dir = vnorm(v3(frustum_x, frustum_y, 1.f));
    M3 rot;
    rot.r0 = v3(cam.rot[0][0], cam.rot[0][1], cam.rot[0][2]);
    rot.r1 = v3(cam.rot[1][0], cam.rot[1][1], cam.rot[1][2]);
    rot.r2 = v3(cam.rot[2][0], cam.rot[2][1], cam.rot[2][2]);
    return mtransform(rot, dir);
}

// The per-call values are loaded where a ray is generated, with volatile loads: hoisted out of a persistent
// kernel's loop they would pin ~25 registers for the whole kernel (the line stays in L1).
__device__ __forceinline__ float tb_ldv(const float* p) { float v; asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p)); return v; }
__device__ __forceinline__ int tb_ldv(const int* p) { int v; asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(p)); return v; }
__device__ __forceinline__ uint32_t tb_ldv(const uint32_t* p) { uint32_t v; asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p)); return v; }
__device__ __forceinline__ uint64_t tb_ldv(const uint64_t* p) { unsigned long long v; asm volatile("ld.global.nc.u64 %0, [%1];" : "=l"(v) : "l"(p)); return (uint64_t)v; }

// Camera ray of wave slot `slot` (Terra.c:553-560): direction, origin (the camera position), and the path key that
// rides in the .w lanes of the ray records.
__device__ __forceinline__ V3 tb_gen_ray(const TbGenArgs& g, int slot, uint64_t* key_out, V3* origin_out) {
    const TbCallParams* c = g.call;
    int rx, ry; uint32_t s_rel;
    tb_slot_to_pixel(g.batch, slot, &rx, &ry, &s_rel);
    const int px = tb_ldv(&c->x) + rx, py = tb_ldv(&c->y) + ry;
    const int fbw = tb_ldv(&c->fbw), fbh = tb_ldv(&c->fbh);
    const uint32_t sample = tb_ldv(&c->sample_base) + s_rel;
    const uint64_t key = tb_path_key(tb_ldv(&c->seed), (uint32_t)(py * fbw + px), sample);
    const float r1 = tb_randf_pcg(key, TB_DIM_JITTER_X), r2 = tb_randf_pcg(key, TB_DIM_JITTER_Y);
    TbCameraFrame cam;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) cam.rot[i][j] = tb_ldv(&c->cam.rot[i][j]);
        cam.pos[i] = tb_ldv(&c->cam.pos[i]);
    }
    cam.tan_half_fov = tb_ldv(&c->cam.tan_half_fov);
    *key_out = key;
    *origin_out = v3(cam.pos[0], cam.pos[1], cam.pos[2]);
    return tb_camera_dir(cam, px, py, fbw, fbh, g.jitter, r1, r2);
}

#include "tb_trace.cuh"
#include "tb_trace_v6.cuh"
#include "tb_trace_v7.cuh"
#include "tb_trace_v8.cuh"
#include "tb_trace_packet.cuh"

// ==================================================================================== K1: ray generation kernel
__global__ void __launch_bounds__(256) k_raygen(TbRaygenArgs a) {
    const int n = a.gen.batch.n_pixels * a.gen.batch.n_samples;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += gridDim.x * blockDim.x) {
        uint64_t key; V3 o;
        const V3 d = tb_gen_ray(a.gen, slot, &key, &o);
        // the path key rides in the .w lanes of the ray so later stages need no extra array
        a.ray_o[slot] = make_float4(o.x, o.y, o.z, __uint_as_float((uint32_t)key));
        a.ray_d[slot] = make_float4(d.x, d.y, d.z, __uint_as_float((uint32_t)(key >> 32)));
        a.thr[slot] = make_float4(1.f, 1.f, 1.f, 0.f);
        a.lo[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
        a.queue[slot] = slot;
    }
}

// Primary rays at jitter 0 for the parity entry point: one float4 origin/direction per pixel.
__global__ void __launch_bounds__(256) k_primary_rays(TbCameraFrame cam, int w, int h, float4* ray_o, float4* ray_d) {
    const int n = w * h;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const V3 d = tb_camera_dir(cam, p % w, p / w, w, h, 0.f, 0.f, 0.f);
        ray_o[p] = make_float4(cam.pos[0], cam.pos[1], cam.pos[2], 0.f);
        ray_d[p] = make_float4(d.x, d.y, d.z, 0.f);
    }
}

__global__ void __launch_bounds__(256) k_pack_rays(size_t n, const float* o3, const float* d3, float4* ray_o, float4* ray_d) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        ray_o[i] = make_float4(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2], 0.f);
        ray_d[i] = make_float4(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2], 0.f);
    }
}

__global__ void __launch_bounds__(256) k_pack_rays_noalloc(size_t n, const float* __restrict__ o3, const float* __restrict__ d3) {
    float4* ray_o = (float4*)(g_noalloc_scratch + TB_NOALLOC_HEADER);
    float4* ray_d = ray_o + TB_NOALLOC_MAX_RAYS;
    if (blockIdx.x == 0 && threadIdx.x == 0) *(int*)g_noalloc_scratch = 0;      // the trace kernel's fetch cursor must start at zero
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        ray_o[i] = make_float4(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2], 0.f);
        ray_d[i] = make_float4(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2], 0.f);
    }
}

// ==================================================================================== K3: surface
// terra_surface_init (Terra.c:1726-1764): barycentrics are re-derived from the hit point with a
// Cramer solve (the intersector's u,v are ignored), normal/texcoord blended as
// c*v + b*u + a*(1-u-v), attributes evaluated, skewed tangent frame built.
struct TbHitInfo { int rank, object, preset; V3 a, b, c; V2 texcoord; };

// `tag` = object | preset << 8: the trace kernels leave it in the .z lane of a wavefront hit record (the intersector's
// u, v are not used by shading), so the material record is requested together with the triangle, not after it.
__device__ __forceinline__ TbSurface tb_surface_init(const TbDevScene& sc, int rank, int tag, V3 point, TbHitInfo* info) {
    const TbDevMaterial* m = sc.materials + (tag & 0xff);
    const int cnt = m->attr_count;
    const float4* tp = sc.tris + (size_t)rank * 3;
    const float4 ta = __ldg(tp), tb = __ldg(tp + 1), tc = __ldg(tp + 2);
    const float4* pp = sc.props + (size_t)rank * 4;
    const float4 p0 = __ldg(pp), p1 = __ldg(pp + 1), p2 = __ldg(pp + 2), p3 = __ldg(pp + 3);
    const V3 a = v3(ta.x, ta.y, ta.z), b = v3(tb.x, tb.y, tb.z), c = v3(tc.x, tc.y, tc.z);
    const V3 na = v3(p0.x, p0.y, p0.z), nb = v3(p0.w, p1.x, p1.y), nc = v3(p1.z, p1.w, p2.x);
    const V2 t_a = { p2.y, p2.z }, t_b = { p2.w, p3.x }, t_c = { p3.y, p3.z };
    info->rank = rank; info->object = tag & 0xff; info->preset = (tag >> 8) & 0xff;
    info->a = a; info->b = b; info->c = c;

    const V3 e0 = vsub(b, a), e1 = vsub(c, a), p = vsub(point, a);
    const float d00 = vdot(e0, e0), d11 = vdot(e1, e1), d01 = vdot(e0, e1);
    const float dp0 = vdot(p, e0), dp1 = vdot(p, e1);
    const float div = d00 * d11 - d01 * d01;
    const float ux = (d11 * dp0 - d01 * dp1) / div;
    const float uy = (d00 * dp1 - d01 * dp0) / div;
    const float uw = 1 - ux - uy;
    TbSurface s;
    s.n = vnorm(vadd(vadd(vmul(nc, uy), vmul(nb, ux)), vmul(na, uw)));
    V2 tex;
    tex.x = (t_c.x * uy + t_b.x * ux) + t_a.x * uw;
    tex.y = (t_c.y * uy + t_b.y * ux) + t_a.y * uw;
    info->texcoord = tex;
    s.a0 = cnt > 0 ? tb_attr_eval(m->attr[0], sc.textures, tex, point) : v3s(0.f);
    s.a1 = cnt > 1 ? tb_attr_eval(m->attr[1], sc.textures, tex, point) : v3s(0.f);
    s.a2 = cnt > 2 ? tb_attr_eval(m->attr[2], sc.textures, tex, point) : v3s(0.f);
    s.state = 0.f;
    if (info->preset == TB_PRESET_PHONG && cnt > 3) s.state = tb_attr_eval(m->attr[3], sc.textures, tex, point).x;
    if (info->preset == TB_PRESET_GLASS) s.state = s.a2.x;
    s.emissive = tb_attr_eval(m->emissive, sc.textures, tex, point);
    s.xf = basis_from_normal(s.n);
    return s;
}

// ==================================================================================== K4: shade
struct TbShadeArgs {
    TbDevScene sc;
    float4 *ray_o, *ray_d, *thr, *lo;
    const float4* hit;
    float4 *aux_o, *aux_d;
    const float4* aux_hit;
    const int* q;              // slots of this preset
    const int* q_count;
    int*  next_q;              // live paths for the next bounce
    int*  next_count;
    int*  auxq;                // visibility rays to trace between the two stages
    int*  aux_count;
    int   bounce;
    int   env_base;            // aux id of slot s's environment visibility ray = env_base + s
    TbGenArgs gen;             // gen.on (first bounce after a fused-raygen trace): the ray is derived from the slot, throughput is 1,
                               // and the radiance is stored, not added (nothing initialised the path state of a path that hit)
};

// Warp-aggregated append; every lane of the warp must call it (full mask).
__device__ __forceinline__ int tb_append(int* counter, bool pred) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (!m) return -1;
    int base = 0;
    if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    return pred ? base + __popc(m & ((1u << lane) - 1u)) : -1;
}

// One explicit light sample: terra_scene_pick_light + terra_light_pick_triangle +
// terra_light_sample_triangle (Terra.c:1592-1621, 1662-1697), randoms as at :1370-1388.
struct TbLightSample { int light, tri, object; uint32_t rank; V3 pos, norm; float tri_area; };

__device__ __forceinline__ TbLightSample tb_sample_light(const TbDevScene& sc, uint64_t key, uint32_t dim0) {
    TbLightSample ls;
    const float e_pick = (float)((double)tb_randf_libc(key, dim0 + TB_DIM_LIGHT_PICK) - TB_EPS_D);
    long long li = (long long)((double)e_pick * (double)sc.n_lights);
    if (li < 0) li = 0;
    if (li > sc.n_lights - 1) li = sc.n_lights - 1;         // e == 1.0 would index past the table in the reference
    const TbDevLight L = sc.lights[li];
    const float e_tri = tb_randf_libc(key, dim0 + TB_DIM_TRI_PICK);
    long long ti = (long long)(e_tri * (float)L.tri_count);
    if (ti > (long long)L.tri_count - 1) ti = (long long)L.tri_count - 1;
    ls.light = (int)li; ls.tri = (int)ti; ls.object = L.object;
    ls.rank = sc.light_tri_rank[L.tri_base + ti];
    ls.tri_area = sc.light_tri_area[L.tri_base + ti];
    const float e1 = tb_randf_libc(key, dim0 + TB_DIM_LIGHT_U), e2 = tb_randf_libc(key, dim0 + TB_DIM_LIGHT_V);
    const float s = sqrtf(e1);
    const float a = 1 - s, b = e2 * s, c = 1 - a - b;
    const float4* tp = sc.tris + (size_t)ls.rank * 3;
    const float4 ta = __ldg(tp), tb = __ldg(tp + 1), tc = __ldg(tp + 2);
    ls.pos = vadd(vadd(vmul(v3(ta.x, ta.y, ta.z), a), vmul(v3(tb.x, tb.y, tb.z), b)), vmul(v3(tc.x, tc.y, tc.z), c));
    const float4* pp = sc.props + (size_t)ls.rank * 4;
    const float4 p0 = __ldg(pp), p1 = __ldg(pp + 1), p2 = __ldg(pp + 2);
    const V3 na = v3(p0.x, p0.y, p0.z), nb = v3(p0.w, p1.x, p1.y), nc = v3(p1.z, p1.w, p2.x);
    ls.norm = vnorm(vadd(vadd(vmul(na, a), vmul(nb, b)), vmul(nc, c)));
    return ls;
}

// terra_surface_ray, Terra.c:1715-1724 (sign = +1)
__device__ __forceinline__ V3 tb_surface_ray_origin(const TbSurface& s, V3 point) { return vadd(point, vmul(s.n, 0.0001f * 1.f)); }

// Debug views, terra_integrate_debug_{mono,depth,normals,mis_weight} (Terra.c:1140-1337). All of them
// return zero past bounce 0, so the host loop runs one bounce only. `ray_origin` is the caller's ray
// origin (what terra_integrate_debug_depth sees), not the one terra_scene_raycast moved by 1e-3 * dir.
__device__ __forceinline__ V3 tb_integrate_debug(const TbDevScene& sc, int integ, TbSurface& s, const TbHitInfo& info,
                                                 V3 point, V3 ray_origin, V3 wo, V3 thr, uint64_t key, uint32_t dim0,
                                                 const float4* aux_hit, int preset) {
    if (integ == kTerraIntegratorDebugMono) return v3s(1.f);
    if (integ == kTerraIntegratorDebugDepth) {                                 // :1148-1157
        const float far_plane = 500.f;
        return v3s(vlen(vsub(ray_origin, point)) / far_plane);
    }
    if (integ == kTerraIntegratorDebugNormals) {                               // :1159-1198
        // terra_clampf3 (TerraMath.inl:290-299): max first, then min, `>` / `<` as written
        V3 p, n;
        p.x = s.n.x > 0.f ? s.n.x : 0.f; p.y = s.n.y > 0.f ? s.n.y : 0.f; p.z = s.n.z > 0.f ? s.n.z : 0.f;
        p.x = p.x < 1.f ? p.x : 1.f;     p.y = p.y < 1.f ? p.y : 1.f;     p.z = p.z < 1.f ? p.z : 1.f;
        n.x = s.n.x > -1.f ? s.n.x : -1.f; n.y = s.n.y > -1.f ? s.n.y : -1.f; n.z = s.n.z > -1.f ? s.n.z : -1.f;
        n.x = n.x < 0.f ? n.x : 0.f;     n.y = n.y < 0.f ? n.y : 0.f;     n.z = n.z < 0.f ? n.z : 0.f;
        n = vmul(n, -1.f);
        V3 c = v3s(0.f);
        c = vadd(c, vmul(v3(1, 0, 0), p.x)); c = vadd(c, vmul(v3(0, 1, 0), p.y)); c = vadd(c, vmul(v3(0, 0, 1), p.z));
        c = vadd(c, vmul(v3(0, 1, 1), n.x)); c = vadd(c, vmul(v3(1, 0, 1), n.y)); c = vadd(c, vmul(v3(1, 1, 0), n.z));
        return c;
    }
    // ---- kTerraIntegratorDebugMisWeights, :1200-1337: blue = weight of the light sample, red = of the BSDF sample
    V3 Lo = v3s(0.f);
    if (sc.n_lights <= 0) return Lo;
    const V3 bsdf_sample = tb_bsdf_sample(preset, s, tb_randf_libc(key, dim0 + TB_DIM_MIS0), tb_randf_libc(key, dim0 + TB_DIM_MIS1),
                                          tb_randf_libc(key, dim0 + TB_DIM_MIS2), wo);
    const TbLightSample ls = tb_sample_light(sc, key, dim0);
    {
        const V3 p_to_light = vsub(ls.pos, point);
        const V3 wi = vnorm(p_to_light);
        const float4 ah = aux_hit[0];
        if (ah.x >= 0.f && (uint32_t)__float_as_int(ah.y) == ls.rank) {
            const float cosl = vdot(ls.norm, vneg(wi));
            if (cosl > 0) {
                const float bsdf_pdf = tb_bsdf_pdf(preset, s, wi, wo);
                const float light_pdf = vsqlen(p_to_light) / fabsf(cosl * ls.tri_area);
                const float weight = (bsdf_pdf * bsdf_pdf) / (light_pdf * light_pdf + bsdf_pdf * bsdf_pdf);
                Lo = vadd(Lo, v3(0.f, 0.f, weight));
            }
        }
    }
    {
        const V3 wi = bsdf_sample;
        const float bsdf_pdf = tb_bsdf_pdf(preset, s, wi, wo);
        const float4 bh = aux_hit[1];
        if (bh.x >= 0.f) {
            const V3 bo = tb_offset_origin(tb_surface_ray_origin(s, point), wi);
            const V3 bpoint = vadd(bo, vmul(wi, bh.x));
            TbHitInfo binfo;
            const TbSurface bsurf = tb_surface_init(sc, __float_as_int(bh.y), __float_as_int(bh.z), bpoint, &binfo);
            if (binfo.object == sc.lights[ls.light].object) {
                const float NoW = vdot(bsurf.n, vneg(wi));
                if (NoW > 0) {
                    const float dist = vsqdist(bpoint, point);
                    const V3 ab = vsub(binfo.b, binfo.a), ac = vsub(binfo.c, binfo.a);
                    const float area = vlen(vcross(ab, ac)) / 2;
                    const float light_pdf = dist / (NoW * area);
                    const float weight = (bsdf_pdf * bsdf_pdf) / (light_pdf * light_pdf + bsdf_pdf * bsdf_pdf);
                    Lo = vadd(Lo, v3(weight, 0.f, 0.f));
                }
            }
        }
    }
    return vpoint(Lo, thr);
}

#include "tb_distribution.cuh"

// MODE: the integrator family the kernel is compiled for — 0 simple (Terra.c:1339-1347), 1 direct / direct MIS
// (:1349-1587, plus the explicit environment samples of env-lit mode 2), 2 the debug views (:1140-1337). One kernel
// for all of them carried the registers of the MIS path (104-118, two CTAs per SM) through the simple integrator
// every benchmark runs: ncu r1w showed it latency-bound at a quarter of the SM's warps.
enum { TB_SHADE_SIMPLE = 0, TB_SHADE_DIRECT = 1, TB_SHADE_DEBUG = 2 };
template <int PRESET, int STAGE, int MODE>
__global__ void __launch_bounds__(TB_SHADE_THREADS, MODE == TB_SHADE_SIMPLE ? TB_SHADE_MIN_BLOCKS_SIMPLE : TB_SHADE_MIN_BLOCKS) k_shade(TbShadeArgs a) {
    const int count = *a.q_count;
    const int stride = gridDim.x * blockDim.x;
    const int rounds = (count + stride - 1) / stride;
    const TbDevScene& sc = a.sc;
    const bool first = a.gen.on != 0;
    // The kernel is bound by the latency of a chain of dependent gathers (ncu r1w / r2d: long-scoreboard 6-35 stalled
    // warps per issue, issue-active 12-36 %): queue entry -> slot -> hit record -> triangle / properties / material.
    // The first two links do not depend on the shading of the previous hit, so they are fetched one and two rounds
    // ahead (5 registers) and only the last link is exposed: -3 % on the non-trace part of the C2 step (r2i A/B;
    // 4 CTAs per SM at 64 registers instead: no gain on C2, -3 % on C5).
    const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
    int slot_c = i0 < count ? a.q[i0] : 0;
    int slot_n = i0 + stride < count ? a.q[i0 + stride] : 0;
    float4 hit_c = i0 < count ? a.hit[slot_c] : make_float4(0, 0, 0, 0);
    for (int it = 0; it < rounds; ++it) {
        const int i = it * stride + i0;
        const bool valid = i < count;
        const float4 hit_n = i + stride < count ? a.hit[slot_n] : make_float4(0, 0, 0, 0);
        const int slot_nn = i + 2 * stride < count ? a.q[i + 2 * stride] : 0;
        bool alive = false, want_aux0 = false, want_aux1 = false, want_aux2 = false;
        int slot = 0;
        V3 next_o = v3s(0.f), next_d = v3s(0.f), aux0_d = v3s(0.f), aux1_d = v3s(0.f), aux2_d = v3s(0.f), aux_o = v3s(0.f);
        float4 o4 = make_float4(0, 0, 0, 0), d4 = o4;
        if (valid) {
            slot = slot_c;
            if (first) {
                uint64_t k0; V3 go;
                const V3 gd = tb_gen_ray(a.gen, slot, &k0, &go);
                o4 = make_float4(go.x, go.y, go.z, __uint_as_float((uint32_t)k0));
                d4 = make_float4(gd.x, gd.y, gd.z, __uint_as_float((uint32_t)(k0 >> 32)));
            } else {
                o4 = a.ray_o[slot]; d4 = a.ray_d[slot];
            }
            const float4 h4 = hit_c;
            const uint64_t key = ((uint64_t)__float_as_uint(d4.w) << 32) | __float_as_uint(o4.w);
            const uint32_t dim0 = 2u + (uint32_t)a.bounce * TB_DIMS_PER_BOUNCE;
            const V3 d = v3(d4.x, d4.y, d4.z);
            const V3 ro = tb_offset_origin(v3(o4.x, o4.y, o4.z), d);
            const V3 point = vadd(ro, vmul(d, h4.x));                        // terra_ray_pos, TerraGeometry.c:9-12
            TbHitInfo info;
            TbSurface s = tb_surface_init(sc, __float_as_int(h4.y), __float_as_int(h4.z), point, &info);
            const V3 wo = vneg(d);
            const int integ = sc.integrator;
            const bool debug = MODE == TB_SHADE_DEBUG;
            const bool dbg_mis = debug && integ == kTerraIntegratorDebugMisWeights;   // bounce 0 only (host loop)
            const bool direct = MODE == TB_SHADE_DIRECT || dbg_mis;
            const bool mis = (MODE == TB_SHADE_DIRECT && integ == kTerraIntegratorDirectMis) || dbg_mis;

            const bool env_nee = sc.env_nee && direct && !dbg_mis;                    // extension: explicit environment samples
            if (STAGE == 0) {
                // ---- emit the visibility rays of terra_integrate_direct / _direct_mis
                if (direct && (sc.n_lights > 0 || env_nee)) {
                    aux_o = tb_surface_ray_origin(s, point);
                    if (mis) {
                        aux1_d = tb_bsdf_sample(PRESET, s, tb_randf_libc(key, dim0 + TB_DIM_MIS0), tb_randf_libc(key, dim0 + TB_DIM_MIS1),
                                                tb_randf_libc(key, dim0 + TB_DIM_MIS2), wo);
                        want_aux1 = true;
                    }
                    if (sc.n_lights > 0) {
                        const TbLightSample ls = tb_sample_light(sc, key, dim0);
                        aux0_d = vnorm(vsub(ls.pos, point));
                        want_aux0 = true;
                    }
                    if (env_nee) {
                        float pdf_e;
                        want_aux2 = tb_env_sample(sc.env_dist, tb_randf_libc(key, dim0 + TB_DIM_ENV0), tb_randf_libc(key, dim0 + TB_DIM_ENV1), &aux2_d, &pdf_e)
                                    && vdot(aux2_d, s.n) > 0.f;
                    }
                }
            } else {
                // ---- terra_integrate (Terra.c:1099-1138)
                V3 thr = v3s(1.f);
                if (!first) { const float4 t4 = a.thr[slot]; thr = v3(t4.x, t4.y, t4.z); }
                V3 Lo = v3s(0.f);
                if (debug) {
                    Lo = tb_integrate_debug(sc, integ, s, info, point, v3(o4.x, o4.y, o4.z), wo, thr, key, dim0, a.aux_hit + (size_t)slot * 2, PRESET);
                } else if (!direct) {                                          // simple, :1339-1347
                    if (vdot(wo, s.n) > 0) Lo = vpoint(s.emissive, thr);
                } else {
                    if (a.bounce == 0 && vdot(wo, s.n) > 0) Lo = vadd(Lo, s.emissive);
                    V3 bsdf_sample = v3s(0.f);
                    if (mis && (sc.n_lights > 0 || env_nee))
                        bsdf_sample = tb_bsdf_sample(PRESET, s, tb_randf_libc(key, dim0 + TB_DIM_MIS0), tb_randf_libc(key, dim0 + TB_DIM_MIS1),
                                                     tb_randf_libc(key, dim0 + TB_DIM_MIS2), wo);
                    if (sc.n_lights > 0) {
                        const TbLightSample ls = tb_sample_light(sc, key, dim0);
                        const float light_pick_pdf = 1.f / (float)sc.lights_triangles_count;
                        const V3 p_to_light = vsub(ls.pos, point);
                        const V3 wi = vnorm(p_to_light);
                        const float4 ah = a.aux_hit[(size_t)slot * 2];
                        if (ah.x >= 0.f && (uint32_t)__float_as_int(ah.y) == ls.rank) {      // same object AND triangle, :1403
                            const float cosl = vdot(vneg(wi), ls.norm);
                            if (cosl > 0) {
                                const V3 lo_ = tb_offset_origin(tb_surface_ray_origin(s, point), wi);
                                const V3 lpoint = vadd(lo_, vmul(wi, ah.x));
                                TbHitInfo linfo;
                                const TbSurface lsurf = tb_surface_init(sc, (int)ls.rank, ls.object, lpoint, &linfo);
                                const float light_pdf = vsqlen(p_to_light) / fabsf(cosl * ls.tri_area);
                                if (!mis) {                                    // :1417-1422
                                    const V3 f = tb_bsdf_eval(PRESET, s, wi, wo);
                                    V3 Ld = vpoint(lsurf.emissive, f);
                                    Ld = vmul(Ld, vdot(wi, s.n) / (light_pdf * light_pick_pdf));
                                    Lo = vadd(Lo, Ld);
                                } else {                                       // :1505-1514
                                    const float bsdf_pdf = tb_bsdf_pdf(PRESET, s, wi, wo);
                                    const float weight = (light_pdf * light_pdf) / (light_pdf * light_pdf + bsdf_pdf * bsdf_pdf);
                                    if (light_pdf != 0) {
                                        const V3 f = tb_bsdf_eval(PRESET, s, wi, wo);
                                        V3 L = vpoint(lsurf.emissive, f);
                                        L = vmul(L, vdot(wi, s.n) * weight / (light_pdf * light_pick_pdf));
                                        Lo = vadd(Lo, L);
                                    }
                                }
                            }
                        }
                        if (mis) {                                             // :1517-1583
                            const V3 wi2 = bsdf_sample;
                            const V3 f = tb_bsdf_eval(PRESET, s, wi2, wo);
                            const float bsdf_pdf = tb_bsdf_pdf(PRESET, s, wi2, wo);
                            const float4 bh = a.aux_hit[(size_t)slot * 2 + 1];
                            if (bh.x >= 0.f) {
                                const V3 bo = tb_offset_origin(tb_surface_ray_origin(s, point), wi2);
                                const V3 bpoint = vadd(bo, vmul(wi2, bh.x));
                                TbHitInfo binfo;
                                const TbSurface bsurf = tb_surface_init(sc, __float_as_int(bh.y), __float_as_int(bh.z), bpoint, &binfo);
                                if (binfo.object == sc.lights[ls.light].object) {
                                    const float NoW = vdot(bsurf.n, vneg(wi2));
                                    if (NoW > 0) {
                                        const float dist = vsqdist(bpoint, point);
                                        const V3 ab = vsub(binfo.b, binfo.a), ac = vsub(binfo.c, binfo.a);
                                        const float area = vlen(vcross(ab, ac)) / 2;            // terra_triangle_area, :1833-1838
                                        const float light_pdf = dist / (NoW * area);
                                        const float weight = (bsdf_pdf * bsdf_pdf) / (light_pdf * light_pdf + bsdf_pdf * bsdf_pdf);
                                        if (bsdf_pdf != 0) {
                                            V3 L = vpoint(bsurf.emissive, f);
                                            L = vmul(L, vdot(wi2, s.n) * weight / bsdf_pdf);
                                            Lo = vadd(Lo, L);
                                        }
                                    }
                                }
                            }
                        }
                    }
                    if (env_nee) {
                        // ---- explicit environment sample (no reference counterpart; same estimator shapes as the
                        //      area-light code above: f * L * cos / pdf, power-heuristic weights under MIS)
                        V3 wi_e; float pdf_e;
                        if (tb_env_sample(sc.env_dist, tb_randf_libc(key, dim0 + TB_DIM_ENV0), tb_randf_libc(key, dim0 + TB_DIM_ENV1), &wi_e, &pdf_e)
                            && vdot(wi_e, s.n) > 0.f && a.aux_hit[(size_t)a.env_base + slot].x < 0.f) {
                            V2 uv; uv.x = wi_e.x; uv.y = wi_e.y;
                            const V3 env = tb_attr_eval(sc.env, sc.textures, uv, wi_e);
                            const V3 f = tb_bsdf_eval(PRESET, s, wi_e, wo);
                            float weight = 1.f;
                            if (mis) {
                                const float bsdf_pdf = tb_bsdf_pdf(PRESET, s, wi_e, wo);
                                weight = (pdf_e * pdf_e) / (pdf_e * pdf_e + bsdf_pdf * bsdf_pdf);
                            }
                            Lo = vadd(Lo, vmul(vpoint(env, f), vdot(wi_e, s.n) * weight / pdf_e));
                        }
                        if (mis && a.aux_hit[(size_t)slot * 2 + 1].x < 0.f) {       // the BSDF sample left the scene
                            const V3 wi2 = bsdf_sample;
                            const float bsdf_pdf = tb_bsdf_pdf(PRESET, s, wi2, wo);
                            if (bsdf_pdf != 0) {
                                // the environment technique cannot produce a direction below the surface: weight 1 there
                                const float pe = vdot(wi2, s.n) > 0.f ? tb_env_pdf(sc.env_dist, wi2) : 0.f;
                                const float weight = (bsdf_pdf * bsdf_pdf) / (pe * pe + bsdf_pdf * bsdf_pdf);
                                V2 uv; uv.x = wi2.x; uv.y = wi2.y;
                                const V3 env = tb_attr_eval(sc.env, sc.textures, uv, wi2);
                                const V3 f = tb_bsdf_eval(PRESET, s, wi2, wo);
                                Lo = vadd(Lo, vmul(vpoint(env, f), vdot(wi2, s.n) * weight / bsdf_pdf));
                            }
                        }
                    }
                    Lo = vpoint(Lo, thr);
                }
                if (first) {
                    a.lo[slot] = make_float4(Lo.x, Lo.y, Lo.z, 0.f);                            // Lo = 0 + radiance
                } else if (Lo.x != 0.f || Lo.y != 0.f || Lo.z != 0.f) {
                    const float4 l4 = a.lo[slot];
                    a.lo[slot] = make_float4(l4.x + Lo.x, l4.y + Lo.y, l4.z + Lo.z, 0.f);   // Lo = Lo + radiance, :1063
                }
                // ---- continue the path (terra_trace, Terra.c:1065-1093)
                // (debug views add nothing after bounce 0, :1140-1337: their paths end here)
                if (a.bounce < sc.bounces && !debug) {
                    const float e0 = tb_randf_libc(key, dim0 + TB_DIM_BSDF0), e1 = tb_randf_libc(key, dim0 + TB_DIM_BSDF1);
                    const float e2 = tb_randf_libc(key, dim0 + TB_DIM_BSDF2);
                    const V3 wi = tb_bsdf_sample(PRESET, s, e0, e1, e2, wo);
                    const float pdf = tmaxf(tb_bsdf_pdf(PRESET, s, wi, wo), TB_EPS_F);
                    V3 f = tb_bsdf_eval(PRESET, s, wi, wo);
                    f = vmul(f, 1.f / pdf);
                    thr = vpoint(thr, f);
                    thr = vmul(thr, vdot(s.n, wi));
                    const float p = tmaxf(thr.x, tmaxf(thr.y, thr.z));
                    const float e3 = tb_randf_libc(key, dim0 + TB_DIM_RR);
                    if (!(e3 > p)) {
                        thr = vmul(thr, (float)(1.f / ((double)p + TB_EPS_D)));
                        a.thr[slot] = make_float4(thr.x, thr.y, thr.z, 0.f);
                        next_o = tb_surface_ray_origin(s, point);
                        next_d = wi;
                        alive = true;
                    }
                }
            }
        }
        if (STAGE == 0) {
            const int i0 = tb_append(a.aux_count, want_aux0);
            if (want_aux0) {
                a.aux_o[(size_t)slot * 2] = make_float4(aux_o.x, aux_o.y, aux_o.z, 0.f);
                a.aux_d[(size_t)slot * 2] = make_float4(aux0_d.x, aux0_d.y, aux0_d.z, 0.f);
                a.auxq[i0] = slot * 2;
            }
            const int i1 = tb_append(a.aux_count, want_aux1);
            if (want_aux1) {
                a.aux_o[(size_t)slot * 2 + 1] = make_float4(aux_o.x, aux_o.y, aux_o.z, 0.f);
                a.aux_d[(size_t)slot * 2 + 1] = make_float4(aux1_d.x, aux1_d.y, aux1_d.z, 0.f);
                a.auxq[i1] = slot * 2 + 1;
            }
            if (sc.env_nee) {                      // warp-uniform (scene constant)
                const int i2 = tb_append(a.aux_count, want_aux2);
                if (want_aux2) {
                    a.aux_o[(size_t)a.env_base + slot] = make_float4(aux_o.x, aux_o.y, aux_o.z, 0.f);
                    a.aux_d[(size_t)a.env_base + slot] = make_float4(aux2_d.x, aux2_d.y, aux2_d.z, 0.f);
                    a.auxq[i2] = a.env_base + slot;
                }
            }
        } else {
            const int at = tb_append(a.next_count, alive);
            if (alive) {
                a.ray_o[slot] = make_float4(next_o.x, next_o.y, next_o.z, o4.w);
                a.ray_d[slot] = make_float4(next_d.x, next_d.y, next_d.z, d4.w);
                a.next_q[at] = slot;
            }
        }
        slot_c = slot_n; slot_n = slot_nn; hit_c = hit_n;
    }
}

// Env-lit extension mode (include/terra_b200.h, terra_b200_set_env_lit): the paths of this bounce that
// left the scene add throughput * environment (Terra.c:1053-1057 with the commented-out line restored).
// terra_attribute_eval(&environment_map, &ray.direction, ...): the direction is the lat-long lookup's
// argument, and what a 2-D texture callback would read as its uv.
__global__ void __launch_bounds__(256) k_miss(TbDevScene sc, const int* q, const int* q_count, const float4* ray_d,
                                               const float4* hit, const float4* thr, float4* lo, TbGenArgs gen) {
    const int count = *q_count;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int slot = q ? q[i] : i;
        if (hit[slot].x >= 0.f) continue;
        float4 d4, t4 = make_float4(1.f, 1.f, 1.f, 0.f);
        if (gen.on) {                   // first bounce after a fused-raygen trace (the packet kernel zeroed lo of these paths)
            uint64_t k0; V3 go;
            const V3 gd = tb_gen_ray(gen, slot, &k0, &go);
            d4 = make_float4(gd.x, gd.y, gd.z, 0.f);
        } else {
            d4 = ray_d[slot]; t4 = thr[slot];
        }
        const float4 l4 = lo[slot];
        const V3 d = v3(d4.x, d4.y, d4.z);
        V2 uv; uv.x = d.x; uv.y = d.y;
        const V3 env = tb_attr_eval(sc.env, sc.textures, uv, d);
        const V3 add = vpoint(v3(t4.x, t4.y, t4.z), env);
        lo[slot] = make_float4(l4.x + add.x, l4.y + add.y, l4.z + add.z, 0.f);
    }
}

// ==================================================================================== K6: accumulate + tonemap
// Terra.c:551-567: acc = sum over the pixel's samples in sample order (one thread per pixel, so
// the float sum order is the reference's and the result is deterministic — no atomics).
__global__ void __launch_bounds__(256) k_accumulate(TbBatch b, const float4* lo, float4* acc) {
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < b.n_pixels; p += gridDim.x * blockDim.x) {
        int rx, ry;
        tb_region_pixel(b, b.p0 + p, &rx, &ry);
        const int at = ry * b.w + rx;                      // acc is region-linear, row-major
        float4 s = acc[at];
        for (int k = 0; k < b.n_samples; ++k) {
            const float4 l = lo[tb_slot_of(b, p, k)];
            s.x = s.x + l.x; s.y = s.y + l.y; s.z = s.z + l.z;
        }
        acc[at] = s;
    }
}

__device__ __forceinline__ V3 tb_uncharted2(V3 x) {                          // Terra.c:1815-1828
    const float A = 0.15f, B = 0.5f, C = 0.1f, D = 0.2f, E = 0.02f, F = 0.3f;
    V3 r;
    r.x = ((x.x * (A * x.x + C * B) + D * E) / (x.x * (A * x.x + B) + D * F)) - E / F;
    r.y = ((x.y * (A * x.y + C * B) + D * E) / (x.y * (A * x.y + B) + D * F)) - E / F;
    r.z = ((x.z * (A * x.z + C * B) + D * E) / (x.z * (A * x.z + B) + D * F)) - E / F;
    return r;
}

__device__ __forceinline__ V3 tb_tonemap(V3 color, int op, float inv_gamma) {    // Terra.c:578-627
    switch (op) {
        case kTerraTonemappingOperatorLinear:
            return v3(powf(color.x, inv_gamma), powf(color.y, inv_gamma), powf(color.z, inv_gamma));
        case kTerraTonemappingOperatorReinhard:
            color.x = color.x / (1.f + color.x);
            color.y = color.y / (1.f + color.y);
            color.z = color.z / (1.f + color.z);
            return v3(powf(color.x, inv_gamma), powf(color.y, inv_gamma), powf(color.z, inv_gamma));
        case kTerraTonemappingOperatorFilmic: {
            V3 x = v3(tmaxf(0.f, color.x - 0.004f), tmaxf(0.f, color.y - 0.004f), tmaxf(0.f, color.z - 0.004f));
            color.x = (x.x * (6.2f * x.x + 0.5f)) / (x.x * (6.2f * x.x + 1.7f) + 0.06f);
            color.y = (x.y * (6.2f * x.y + 0.5f)) / (x.y * (6.2f * x.y + 1.7f) + 0.06f);
            color.x = (x.z * (6.2f * x.z + 0.5f)) / (x.z * (6.2f * x.z + 1.7f) + 0.06f);   // sic: .x twice, .z untouched
            return color;
        }
        case kTerraTonemappingOperatorUncharted2: {
            V3 white_scale = tb_uncharted2(v3s(11.2f));
            white_scale = v3(1.f / white_scale.x, 1.f / white_scale.y, 1.f / white_scale.z);
            V3 t = tb_uncharted2(vmul(color, 2.f));
            color = vpoint(t, white_scale);
            return v3(powf(color.x, inv_gamma), powf(color.y, inv_gamma), powf(color.z, inv_gamma));
        }
        default:
            return color;
    }
}

// Terra.c:569-630 for a w x h region: results += (acc, spp); pixels = tonemap(acc / samples * exposure).
struct TbFinalizeArgs {
    const float4* acc;         // region-linear
    TerraRawIntegrationResult* results; size_t results_pitch;   // element pitch; points at the region's first pixel
    float* pixels; size_t pixels_pitch;                         // may be NULL
    int w, h, spp, tonemap;
    float exposure, inv_gamma;
    const TbCallParams* call;  // non-NULL (graph replay of a device-resident call): results / pitch / region origin come from here
};
__global__ void __launch_bounds__(256) k_finalize(TbFinalizeArgs a) {
    if (a.call) {
        const TbCallParams c = *a.call;
        a.results = (TerraRawIntegrationResult*)c.results + (size_t)c.y * (size_t)c.fbw + (size_t)c.x;
        a.results_pitch = (size_t)c.fbw;
    }
    const int n = a.w * a.h;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const int i = p / a.w, j = p - i * a.w;
        TerraRawIntegrationResult* r = a.results + (size_t)i * a.results_pitch + j;
        const float4 s = a.acc[p];
        const V3 acc = v3(s.x + r->acc.x, s.y + r->acc.y, s.z + r->acc.z);
        const int samples = r->samples + a.spp;
        r->acc.x = acc.x; r->acc.y = acc.y; r->acc.z = acc.z;
        r->samples = samples;
        if (a.pixels) {
            V3 color = vmul(vdiv(acc, (float)samples), a.exposure);
            color = tb_tonemap(color, a.tonemap, a.inv_gamma);
            float* px = a.pixels + ((size_t)i * a.pixels_pitch + j) * 3;
            px[0] = color.x; px[1] = color.y; px[2] = color.z;
        }
    }
}
// Multi-device terra_render: every GPU rendered its own sample range of the WHOLE region into its own accumulator;
// GPU g then finalises ITS row band — it sums the band's pixels over all accumulators (its own and, through
// peer-mapped pointers, the others': loads over NVLink), in device order so the result does not depend on which GPU
// finishes first, adds the caller's running results[] and tonemaps. Reduce-scatter and Terra.c:569-630 in one
// kernel: no intermediate buffer, no second pass, and each GPU then copies its band back over its own PCIe link.
#define TB_MAX_DEVICES 16
struct TbPeerFinalizeArgs {
    const float4* acc[TB_MAX_DEVICES];   // region-linear accumulators, one per device
    int n_dev;
    int p0, n_px;                        // the band: region-linear pixel range [p0, p0 + n_px)
    TerraRawIntegrationResult* results;  // band-linear staging of the caller's results[] (read-modify-write)
    float* pixels;                       // band-linear staging of pixels[]
    int spp, tonemap;
    float exposure, inv_gamma;
};
__global__ void __launch_bounds__(256) k_finalize_peers(TbPeerFinalizeArgs a) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n_px; i += gridDim.x * blockDim.x) {
        const int p = a.p0 + i;
        float4 s = a.acc[0][p];
        for (int g = 1; g < a.n_dev; ++g) {
            const float4 o = a.acc[g][p];
            s.x = s.x + o.x; s.y = s.y + o.y; s.z = s.z + o.z;
        }
        TerraRawIntegrationResult* r = a.results + i;
        const V3 acc = v3(s.x + r->acc.x, s.y + r->acc.y, s.z + r->acc.z);
        const int samples = r->samples + a.spp;
        r->acc.x = acc.x; r->acc.y = acc.y; r->acc.z = acc.z;
        r->samples = samples;
        V3 color = vmul(vdiv(acc, (float)samples), a.exposure);
        color = tb_tonemap(color, a.tonemap, a.inv_gamma);
        float* px = a.pixels + (size_t)i * 3;
        px[0] = color.x; px[1] = color.y; px[2] = color.z;
    }
}
__global__ void __launch_bounds__(256) k_resolve(const TerraRawIntegrationResult* results, float* pixels, size_t n,
                                                 int tonemap, float exposure, float inv_gamma) {
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
        const TerraRawIntegrationResult r = results[p];
        V3 color = vmul(vdiv(v3(r.acc.x, r.acc.y, r.acc.z), (float)r.samples), exposure);
        color = tb_tonemap(color, tonemap, inv_gamma);
        pixels[3 * p] = color.x; pixels[3 * p + 1] = color.y; pixels[3 * p + 2] = color.z;
    }
}

// ==================================================================================== BSDF known-answer kernel
__global__ void k_bsdf_kat(TbDevScene sc, int object, size_t n, const float* e123, const float* wo3, const float* n3, float* out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const TbDevMaterial* m = sc.materials + object;
        TbSurface s;
        s.n = v3(n3[3 * i], n3[3 * i + 1], n3[3 * i + 2]);
        s.xf = basis_from_normal(s.n);
        V2 uv = { 0.f, 0.f };
        s.a0 = tb_attr_eval(m->attr[0], sc.textures, uv, s.n);
        s.a1 = tb_attr_eval(m->attr[1], sc.textures, uv, s.n);
        s.a2 = tb_attr_eval(m->attr[2], sc.textures, uv, s.n);
        s.emissive = v3s(0.f);
        s.state = 0.f;
        if (m->preset == TB_PRESET_PHONG) s.state = tb_attr_eval(m->attr[3], sc.textures, uv, s.n).x;
        if (m->preset == TB_PRESET_GLASS) s.state = s.a2.x;
        const V3 wo = v3(wo3[3 * i], wo3[3 * i + 1], wo3[3 * i + 2]);
        const V3 wi = tb_bsdf_sample(m->preset, s, e123[3 * i], e123[3 * i + 1], e123[3 * i + 2], wo);
        const float pdf = tb_bsdf_pdf(m->preset, s, wi, wo);
        const V3 f = tb_bsdf_eval(m->preset, s, wi, wo);
        float* o = out + 8 * i;
        o[0] = wi.x; o[1] = wi.y; o[2] = wi.z; o[3] = pdf; o[4] = f.x; o[5] = f.y; o[6] = f.z; o[7] = s.state;
    }
}

// ==================================================================================== host-side launchers
// counters layout (ints): [0] trace fetch cursor  [1] live paths this bounce  [2] live paths next bounce
// [3] visibility rays  [4..7] per-preset hit counts  [8..9] u64 rays traced (statistics)
// [10..13] u64 nodes fetched, triangles tested by the incoherent-ray launches (counting variants only)
// [14..15] u64 camera rays  [16..19] u64 nodes, triangles of the camera-ray launches
// [32..51] u64 x 10: lane states of k_trace_v6's rounds (counting variants only, TbTraceArgs::round_stats)
#define TB_COUNTER_INTS 64
__global__ void k_counters(int* c, int mode, int value) {
    unsigned long long* rays = (unsigned long long*)(c + 8);
    unsigned long long* cam_rays = (unsigned long long*)(c + 14);
    switch (mode) {
        case 0: c[0] = 0; c[1] = value; c[2] = 0; c[3] = 0; c[4] = c[5] = c[6] = c[7] = 0;            // new wave = its first bounce begins
                *rays += (unsigned long long)value; *cam_rays += (unsigned long long)value; break;
        case 1: c[0] = 0; c[2] = 0; c[3] = 0; c[4] = c[5] = c[6] = c[7] = 0; *rays += (unsigned long long)c[1]; break;   // bounce begins
        case 2: c[0] = 0; *rays += (unsigned long long)c[3]; break;                                  // before the visibility trace
        case 3: c[1] = c[2]; c[0] = 0; c[2] = 0; c[3] = 0; c[4] = c[5] = c[6] = c[7] = 0; *rays += (unsigned long long)c[1]; break;   // bounce ends, the next one begins
        case 4: *rays += (unsigned long long)value; c[0] = 0; break;                                 // stand-alone trace of `value` rays
        case 5: *rays += (unsigned long long)value; *cam_rays += (unsigned long long)value; c[0] = 0; break;   // ... of camera rays
        default: for (int k = 8; k < 20; ++k) c[k] = 0;
                 for (int k = 32; k < 52; ++k) c[k] = 0; break;
    }
}

// Path slots per wave. A wave's bounce launches shrink (C2: 100 % -> 30 % -> 12 % -> 5 % -> 2 % of its paths) and every
// launch ends in a tail of half-empty warps, so larger waves mean fewer, longer launches: 64M / 128M / 256M slots =
// 5947 / 6059 / 6075 Mpaths/s on the C2 frame (profiles/r3n). 128M slots = 10 GB of path state for the simple integrator,
// 24 GB for direct / MIS; a render only reserves what its region x samples need.
static size_t tb_batch_capacity() {
    const char* e = getenv("TERRA_B200_BATCH");
    size_t cap = e ? (size_t)strtoull(e, NULL, 10) : ((size_t)1 << 27);
    if (cap < 1024) cap = 1024;
    if (cap > ((size_t)1 << 28)) cap = (size_t)1 << 28;
    return cap;
}

static TbEventPair* tb_event_pair(TbDevice* d, int coherent) {
    if (d->ev_used == d->ev_pool.size()) {
        TbEventPair e;
        TB_CUDA(cudaEventCreate(&e.a));
        TB_CUDA(cudaEventCreate(&e.b));
        e.coherent = 0;
        d->ev_pool.push_back(e);
    }
    d->ev_pool[d->ev_used].coherent = coherent;
    return &d->ev_pool[d->ev_used++];
}

// Sum the device times of the trace launches recorded since the last drain, per kind of launch (camera rays /
// everything else: they run different kernels with different roofs). Stream must be idle.
static void tb_drain_events(TbDevice* d) {
    for (size_t i = 0; i < d->ev_used; ++i) {
        float ms = 0.f;
        TB_CUDA(cudaEventElapsedTime(&ms, d->ev_pool[i].a, d->ev_pool[i].b));
        d->stats.trace_kernel_ms += ms;
        d->stats.trace_kernel_launches += 1;
        if (d->ev_pool[i].coherent) { d->stats.coherent_kernel_ms += ms; d->stats.coherent_kernel_launches += 1; }
    }
    d->ev_used = 0;
}

// Which closest-hit kernel a launch uses (measured on B200, profiles/): camera rays are coherent — one warp walks
// the tree once for its 32 rays (k_trace_packet); bounce / visibility / API rays are incoherent — per-lane rays in
// warp-synchronous rounds over 32-byte quantised nodes (k_trace_v6). The work counters of SURVEY §8(d) mean the
// per-ray walk: a counting camera launch runs k_trace<COUNT> (a packet would count the union over its 32 rays).
// Earlier structures and the options that lost their A/B live in experiments/ (not built).
enum { TB_K_AUTO = 0, TB_K_WALK, TB_K_V6, TB_K_PACKET, TB_K_V7, TB_K_V8 };
static_assert(TB_K_V6 == TB_K_V6_ID && TB_K_PACKET == TB_K_PACKET_ID && TB_K_V7 == TB_K_V7_ID && TB_K_V8 == TB_K_V8_ID, "kernel ids used by tb_device_create");

template <bool COUNT>
static void tb_launch_trace_kernel(TbDevice* d, cudaStream_t st, TbTraceArgs& c, int kind, const TbGenArgs* gen) {
    const int grid = d->sm_count * TB_TRACE_CTAS_PER_SM;
    const bool exact = d->exact_boxes != 0;
    if (d->d_qnodes) {
        c.qnodes = d->d_qnodes;
        memcpy(c.grid_lo, d->grid_lo, sizeof c.grid_lo);
        memcpy(c.grid_cell, d->grid_cell, sizeof c.grid_cell);
    }
    if (kind == TB_K_PACKET && !COUNT) {
        TbGenArgs g;
        memset(&g, 0, sizeof g);
        if (gen) {
            g = *gen;
            if (exact) k_trace_packet<true, true><<<grid, TB_TRACE_THREADS, 0, st>>>(c, g);
            else       k_trace_packet<false, true><<<grid, TB_TRACE_THREADS, 0, st>>>(c, g);
        } else {
            if (exact) k_trace_packet<true, false><<<grid, TB_TRACE_THREADS, 0, st>>>(c, g);
            else       k_trace_packet<false, false><<<grid, TB_TRACE_THREADS, 0, st>>>(c, g);
        }
        return;
    }
    if (kind == TB_K_PACKET || kind == TB_K_WALK) {
        if (exact) k_trace<COUNT, true><<<grid, TB_TRACE_THREADS, 0, st>>>(c);
        else       k_trace<COUNT, false><<<grid, TB_TRACE_THREADS, 0, st>>>(c);
        return;
    }
    if (kind == TB_K_V7 && d->d_q4nodes && !exact && (!COUNT || d->counting == 2)) {
        // the rounds kernel on the 4-wide tree. Work counters (counting == 1) mean the canonical per-ray walk of the
        // binary tree, as for camera rays: a 4-wide node is not the 64-byte node SURVEY 8(d) counts. counting == 2
        // counts this kernel's own fetches and lane states (terra_b200_round_stats).
        const int forced7 = getenv("TERRA_B200_STACK_SMEM") ? atoi(getenv("TERRA_B200_STACK_SMEM")) : 0;   // A/B: rows of shared stack
        int rows7 = forced7 > 0 ? forced7 : TB_V7_SMEM_DEPTH;
        if (rows7 > TB_V6_MAX_SMEM_DEPTH) rows7 = TB_V6_MAX_SMEM_DEPTH;
        if (rows7 > d->stack_bound4) rows7 = d->stack_bound4 > 1 ? d->stack_bound4 : 1;
        const bool spill7 = rows7 < d->stack_bound4;
        c.stack_smem = rows7;
        c.q4nodes = d->d_q4nodes;
        const size_t smem7 = (size_t)(rows7 + 3) * TB_TRACE_THREADS * 8 + (size_t)TB_LEAN_WORDS * TB_TRACE_THREADS * 4 + (TB_TRACE_THREADS / 32) * 8;
        const int grid7 = d->sm_count * TB_V7_CTAS_PER_SM;
        if (spill7) k_trace_v7<COUNT, true><<<grid7, TB_TRACE_THREADS, smem7, st>>>(c);
        else        k_trace_v7<COUNT, false><<<grid7, TB_TRACE_THREADS, smem7, st>>>(c);
        return;
    }
    if ((kind == TB_K_V7 || kind == TB_K_V8) && COUNT && !exact && d->counting != 2) {
        // work counters = the canonical per-ray walk of the binary tree (k_trace_v8 also fetches the ~5 % of nodes whose
        // stack entry the closest hit had culled: its own fetches are counting == 2)
        k_trace<COUNT, false><<<grid, TB_TRACE_THREADS, 0, st>>>(c);
        return;
    }
    const int forced = getenv("TERRA_B200_STACK_SMEM") ? atoi(getenv("TERRA_B200_STACK_SMEM")) : 0;   // A/B: rows of shared stack
    if (kind != TB_K_V6 && !exact) {
        // TB_K_V8: the rounds kernel on the binary tree, 8 CTAs per SM (also what TB_K_V7 falls back to without a 4-wide
        // tree). Its stack entries are 4 bytes, so trees up to 40 levels keep the whole stack in shared memory.
        int rows8 = d->stack_bound > 1 ? d->stack_bound : 1;
        const int cap8 = forced > 0 ? forced : TB_V8_SMEM_DEPTH;
        if (rows8 > cap8) rows8 = cap8;
        const bool spill8 = rows8 < d->stack_bound;
        c.stack_smem = rows8;
        const size_t smem8 = (size_t)(rows8 + 3) * TB_TRACE_THREADS * 4 + (size_t)TB_LEAN_WORDS * TB_TRACE_THREADS * 4 + (TB_TRACE_THREADS / 32) * 8;
        const int grid8 = d->sm_count * TB_V8_CTAS_PER_SM;
        if (spill8) k_trace_v8<COUNT, true><<<grid8, TB_TRACE_THREADS, smem8, st>>>(c);
        else        k_trace_v8<COUNT, false><<<grid8, TB_TRACE_THREADS, smem8, st>>>(c);
        return;
    }
    // TB_K_V6: round 2's first rounds kernel — the one with the bit-exact float boxes (exact_boxes) and the A/B baseline
    int rows = d->stack_bound > 1 ? d->stack_bound : 1;
    if (forced > 0 && forced < rows) rows = forced;
    const bool spill = rows > TB_V6_MAX_SMEM_DEPTH || rows < d->stack_bound;
    if (rows > TB_V6_MAX_SMEM_DEPTH) rows = TB_V6_SPILL_SMEM_DEPTH;
    c.stack_smem = rows;
    const size_t smem = (size_t)(rows + 3) * TB_TRACE_THREADS * 8;
    const int grid6 = d->sm_count * TB_V6_CTAS_PER_SM;
    if (exact) {
        if (spill) k_trace_v6<COUNT, true, true><<<grid6, TB_TRACE_THREADS, smem, st>>>(c);
        else       k_trace_v6<COUNT, true, false><<<grid6, TB_TRACE_THREADS, smem, st>>>(c);
    } else {
        if (spill) k_trace_v6<COUNT, false, true><<<grid6, TB_TRACE_THREADS, smem, st>>>(c);
        else       k_trace_v6<COUNT, false, false><<<grid6, TB_TRACE_THREADS, smem, st>>>(c);
    }
}

// Kernel a launch of this kind will run (TB_K_*), after the A/B overrides.
static int tb_trace_kind(const TbDevice* d, bool coherent) {
    int kind = d->trace_kind;
    if (kind == TB_K_AUTO) kind = coherent ? d->coherent_kind : d->incoherent_kind;
    // incoherent rays: the 4-wide tree pays when the geometry is several times L2 and most node fetches are DRAM round
    // trips (v7 against v8 at 1M / 3M / 10M triangles: -14 % / -5 % / +29 %, profiles/r3e, r3f)
    if (kind == TB_K_AUTO && !coherent)
        kind = d->d_q4nodes && (size_t)d->sc.n_nodes * 32 + (size_t)d->sc.n_tris * 48 >= d->v7_min_bytes ? TB_K_V7 : TB_K_V8;
    if (kind == TB_K_AUTO) kind = TB_K_PACKET;
    return kind;
}

static void tb_launch_trace(TbDevice* d, cudaStream_t st, const TbTraceArgs& a, bool coherent, const TbGenArgs* gen = NULL) {
    TbEventPair* e = d->capturing ? NULL : tb_event_pair(d, coherent ? 1 : 0);     // a captured launch is timed as part of the whole call only
    if (e) TB_CUDA(cudaEventRecord(e->a, st));
    TbTraceArgs c = a;
    c.visit_counters = (unsigned long long*)(d->wave.counters + (coherent ? 16 : 10));
    c.round_stats = coherent ? NULL : (unsigned long long*)(d->wave.counters + 32);
    c.tri_min = d->tri_min; c.refetch_below = d->refetch_below;
    const int kind = tb_trace_kind(d, coherent);
    if (d->counting) tb_launch_trace_kernel<true>(d, st, c, kind, NULL);
    else             tb_launch_trace_kernel<false>(d, st, c, kind, gen);
    if (e) TB_CUDA(cudaEventRecord(e->b, st));
    d->stats.kernel_launches += 1;
}

// Camera rays can be generated inside the trace kernel when that kernel is the packet kernel.
static bool tb_fused_raygen(const TbDevice* d) {
    return d->fused_raygen && !d->counting && tb_trace_kind(d, true) == TB_K_PACKET;
}

typedef void (*TbShadeKernel)(TbShadeArgs);
static TbShadeKernel tb_shade_kernel(int preset, int stage, int mode) {
    static const TbShadeKernel simple[TB_PRESET_COUNT] = {
        k_shade<0, 1, TB_SHADE_SIMPLE>, k_shade<1, 1, TB_SHADE_SIMPLE>, k_shade<2, 1, TB_SHADE_SIMPLE>, k_shade<3, 1, TB_SHADE_SIMPLE> };
    static const TbShadeKernel direct[TB_PRESET_COUNT][2] = {
        { k_shade<0, 0, TB_SHADE_DIRECT>, k_shade<0, 1, TB_SHADE_DIRECT> }, { k_shade<1, 0, TB_SHADE_DIRECT>, k_shade<1, 1, TB_SHADE_DIRECT> },
        { k_shade<2, 0, TB_SHADE_DIRECT>, k_shade<2, 1, TB_SHADE_DIRECT> }, { k_shade<3, 0, TB_SHADE_DIRECT>, k_shade<3, 1, TB_SHADE_DIRECT> } };
    static const TbShadeKernel debug[TB_PRESET_COUNT][2] = {
        { k_shade<0, 0, TB_SHADE_DEBUG>, k_shade<0, 1, TB_SHADE_DEBUG> }, { k_shade<1, 0, TB_SHADE_DEBUG>, k_shade<1, 1, TB_SHADE_DEBUG> },
        { k_shade<2, 0, TB_SHADE_DEBUG>, k_shade<2, 1, TB_SHADE_DEBUG> }, { k_shade<3, 0, TB_SHADE_DEBUG>, k_shade<3, 1, TB_SHADE_DEBUG> } };
    if (mode == TB_SHADE_SIMPLE) return simple[preset];
    return mode == TB_SHADE_DIRECT ? direct[preset][stage] : debug[preset][stage];
}

// All waves of one region: adds `spp` samples per pixel into d_acc (region-linear float4, zeroed by the caller).
// Region origin, framebuffer size, sample base, seed and camera reach the kernels through d->d_call (tb_set_call),
// so the same launches serve a direct call and a CUDA-graph replay.
static void tb_render_region(TbDevice* d, cudaStream_t st, size_t w, size_t h, size_t spp, float4* d_acc) {
    const size_t npx = w * h;
    const size_t cap = tb_batch_capacity();
    const size_t px_per_wave = npx < cap ? npx : cap;
    size_t s_per_wave = npx < cap ? cap / npx : 1;
    // a warp of the camera-ray packet kernel holds `sub` samples of one pixel, a power of two that divides the wave's
    // sample count: keep that count a multiple of 32 (or a power of two below 32). 37 samples per wave instead of 32
    // cost the C2 frame 9 % (r2o: packets of 32 different pixels instead of 32 samples of one)
    if (s_per_wave >= 32) s_per_wave &= ~(size_t)31;
    else while (s_per_wave & (s_per_wave - 1)) s_per_wave &= s_per_wave - 1;
    if (s_per_wave > spp) s_per_wave = spp;
    if (s_per_wave < 1) s_per_wave = 1;
    if (d->wave.capacity < px_per_wave * s_per_wave) d->graph_epoch++;       // the wave buffers move: graphs that baked their addresses are stale
    tb_wave_reserve(&d->wave, px_per_wave * s_per_wave);
    TbWave& wv = d->wave;
    const TbDevScene& sc = d->sc;
    if (sc.integrator < 0 || sc.integrator > kTerraIntegratorDebugMisWeights) {
        terra_log("terra_b200: unknown integrator %d\n", sc.integrator);      // assert(false), Terra.c:1134
        abort();
    }
    const int mode = sc.integrator == kTerraIntegratorSimple ? TB_SHADE_SIMPLE
                   : sc.integrator <= kTerraIntegratorDirectMis ? TB_SHADE_DIRECT : TB_SHADE_DEBUG;
    const bool direct = mode == TB_SHADE_DIRECT || sc.integrator == kTerraIntegratorDebugMisWeights;
    const bool env_nee = sc.env_nee && mode == TB_SHADE_DIRECT;
    // debug views (Terra.c:1140-1337) contribute at bounce 0 only: one trace + one shade per wave
    const int last_bounce = mode == TB_SHADE_DEBUG ? 0 : sc.bounces;
    const bool fused = tb_fused_raygen(d);
    for (size_t p0 = 0; p0 < npx; p0 += px_per_wave) {
        const size_t np = npx - p0 < px_per_wave ? npx - p0 : px_per_wave;
        for (size_t s0 = 0; s0 < spp; s0 += s_per_wave) {
            const size_t ns = spp - s0 < s_per_wave ? spp - s0 : s_per_wave;
            TbGenArgs gen;
            memset(&gen, 0, sizeof gen);
            TbBatch& b = gen.batch;
            b.w = (int)w; b.h = (int)h;
            b.p0 = (int)p0; b.n_pixels = (int)np; b.s0 = (int)s0; b.n_samples = (int)ns;
            b.tiled = (d->tile_pixels && w % 8 == 0 && h % 4 == 0 && np % 32 == 0 && p0 % 32 == 0) ? 1 : 0;
            b.sub = 1;
            while (b.sub < 32 && b.sub < d->warp_samples && ns % (size_t)(b.sub * 2) == 0) b.sub *= 2;
            gen.jitter = sc.jitter; gen.on = 1; gen.call = d->d_call;
            const int n_paths = (int)(np * ns);
            k_counters<<<1, 1, 0, st>>>(wv.counters, 0, n_paths);
            d->stats.kernel_launches += 1;
            if (!fused) {
                TbRaygenArgs rg;
                rg.gen = gen;
                rg.ray_o = wv.ray_o; rg.ray_d = wv.ray_d; rg.thr = wv.thr; rg.lo = wv.lo; rg.queue = wv.queue[0];
                k_raygen<<<d->sm_count * 8, 256, 0, st>>>(rg);
                d->stats.kernel_launches += 1;
            }
            int cur = 0;
            for (int bounce = 0; bounce <= last_bounce; ++bounce) {
                const bool implicit = fused && bounce == 0;          // this bounce's rays exist in no array: they are functions of the slot
                TbGenArgs off = gen;
                off.on = 0;
                TbTraceArgs ta;
                memset(&ta, 0, sizeof ta);
                ta.nodes = sc.nodes; ta.tris = sc.tris;
                ta.ray_o = wv.ray_o; ta.ray_d = wv.ray_d;
                ta.queue = implicit ? NULL : wv.queue[cur]; ta.count_ptr = wv.counters + 1; ta.cursor = wv.counters;
                ta.hit_out = wv.hit; ta.lo = wv.lo;
                ta.matq = wv.matq; ta.matq_count = wv.counters + 4; ta.matq_stride = (int)wv.capacity;
                ta.emit_prim = 0;
                tb_launch_trace(d, st, ta, bounce == 0, implicit ? &gen : NULL);
                if (sc.env_lit && last_bounce == sc.bounces && (!env_nee || bounce == 0)) {
                    k_miss<<<d->sm_count * 4, 256, 0, st>>>(sc, implicit ? NULL : wv.queue[cur], wv.counters + 1, wv.ray_d, wv.hit, wv.thr, wv.lo,
                                                            implicit ? gen : off);
                    d->stats.kernel_launches += 1;
                }
                TbShadeArgs sa;
                sa.sc = sc;
                sa.ray_o = wv.ray_o; sa.ray_d = wv.ray_d; sa.thr = wv.thr; sa.lo = wv.lo; sa.hit = wv.hit;
                sa.aux_o = wv.aux_o; sa.aux_d = wv.aux_d; sa.aux_hit = wv.aux_hit;
                sa.next_q = wv.queue[cur ^ 1]; sa.next_count = wv.counters + 2;
                sa.auxq = wv.auxq; sa.aux_count = wv.counters + 3;
                sa.bounce = bounce;
                sa.env_base = (int)(2 * wv.capacity);
                sa.gen = implicit ? gen : off;
                // enough CTAs for the queue when it is short, a grid-stride loop when it is long
                const int shade_cap = d->sm_count * 8;
                int shade_grid = (int)((np * ns + TB_SHADE_THREADS - 1) / TB_SHADE_THREADS);
                if (shade_grid > shade_cap) shade_grid = shade_cap;
                if (shade_grid < 1) shade_grid = 1;
                if (direct && (sc.n_lights > 0 || env_nee)) {
                    for (int p = 0; p < TB_PRESET_COUNT; ++p) {
                        if (!(sc.preset_mask & (1 << p))) continue;
                        sa.q = wv.matq + (size_t)p * wv.capacity; sa.q_count = wv.counters + 4 + p;
                        tb_shade_kernel(p, 0, mode)<<<shade_grid, TB_SHADE_THREADS, 0, st>>>(sa);
                        d->stats.kernel_launches += 1;
                    }
                    k_counters<<<1, 1, 0, st>>>(wv.counters, 2, 0);
                    TbTraceArgs xa;
                    memset(&xa, 0, sizeof xa);
                    xa.nodes = sc.nodes; xa.tris = sc.tris;
                    xa.ray_o = wv.aux_o; xa.ray_d = wv.aux_d;
                    xa.queue = wv.auxq; xa.count_ptr = wv.counters + 3; xa.cursor = wv.counters;
                    xa.hit_out = wv.aux_hit;
                    xa.emit_prim = 0;
                    tb_launch_trace(d, st, xa, false);
                    d->stats.kernel_launches += 1;
                }
                for (int p = 0; p < TB_PRESET_COUNT; ++p) {
                    if (!(sc.preset_mask & (1 << p))) continue;
                    sa.q = wv.matq + (size_t)p * wv.capacity; sa.q_count = wv.counters + 4 + p;
                    tb_shade_kernel(p, 1, mode)<<<shade_grid, TB_SHADE_THREADS, 0, st>>>(sa);
                    d->stats.kernel_launches += 1;
                }
                if (bounce == last_bounce) break;                    // nothing reads the queues of a bounce that does not happen
                k_counters<<<1, 1, 0, st>>>(wv.counters, 3, 0);
                d->stats.kernel_launches += 1;
                cur ^= 1;
            }
            k_accumulate<<<d->sm_count * 4, 256, 0, st>>>(gen.batch, wv.lo, d_acc);
            d->stats.kernel_launches += 1;
            d->stats.paths_traced += (uint64_t)n_paths;
        }
    }
    TB_CUDA(cudaGetLastError());
}

// The per-call values of a render call, written to device memory before its kernels (direct launches and graph
// replays alike). h_call is page-locked and reused: the calls are synchronous, so the copy has run before the next one
// (the chunks of a fanned-out call are not: each takes its own slot of h_call).
static void tb_set_call(TbDevice* d, cudaStream_t st, const TbCameraFrame* cam, size_t fbw, size_t fbh, size_t x, size_t y,
                        size_t sample_base, uint64_t seed, void* d_results, int slot = 0) {
    TbCallParams* hp = d->h_call + slot;
    memset(hp, 0, sizeof *hp);
    hp->cam = *cam;
    hp->x = (int)x; hp->y = (int)y; hp->fbw = (int)fbw; hp->fbh = (int)fbh;
    hp->sample_base = (uint32_t)sample_base; hp->seed = seed; hp->results = d_results;
    TB_CUDA(cudaMemcpyAsync(d->d_call, hp, sizeof *hp, cudaMemcpyHostToDevice, st));
}

// ==================================================================================== entry points
static void tb_finish_call(TbDevice* d, cudaStream_t st, bool whole_render) {
    if (whole_render) TB_CUDA(cudaEventRecord(d->ev_call_b, st));
    TB_CUDA(cudaStreamSynchronize(st));
    TB_CUDA(cudaGetLastError());
    if (whole_render) {
        float ms = 0.f;
        TB_CUDA(cudaEventElapsedTime(&ms, d->ev_call_a, d->ev_call_b));
        d->stats.render_ms += ms;
    }
    tb_drain_events(d);
}

// Graph of one terra_render call of shape (w, h, spp) on the current scene. The first call of a shape runs
// directly (a one-off call must not pay for capture + instantiation); the second captures; later ones replay.
// Everything a graph bakes in — buffer addresses, scene constants, kernel choices — is covered by graph_epoch.
template <typename Body>
static TbDevice::Graph* tb_graph_for(TbDevice* d, cudaStream_t st, size_t w, size_t h, size_t spp, int kind, Body& body) {   // kind: 0 terra_render, 1 device-resident
    ++d->graph_clock;
    for (TbDevice::Graph& g : d->graphs)
        if (g.w == w && g.h == h && g.spp == spp && g.kind == kind && g.epoch == d->graph_epoch) { g.last_use = d->graph_clock; return &g; }
    const uint64_t tag = ((uint64_t)w | ((uint64_t)h << 20) | ((uint64_t)spp << 40) | ((uint64_t)kind << 63)) ^ (d->graph_epoch * 0x9E3779B97F4A7C15ull);
    bool seen = false;
    for (uint64_t t : d->graph_seen) seen = seen || t == tag;
    if (!seen) {
        if (d->graph_seen.size() >= 64) d->graph_seen.clear();
        d->graph_seen.push_back(tag);
        return NULL;
    }
    // drop stale graphs, and the least recently used one beyond 8
    for (size_t i = 0; i < d->graphs.size();) {
        if (d->graphs[i].epoch != d->graph_epoch) { cudaGraphExecDestroy(d->graphs[i].exec); d->graphs.erase(d->graphs.begin() + i); }
        else ++i;
    }
    if (d->graphs.size() >= 8) {
        size_t lru = 0;
        for (size_t i = 1; i < d->graphs.size(); ++i) if (d->graphs[i].last_use < d->graphs[lru].last_use) lru = i;
        cudaGraphExecDestroy(d->graphs[lru].exec);
        d->graphs.erase(d->graphs.begin() + lru);
    }
    // buffers must exist before the capture starts (no allocation inside it): size the wave for this shape
    {
        const size_t npx = w * h, cap = tb_batch_capacity();
        const size_t px_per_wave = npx < cap ? npx : cap;
        size_t s_per_wave = npx < cap ? cap / npx : 1;
        if (s_per_wave > spp) s_per_wave = spp;
        if (s_per_wave < 1) s_per_wave = 1;
        if (d->wave.capacity < px_per_wave * s_per_wave) { d->graph_epoch++; tb_wave_reserve(&d->wave, px_per_wave * s_per_wave); }
    }
    const TerraB200Stats before = d->stats;
    cudaGraph_t graph = NULL;
    d->capturing = true;
    TB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    body();
    TB_CUDA(cudaStreamEndCapture(st, &graph));
    d->capturing = false;
    TbDevice::Graph g;
    g.w = w; g.h = h; g.spp = spp; g.kind = kind; g.epoch = d->graph_epoch; g.last_use = d->graph_clock;
    g.launches = d->stats.kernel_launches - before.kernel_launches;
    g.paths = d->stats.paths_traced - before.paths_traced;
    d->stats = before;                                     // the capture pass executed nothing
    TB_CUDA(cudaGraphInstantiate(&g.exec, graph, 0));
    TB_CUDA(cudaGraphDestroy(graph));
    d->graphs.push_back(g);
    return &d->graphs.back();
}

// d_acc is shared by both render entry points: one capacity, never shrunk.
static void tb_acc_reserve(TbDevice* d, size_t npx) {
    if (npx <= d->acc_capacity) return;
    d->graph_epoch++;
    tb_free(d->d_acc);
    d->d_acc = tb_alloc<float4>(npx);
    d->acc_capacity = npx;
}

// A caller that passes no stream works on the legacy default stream: order this library's own (non-blocking)
// stream after whatever the caller has queued there, so that e.g. a memset of the results array is seen.
static cudaStream_t tb_call_stream(TbDevice* d, void* stream) {
    if (stream) return (cudaStream_t)stream;
    TB_CUDA(cudaEventRecord(d->ev_legacy, cudaStreamLegacy));
    TB_CUDA(cudaStreamWaitEvent(d->stream, d->ev_legacy, 0));
    return d->stream;
}

extern "C" void tb_device_render(TbDevice* d, const TbCameraFrame* cam, const TerraFramebuffer* fb,
                                 size_t x, size_t y, size_t w, size_t h, size_t spp, uint64_t seed) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = d->stream;
    const size_t npx = w * h;
    if (npx > d->fb_capacity) {
        d->graph_epoch++;
        tb_free(d->d_results); tb_free(d->d_pixels);
        d->d_results = tb_alloc<TerraRawIntegrationResult>(npx);
        d->d_pixels = tb_alloc<float>(npx * 3);
        d->fb_capacity = npx;
    }
    tb_acc_reserve(d, npx);
    TB_CUDA(cudaEventRecord(d->ev_call_a, st));
    // progressive accumulation: the caller's results[] are the state (Terra.c:570-572), so they go up first
    TB_CUDA(cudaMemcpy2DAsync(d->d_results, w * sizeof(TerraRawIntegrationResult),
                              fb->results + y * fb->width + x, fb->width * sizeof(TerraRawIntegrationResult),
                              w * sizeof(TerraRawIntegrationResult), h, cudaMemcpyHostToDevice, st));
    d->stats.h2d_bytes += npx * sizeof(TerraRawIntegrationResult);
    // ---- the kernels of the call: launched one by one, or — for a small call whose shape has been seen
    //      before — replayed from a CUDA graph (a 128x128 tile is ~45 launches of a few microseconds each)
    auto body = [&]() {
        TB_CUDA(cudaMemsetAsync(d->d_acc, 0, npx * sizeof(float4), st));
        tb_render_region(d, st, w, h, spp, d->d_acc);
        TbFinalizeArgs fa;
        fa.acc = d->d_acc;
        fa.results = d->d_results; fa.results_pitch = w;
        fa.pixels = d->d_pixels; fa.pixels_pitch = w;
        fa.w = (int)w; fa.h = (int)h; fa.spp = (int)spp; fa.tonemap = d->sc.tonemap;
        fa.exposure = d->sc.exposure; fa.inv_gamma = d->sc.inv_gamma;
        fa.call = NULL;
        k_finalize<<<d->sm_count * 4, 256, 0, st>>>(fa);
        d->stats.kernel_launches += 1;
    };
    TbDevice::Graph* g = NULL;
    if (d->graph_max_paths && npx * spp <= d->graph_max_paths && !d->counting) g = tb_graph_for(d, st, w, h, spp, 0, body);
    tb_set_call(d, st, cam, fb->width, fb->height, x, y, 0, seed, NULL);
    if (g) {
        TB_CUDA(cudaGraphLaunch(g->exec, st));
        d->stats.kernel_launches += g->launches;
        d->stats.paths_traced += g->paths;
        d->stats.graph_replays += 1;
    } else {
        body();
    }
    TB_CUDA(cudaMemcpy2DAsync(fb->results + y * fb->width + x, fb->width * sizeof(TerraRawIntegrationResult),
                              d->d_results, w * sizeof(TerraRawIntegrationResult),
                              w * sizeof(TerraRawIntegrationResult), h, cudaMemcpyDeviceToHost, st));
    TB_CUDA(cudaMemcpy2DAsync(fb->pixels + y * fb->width + x, fb->width * sizeof(TerraFloat3),
                              d->d_pixels, w * sizeof(TerraFloat3), w * sizeof(TerraFloat3), h, cudaMemcpyDeviceToHost, st));
    d->stats.d2h_bytes += npx * (sizeof(TerraRawIntegrationResult) + sizeof(TerraFloat3));
    tb_finish_call(d, st, true);
}

extern "C" void tb_device_render_resident(TbDevice* d, const TbCameraFrame* cam, size_t fbw, size_t fbh,
                                          size_t x, size_t y, size_t w, size_t h, size_t sample_base, size_t spp,
                                          uint64_t seed, void* d_results, void* stream) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = tb_call_stream(d, stream);
    const size_t npx = w * h;
    tb_acc_reserve(d, npx);
    TB_CUDA(cudaEventRecord(d->ev_call_a, st));
    auto body = [&]() {
        TB_CUDA(cudaMemsetAsync(d->d_acc, 0, npx * sizeof(float4), st));
        tb_render_region(d, st, w, h, spp, d->d_acc);
        TbFinalizeArgs fa;
        fa.acc = d->d_acc;
        fa.results = NULL; fa.results_pitch = 0;              // region origin, pitch and the results pointer come from d_call
        fa.pixels = NULL; fa.pixels_pitch = 0;
        fa.w = (int)w; fa.h = (int)h; fa.spp = (int)spp; fa.tonemap = 0; fa.exposure = 1.f; fa.inv_gamma = 1.f;
        fa.call = d->d_call;
        k_finalize<<<d->sm_count * 4, 256, 0, st>>>(fa);
        d->stats.kernel_launches += 1;
    };
    TbDevice::Graph* g = NULL;
    if (d->graph_max_paths && npx * spp <= d->graph_max_paths && !d->counting) g = tb_graph_for(d, st, w, h, spp, 1, body);
    tb_set_call(d, st, cam, fbw, fbh, x, y, sample_base, seed, d_results);
    if (g) {
        TB_CUDA(cudaGraphLaunch(g->exec, st));
        d->stats.kernel_launches += g->launches;
        d->stats.paths_traced += g->paths;
        d->stats.graph_replays += 1;
    } else {
        body();
    }
    tb_finish_call(d, st, true);
}

// The copy-in and copy-out streams beside the compute stream (host-buffer ray batches, fanned-out terra_render), made on first use.
static void tb_pipe_streams(TbDevice* d) {
    if (d->st_in) return;
    {
        TB_CUDA(cudaStreamCreateWithFlags(&d->st_in, cudaStreamNonBlocking));
        TB_CUDA(cudaStreamCreateWithFlags(&d->st_out, cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) {
            TB_CUDA(cudaEventCreateWithFlags(&d->ev_in[k], cudaEventDisableTiming));
            TB_CUDA(cudaEventCreateWithFlags(&d->ev_pack[k], cudaEventDisableTiming));
            TB_CUDA(cudaEventCreateWithFlags(&d->ev_trace[k], cudaEventDisableTiming));
            TB_CUDA(cudaEventCreateWithFlags(&d->ev_out[k], cudaEventDisableTiming));
        }
    }
}

// ---- multi-device terra_render (tb_host.cpp fans a large call out over the scene's devices) -----------------
// The region is cut into row CHUNKS; inside a chunk, device g owns SLICE g of the rows. Per chunk:
// 1. tb_device_partial_begin on every device: upload the slice of the caller's results[] (copy-in stream), render the
//    device's sample range of the CHUNK into its accumulator (compute stream), record an event. Nothing blocks.
// 2. tb_device_partial_finish on every device: on the copy-out stream, wait for ALL devices' events, finalise the slice
//    from all accumulators (k_finalize_peers) and copy it back to the caller's framebuffer — while the compute stream
//    is already rendering the next chunk. Only the last chunk's transfers are exposed (r2n: one chunk per call left
//    0.85 ms of PCIe time behind 12.3 ms of rendering on C2 at N = 8).
// 3. tb_device_partial_wait on every device, once per call.
#define TB_CALL_SLOTS 8      /* page-locked TbCallParams of the chunks of one call in flight (tb_set_call) */
extern "C" void tb_device_partial_begin(TbDevice* d, const TbCameraFrame* cam, const TerraFramebuffer* fb, size_t x, size_t y, size_t w, size_t h,
                                        size_t chunk, size_t chunk_row0, size_t chunk_rows, size_t slice_row0, size_t slice_rows,
                                        size_t stage_px, size_t sample_base, size_t spp, uint64_t seed) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = d->stream;
    const size_t npx = w * h, slice_px = slice_rows * w;
    if (chunk == 0) {
        if (stage_px > d->fb_capacity) {
            d->graph_epoch++;
            tb_free(d->d_results); tb_free(d->d_pixels);
            d->d_results = tb_alloc<TerraRawIntegrationResult>(stage_px);
            d->d_pixels = tb_alloc<float>(stage_px * 3);
            d->fb_capacity = stage_px;
        }
        tb_acc_reserve(d, npx);
        tb_pipe_streams(d);
        d->stage_off = 0;
        TB_CUDA(cudaEventRecord(d->ev_call_a, st));
        TB_CUDA(cudaStreamWaitEvent(d->st_in, d->ev_call_a, 0));          // the staging buffers may just have been reallocated on st's clock
    }
    if (slice_px) {
        TB_CUDA(cudaMemcpy2DAsync(d->d_results + d->stage_off, w * sizeof(TerraRawIntegrationResult),
                                  fb->results + (y + slice_row0) * fb->width + x, fb->width * sizeof(TerraRawIntegrationResult),
                                  w * sizeof(TerraRawIntegrationResult), slice_rows, cudaMemcpyHostToDevice, d->st_in));
        d->stats.h2d_bytes += slice_px * sizeof(TerraRawIntegrationResult);
    }
    TB_CUDA(cudaEventRecord(d->ev_slice_in, d->st_in));
    float4* acc = d->d_acc + chunk_row0 * w;
    TB_CUDA(cudaMemsetAsync(acc, 0, chunk_rows * w * sizeof(float4), st));
    tb_set_call(d, st, cam, fb->width, fb->height, x, y + chunk_row0, sample_base, seed, NULL, (int)(chunk % TB_CALL_SLOTS));
    if (spp && chunk_rows) tb_render_region(d, st, w, chunk_rows, spp, acc);
    TB_CUDA(cudaEventRecord(d->ev_partial, st));
}

extern "C" void tb_device_partial_finish(TbDevice* d, TbDevice* const* devs, int n_dev, const TerraFramebuffer* fb, size_t x, size_t y, size_t w, size_t h,
                                         size_t slice_row0, size_t slice_rows, size_t spp_total, int last) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t so = d->st_out;
    for (int g = 0; g < n_dev; ++g) TB_CUDA(cudaStreamWaitEvent(so, devs[g]->ev_partial, 0));
    TB_CUDA(cudaStreamWaitEvent(so, d->ev_slice_in, 0));
    if (slice_rows) {
        TbPeerFinalizeArgs fa;
        memset(&fa, 0, sizeof fa);
        for (int g = 0; g < n_dev; ++g) fa.acc[g] = devs[g]->d_acc;
        fa.n_dev = n_dev;
        fa.p0 = (int)(slice_row0 * w); fa.n_px = (int)(slice_rows * w);
        fa.results = d->d_results + d->stage_off; fa.pixels = d->d_pixels + d->stage_off * 3;
        fa.spp = (int)spp_total; fa.tonemap = d->sc.tonemap; fa.exposure = d->sc.exposure; fa.inv_gamma = d->sc.inv_gamma;
        k_finalize_peers<<<d->sm_count * 4, 256, 0, so>>>(fa);
        d->stats.kernel_launches += 1;
        TB_CUDA(cudaMemcpy2DAsync(fb->results + (y + slice_row0) * fb->width + x, fb->width * sizeof(TerraRawIntegrationResult),
                                  d->d_results + d->stage_off, w * sizeof(TerraRawIntegrationResult),
                                  w * sizeof(TerraRawIntegrationResult), slice_rows, cudaMemcpyDeviceToHost, so));
        TB_CUDA(cudaMemcpy2DAsync(fb->pixels + (y + slice_row0) * fb->width + x, fb->width * sizeof(TerraFloat3),
                                  d->d_pixels + d->stage_off * 3, w * sizeof(TerraFloat3), w * sizeof(TerraFloat3), slice_rows, cudaMemcpyDeviceToHost, so));
        d->stats.d2h_bytes += slice_rows * w * (sizeof(TerraRawIntegrationResult) + sizeof(TerraFloat3));
        d->stage_off += slice_rows * w;
    }
    if (last) {
        // the other devices read this device's accumulator: the next call must not reset it before they are done
        TB_CUDA(cudaEventRecord(d->ev_band, so));
        TB_CUDA(cudaStreamWaitEvent(d->stream, d->ev_band, 0));      // the call ends on the compute stream (tb_finish_call)
    }
}

extern "C" void tb_device_partial_wait(TbDevice* d, TbDevice* const* devs, int n_dev) {
    TB_CUDA(cudaSetDevice(d->index));
    tb_finish_call(d, d->stream, true);
    for (int g = 0; g < n_dev; ++g) if (devs[g] != d) TB_CUDA(cudaEventSynchronize(devs[g]->ev_band));
}

// Peer access between every pair of the scene's devices (NVLink / NVSwitch on an HGX box). Returns 0 when some pair
// cannot map the other's memory: the caller then keeps to one device.
extern "C" int tb_device_enable_peers(TbDevice* const* devs, int n_dev) {
    for (int a = 0; a < n_dev; ++a)
        for (int b = 0; b < n_dev; ++b) {
            if (a == b) continue;
            int can = 0;
            TB_CUDA(cudaDeviceCanAccessPeer(&can, devs[a]->index, devs[b]->index));
            if (!can) return 0;
        }
    for (int a = 0; a < n_dev; ++a) {
        TB_CUDA(cudaSetDevice(devs[a]->index));
        for (int b = 0; b < n_dev; ++b) {
            if (a == b) continue;
            cudaError_t e = cudaDeviceEnablePeerAccess(devs[b]->index, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) TB_CUDA(e);
            cudaGetLastError();
        }
    }
    return 1;
}

// Replica of `src`'s geometry on `dst`, copied device to device (NVLink) instead of a second trip over PCIe:
// a 10M-triangle scene is 1.8 GB per GPU. Tables (materials, lights, textures) are small and hold device pointers:
// they go up from the host per device (tb_device_upload_tables).
extern "C" void tb_device_clone_geometry(TbDevice* dst, const TbDevice* src) {
    TB_CUDA(cudaSetDevice(src->index));
    TB_CUDA(cudaStreamSynchronize(src->stream));
    TB_CUDA(cudaSetDevice(dst->index));
    TB_CUDA(cudaStreamSynchronize(dst->stream));
    dst->graph_epoch++;
    tb_free(dst->d_nodes); tb_free(dst->d_tris); tb_free(dst->d_props); tb_free(dst->d_qnodes);
    dst->d_qnodes = NULL;
    const size_t n_nodes = (size_t)src->sc.n_nodes, n_tris = (size_t)src->sc.n_tris;
    dst->d_nodes = tb_alloc<float4>(n_nodes * 4);
    dst->d_tris = tb_alloc<float4>(n_tris * 3);
    dst->d_props = tb_alloc<float4>(n_tris * 4);
    TB_CUDA(cudaMemcpyPeerAsync(dst->d_nodes, dst->index, src->d_nodes, src->index, n_nodes * 64, dst->stream));
    TB_CUDA(cudaMemcpyPeerAsync(dst->d_tris, dst->index, src->d_tris, src->index, n_tris * 48, dst->stream));
    TB_CUDA(cudaMemcpyPeerAsync(dst->d_props, dst->index, src->d_props, src->index, n_tris * 64, dst->stream));
    if (src->d_qnodes) {
        dst->d_qnodes = tb_alloc<uint4>(n_nodes * 2);
        TB_CUDA(cudaMemcpyPeerAsync(dst->d_qnodes, dst->index, src->d_qnodes, src->index, n_nodes * 32, dst->stream));
        memcpy(dst->grid_lo, src->grid_lo, sizeof dst->grid_lo);
        memcpy(dst->grid_cell, src->grid_cell, sizeof dst->grid_cell);
    }
    tb_free(dst->d_q4nodes); tb_free(dst->d_q4src);
    dst->d_q4nodes = NULL; dst->d_q4src = NULL; dst->n_q4 = 0; dst->stack_bound4 = 0;
    if (src->d_q4nodes) {
        dst->d_q4nodes = tb_alloc<uint4>(src->n_q4 * 4);
        dst->d_q4src = tb_alloc<int>(src->n_q4 * 4);
        TB_CUDA(cudaMemcpyPeerAsync(dst->d_q4nodes, dst->index, src->d_q4nodes, src->index, src->n_q4 * 64, dst->stream));
        TB_CUDA(cudaMemcpyPeerAsync(dst->d_q4src, dst->index, src->d_q4src, src->index, src->n_q4 * 16, dst->stream));
        dst->n_q4 = src->n_q4; dst->stack_bound4 = src->stack_bound4;
        dst->stats.p2p_bytes += src->n_q4 * 80;
    }
    TB_CUDA(cudaStreamSynchronize(dst->stream));
    dst->stats.p2p_bytes += n_nodes * (64 + (src->d_qnodes ? 32 : 0)) + n_tris * 112;
    dst->stack_bound = src->stack_bound;
    dst->exact_boxes = src->exact_boxes;
    TbDevScene& sc = dst->sc;
    sc.nodes = dst->d_nodes; sc.tris = dst->d_tris; sc.props = dst->d_props;
    sc.n_nodes = src->sc.n_nodes; sc.n_tris = src->sc.n_tris;
}

extern "C" void tb_device_resolve(TbDevice* d, size_t fbw, size_t fbh, const void* d_results, void* d_pixels, void* stream) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = tb_call_stream(d, stream);
    k_resolve<<<d->sm_count * 4, 256, 0, st>>>((const TerraRawIntegrationResult*)d_results, (float*)d_pixels, fbw * fbh,
                                               d->sc.tonemap, d->sc.exposure, d->sc.inv_gamma);
    d->stats.kernel_launches += 1;
    TB_CUDA(cudaStreamSynchronize(st));
}

static void tb_trace_generic(TbDevice* d, cudaStream_t st, size_t n, const float4* d_o, const float4* d_d, float4* d_hit, bool coherent = false) {
    k_counters<<<1, 1, 0, st>>>(d->wave.counters, coherent ? 5 : 4, (int)n);
    TbTraceArgs ta;
    memset(&ta, 0, sizeof ta);
    ta.nodes = d->sc.nodes; ta.tris = d->sc.tris;
    ta.ray_o = d_o; ta.ray_d = d_d;
    ta.count = (int)n; ta.cursor = d->wave.counters;
    ta.hit_out = d_hit;
    ta.emit_prim = 1;
    tb_launch_trace(d, st, ta, coherent);
    d->stats.kernel_launches += 1;
}

static void tb_io_reserve(TbDevice* d, size_t n) {
    if (n <= d->io_capacity) return;
    tb_free(d->d_io_o); tb_free(d->d_io_d); tb_free(d->d_io_hit);
    d->d_io_o = tb_alloc<float4>(n);
    d->d_io_d = tb_alloc<float4>(n);
    d->d_io_hit = tb_alloc<TerraB200Hit>(n);
    d->io_capacity = n;
}

extern "C" void tb_device_trace_resident(TbDevice* d, size_t n, const void* d_o, const void* d_d, void* d_hits, void* stream) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = tb_call_stream(d, stream);
    const size_t chunk = (size_t)1 << 30;
    for (size_t at = 0; at < n; at += chunk) {
        const size_t m = n - at < chunk ? n - at : chunk;
        tb_trace_generic(d, st, m, (const float4*)d_o + at, (const float4*)d_d + at, (float4*)d_hits + at);
    }
    tb_finish_call(d, st, false);
}

// terra_b200_trace_rays_packed_noalloc: closest hits of a DEVICE batch of packed xyz rays (the layout terra_b200_trace_rays
// takes from the host) without a single allocation — for callers that cannot allocate (CUDA graph capture, serving loops).
// k_pack_rays_noalloc is k_pack_rays writing into g_noalloc_scratch (it names the array directly: device code may) and
// clearing the fetch cursor the trace kernel starts from; the trace kernel gets the array through the address
// tb_device_create took with cudaGetSymbolAddress. No cudaMalloc / cudaFree / event creation on this path: it does not
// go through tb_launch_trace (per-launch timing events) and touches none of the growable buffers.
extern "C" int tb_device_trace_packed_noalloc(TbDevice* d, size_t n, const float* d_o3, const float* d_d3, void* d_hits, void* stream) {
    if (n > TB_NOALLOC_MAX_RAYS) {
        terra_log("terra_b200: terra_b200_trace_rays_packed_noalloc: %zu rays exceed the static scratch (TERRA_B200_NOALLOC_MAX_RAYS = %d); "
                  "split the batch or use terra_b200_trace_rays_device\n", n, TB_NOALLOC_MAX_RAYS);
        return -1;
    }
    if (n == 0) return 0;
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = tb_call_stream(d, stream);
    float4* ray_o = (float4*)(d->noalloc_scratch + TB_NOALLOC_HEADER);
    float4* ray_d = ray_o + TB_NOALLOC_MAX_RAYS;
    int grid = (int)((n + 255) / 256);
    if (grid > d->sm_count * 8) grid = d->sm_count * 8;
    k_pack_rays_noalloc<<<grid, 256, 0, st>>>(n, d_o3, d_d3);
    TbTraceArgs ta;
    memset(&ta, 0, sizeof ta);
    ta.nodes = d->sc.nodes; ta.tris = d->sc.tris;
    ta.ray_o = ray_o; ta.ray_d = ray_d;
    ta.count = (int)n; ta.cursor = (int*)d->noalloc_scratch;
    ta.hit_out = (float4*)d_hits;
    ta.emit_prim = 1;
    ta.tri_min = d->tri_min; ta.refetch_below = d->refetch_below;
    tb_launch_trace_kernel<false>(d, st, ta, tb_trace_kind(d, false), NULL);
    d->stats.kernel_launches += 2;
    tb_finish_call(d, st, false);
    return 0;
}

// Host-buffer ray batches (terra_b200_trace_rays). The batch moves through three streams in chunks:
//   st_in:   H2D of chunk k+1's origins / directions (12 B + 12 B per ray, as the caller holds them)
//   stream:  pack to float4 + closest hit of chunk k
//   st_out:  D2H of chunk k-1's hits (16 B per ray)
// with two slots of staging, so that a batch costs max(copy in, trace, copy out) instead of their sum: PCIe
// is full duplex and the copy engines run beside the SMs. The overlap needs page-locked caller buffers
// (terra_b200_host_alloc); with pageable memory the driver stages every copy and the calls serialise, the
// results are the same. TERRA_B200_IO_CHUNK sets the chunk (rays).
static void tb_pipe_reserve(TbDevice* d, size_t chunk) {
    tb_pipe_streams(d);
    if (chunk <= d->pipe_capacity) return;
    for (int k = 0; k < 2; ++k) {
        tb_free(d->d_pipe_o3[k]); tb_free(d->d_pipe_d3[k]); tb_free(d->d_pipe_hit[k]);
        d->d_pipe_o3[k] = tb_alloc<float>(chunk * 3);
        d->d_pipe_d3[k] = tb_alloc<float>(chunk * 3);
        d->d_pipe_hit[k] = tb_alloc<TerraB200Hit>(chunk);
    }
    d->pipe_capacity = chunk;
}

static bool tb_is_pageable(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return at.type == cudaMemoryTypeUnregistered;
}

// memcpy on several host threads (one thread moves ~10 GB/s, the PCIe link wants 50)
static void tb_memcpy_mt(void* dst, const void* src, size_t bytes) {
    const size_t min_per_thread = (size_t)4 << 20;
    unsigned hw = std::thread::hardware_concurrency();
    size_t threads = hw ? hw : 1;
    if (threads > 16) threads = 16;
    if (threads > bytes / min_per_thread) threads = bytes / min_per_thread;
    if (threads <= 1) { memcpy(dst, src, bytes); return; }
    std::vector<std::thread> pool;
    const size_t per = ((bytes + threads - 1) / threads + 63) & ~(size_t)63;
    for (size_t t = 0; t < threads; ++t) {
        const size_t lo = t * per, hi = lo + per < bytes ? lo + per : bytes;
        if (lo >= hi) break;
        pool.emplace_back([=]() { memcpy((char*)dst + lo, (const char*)src + lo, hi - lo); });
    }
    for (std::thread& th : pool) th.join();
}

static void tb_stage_reserve(TbDevice* d, size_t chunk) {
    if (chunk <= d->stage_capacity) return;
    for (int k = 0; k < 2; ++k) {
        if (d->h_stage_o3[k]) cudaFreeHost(d->h_stage_o3[k]);
        if (d->h_stage_d3[k]) cudaFreeHost(d->h_stage_d3[k]);
        if (d->h_stage_hit[k]) cudaFreeHost(d->h_stage_hit[k]);
        TB_CUDA(cudaHostAlloc((void**)&d->h_stage_o3[k], chunk * 12, cudaHostAllocDefault));
        TB_CUDA(cudaHostAlloc((void**)&d->h_stage_d3[k], chunk * 12, cudaHostAllocDefault));
        TB_CUDA(cudaHostAlloc((void**)&d->h_stage_hit[k], chunk * sizeof(TerraB200Hit), cudaHostAllocDefault));
    }
    d->stage_capacity = chunk;
}

extern "C" void tb_device_trace_host(TbDevice* d, size_t n, const float* origins, const float* dirs, TerraB200Hit* out) {
    TB_CUDA(cudaSetDevice(d->index));
    if (n == 0) return;
    cudaStream_t st = d->stream;
    size_t chunk = (size_t)1 << 21;
    if (const char* e = getenv("TERRA_B200_IO_CHUNK")) { const long long v = atoll(e); if (v >= 1024) chunk = (size_t)v; }
    if (chunk > n) chunk = n;
    tb_io_reserve(d, chunk);
    tb_pipe_reserve(d, chunk);
    // Pageable caller memory: a cudaMemcpyAsync from it is staged by the driver through one small buffer and
    // blocks the calling thread (~6 GB/s, no overlap). Large pageable batches go through the library's own
    // page-locked bounce buffers instead, filled and drained by host threads while the GPU works on the
    // previous chunk. Small batches are not worth the extra copy.
    size_t staging_min = (size_t)1 << 18;                               // TERRA_B200_STAGING_MIN (rays; 0 = never stage)
    if (const char* e = getenv("TERRA_B200_STAGING_MIN")) { staging_min = (size_t)strtoull(e, NULL, 10); if (!staging_min) staging_min = ~(size_t)0; }
    const bool staged = n >= staging_min &&
                        (tb_is_pageable(origins) || tb_is_pageable(dirs) || tb_is_pageable(out));
    if (staged) tb_stage_reserve(d, chunk);
    size_t k = 0;
    size_t pending_at[2] = { 0, 0 }, pending_m[2] = { 0, 0 };           // staged: hits waiting in h_stage_hit[slot]
    for (size_t at = 0; at < n; at += chunk, ++k) {
        const size_t m = n - at < chunk ? n - at : chunk;
        const int slot = (int)(k & 1);
        const float* src_o = origins + 3 * at;
        const float* src_d = dirs + 3 * at;
        TerraB200Hit* dst_hit = out + at;
        if (staged) {
            if (k >= 2) {                                                 // chunk k-2 has come back: hand it to the caller, slot is free
                TB_CUDA(cudaEventSynchronize(d->ev_out[slot]));
                tb_memcpy_mt(out + pending_at[slot], d->h_stage_hit[slot], pending_m[slot] * sizeof(TerraB200Hit));
            }
            tb_memcpy_mt(d->h_stage_o3[slot], src_o, m * 12);
            tb_memcpy_mt(d->h_stage_d3[slot], src_d, m * 12);
            src_o = d->h_stage_o3[slot]; src_d = d->h_stage_d3[slot]; dst_hit = d->h_stage_hit[slot];
            pending_at[slot] = at; pending_m[slot] = m;
        }
        if (k >= 2) TB_CUDA(cudaStreamWaitEvent(d->st_in, d->ev_pack[slot], 0));       // chunk k-2 has been packed: its raw input is free
        TB_CUDA(cudaMemcpyAsync(d->d_pipe_o3[slot], src_o, m * 12, cudaMemcpyHostToDevice, d->st_in));
        TB_CUDA(cudaMemcpyAsync(d->d_pipe_d3[slot], src_d, m * 12, cudaMemcpyHostToDevice, d->st_in));
        TB_CUDA(cudaEventRecord(d->ev_in[slot], d->st_in));
        TB_CUDA(cudaStreamWaitEvent(st, d->ev_in[slot], 0));
        k_pack_rays<<<d->sm_count * 8, 256, 0, st>>>(m, d->d_pipe_o3[slot], d->d_pipe_d3[slot], d->d_io_o, d->d_io_d);
        TB_CUDA(cudaEventRecord(d->ev_pack[slot], st));
        d->stats.kernel_launches += 1;
        if (k >= 2) TB_CUDA(cudaStreamWaitEvent(st, d->ev_out[slot], 0));              // chunk k-2's hits have left the slot
        tb_trace_generic(d, st, m, d->d_io_o, d->d_io_d, (float4*)d->d_pipe_hit[slot]);
        TB_CUDA(cudaEventRecord(d->ev_trace[slot], st));
        TB_CUDA(cudaStreamWaitEvent(d->st_out, d->ev_trace[slot], 0));
        TB_CUDA(cudaMemcpyAsync(dst_hit, d->d_pipe_hit[slot], m * sizeof(TerraB200Hit), cudaMemcpyDeviceToHost, d->st_out));
        TB_CUDA(cudaEventRecord(d->ev_out[slot], d->st_out));
        d->stats.h2d_bytes += m * 24;
        d->stats.d2h_bytes += m * sizeof(TerraB200Hit);
    }
    if (staged) {                                                         // the last one or two chunks, oldest first
        const size_t first = k >= 2 ? k - 2 : 0;
        for (size_t j = first; j < k; ++j) {
            const int slot = (int)(j & 1);
            TB_CUDA(cudaEventSynchronize(d->ev_out[slot]));
            tb_memcpy_mt(out + pending_at[slot], d->h_stage_hit[slot], pending_m[slot] * sizeof(TerraB200Hit));
        }
    }
    TB_CUDA(cudaStreamSynchronize(d->st_out));
    tb_finish_call(d, st, false);
}

extern "C" void tb_device_primary_hits(TbDevice* d, const TbCameraFrame* cam, size_t w, size_t h, TerraB200Hit* out) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = d->stream;
    const size_t n = w * h;
    tb_io_reserve(d, n);
    k_primary_rays<<<d->sm_count * 8, 256, 0, st>>>(*cam, (int)w, (int)h, d->d_io_o, d->d_io_d);
    d->stats.kernel_launches += 1;
    tb_trace_generic(d, st, n, d->d_io_o, d->d_io_d, (float4*)d->d_io_hit, true);      // camera rays
    TB_CUDA(cudaMemcpyAsync(out, d->d_io_hit, n * sizeof(TerraB200Hit), cudaMemcpyDeviceToHost, st));
    d->stats.d2h_bytes += n * sizeof(TerraB200Hit);
    tb_finish_call(d, st, false);
}

extern "C" void tb_device_bsdf_kat(TbDevice* d, size_t object_idx, size_t n, const float* e123, const float* wo,
                                   const float* normal, float* out) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = d->stream;
    float* d_e = tb_alloc<float>(n * 3); float* d_w = tb_alloc<float>(n * 3); float* d_n = tb_alloc<float>(n * 3);
    float* d_out = tb_alloc<float>(n * 8);
    TB_CUDA(cudaMemcpyAsync(d_e, e123, n * 12, cudaMemcpyHostToDevice, st));
    TB_CUDA(cudaMemcpyAsync(d_w, wo, n * 12, cudaMemcpyHostToDevice, st));
    TB_CUDA(cudaMemcpyAsync(d_n, normal, n * 12, cudaMemcpyHostToDevice, st));
    k_bsdf_kat<<<d->sm_count, 256, 0, st>>>(d->sc, (int)object_idx, n, d_e, d_w, d_n, d_out);
    d->stats.kernel_launches += 1;
    TB_CUDA(cudaMemcpyAsync(out, d_out, n * 32, cudaMemcpyDeviceToHost, st));
    tb_finish_call(d, st, false);
    tb_free(d_e); tb_free(d_w); tb_free(d_n); tb_free(d_out);
}

// ==================================================================================== memory-system probes
// Roofline denominators that MEASURED_PEAKS.json does not carry: L2 bandwidth for a
// resident buffer, (mode 0) streaming 128-bit reads, (mode 1) random 64-byte-aligned gathers of one
// node-sized record (4 x 128-bit), the access pattern of k_trace. Returns GB/s.
__global__ void __launch_bounds__(256) k_probe_stream(const float4* buf, size_t n_vec, int reps, float4* sink) {
    float4 acc = make_float4(0, 0, 0, 0);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
            const float4 v = __ldcg(buf + i);      // L2 only: a per-SM slice of a small buffer would otherwise sit in L1
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
    if (acc.x == 12345.678f) sink[0] = acc;
}
__global__ void __launch_bounds__(256) k_probe_gather(const float4* buf, size_t n_rec, int per_thread, float4* sink) {
    float4 acc = make_float4(0, 0, 0, 0);
    uint64_t st = tb_mix64(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 1);
    for (int k = 0; k < per_thread; ++k) {
        st = st * 6364136223846793005ull + 1442695040888963407ull;
        const size_t rec = (size_t)((st >> 33) % n_rec);
        const float4* p = buf + rec * 4;
        const float4 a = __ldcg(p), b = __ldcg(p + 1), c = __ldcg(p + 2), d = __ldcg(p + 3);
        acc.x += a.x + b.x + c.x + d.x; acc.y += a.y + b.y; acc.z += c.z; acc.w += d.w;
    }
    if (acc.x == 12345.678f) sink[0] = acc;
}
extern "C" double tb_device_probe_bandwidth(int device_index, size_t bytes, int mode) {
    TB_CUDA(cudaSetDevice(device_index));
    cudaDeviceProp prop;
    TB_CUDA(cudaGetDeviceProperties(&prop, device_index));
    float4* buf = tb_alloc<float4>(bytes / 16 + 4);
    float4* sink = tb_alloc<float4>(1);
    TB_CUDA(cudaMemset(buf, 0, bytes));
    cudaEvent_t a, b;
    TB_CUDA(cudaEventCreate(&a)); TB_CUDA(cudaEventCreate(&b));
    const int grid = prop.multiProcessorCount * 8;
    double moved = 0;
    float best = 1e30f;
    for (int it = 0; it < 4; ++it) {                // first pass warms L2; keep the best of the rest
        TB_CUDA(cudaEventRecord(a));
        if (mode == 0) {
            const int reps = (int)(((size_t)8 << 30) / bytes) + 1;
            k_probe_stream<<<grid, 256>>>(buf, bytes / 16, reps, sink);
            moved = (double)reps * (double)(bytes / 16) * 16.0;
        } else {
            const int per_thread = 2048;
            k_probe_gather<<<grid, 256>>>(buf, bytes / 64, per_thread, sink);
            moved = (double)grid * 256.0 * per_thread * 64.0;
        }
        TB_CUDA(cudaEventRecord(b));
        TB_CUDA(cudaEventSynchronize(b));
        float ms = 0.f;
        TB_CUDA(cudaEventElapsedTime(&ms, a, b));
        if (it > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    tb_free(buf); tb_free(sink);
    return moved / (best * 1e-3) / 1e9;
}

// ==================================================================================== refit
// Vertex positions (and normals / texcoords) edited in place, same topology: the tree keeps its shape and every
// box is re-fitted on the GPU, bottom-up, then re-quantised — no rebuild, no flatten. The boxes are what a fresh
// sah3 build gives for the same tree (a child box = exact union of its padded triangle boxes, min / max only), so
// closest hits stay bit-identical to a fresh commit of the edited geometry.
//   k_refit_gather    rank-ordered triangle / property records <- the objects' raw arrays (prim id rides in tris[].w)
//   k_refit_parents   parent index + number of inner children per node (once per refit: the tree is immutable)
//   k_refit_fit       one thread per node without inner children starts, the last child to arrive continues upwards
//   k_refit_quantise  the host quantiser's arithmetic (double, outward rounding + one cell) per node
__global__ void __launch_bounds__(256) k_refit_gather(const float* __restrict__ raw_tris, const float* __restrict__ raw_props,
                                                      const uint32_t* __restrict__ obj_base, float4* tris, float4* props, int n_tris) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_tris; r += gridDim.x * blockDim.x) {
        const float4 a = tris[(size_t)r * 3], b = tris[(size_t)r * 3 + 1], c = tris[(size_t)r * 3 + 2];
        const uint32_t prim = __float_as_uint(a.w);
        const size_t src = (size_t)obj_base[prim & 0xffu] + (prim >> 8);
        const float* t = raw_tris + src * 9;
        tris[(size_t)r * 3]     = make_float4(t[0], t[1], t[2], a.w);
        tris[(size_t)r * 3 + 1] = make_float4(t[3], t[4], t[5], b.w);
        tris[(size_t)r * 3 + 2] = make_float4(t[6], t[7], t[8], c.w);
        const float* p = raw_props + src * 15;
        float* q = (float*)(props + (size_t)r * 4);
#pragma unroll
        for (int k = 0; k < 15; ++k) q[k] = p[k];
    }
}
__global__ void __launch_bounds__(256) k_refit_parents(const float4* __restrict__ nodes, int n_nodes, int* parent, int* need, int* arrive) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += gridDim.x * blockDim.x) {
        const float4 q3 = nodes[(size_t)i * 4 + 3];
        const int r0 = __float_as_int(q3.x), r1 = __float_as_int(q3.y);
        need[i] = (r0 >= 0 ? 1 : 0) + (r1 >= 0 ? 1 : 0);
        arrive[i] = 0;
        if (r0 >= 0) parent[r0] = i;
        if (r1 >= 0) parent[r1] = i;
        if (i == 0) parent[0] = -1;
    }
}
__device__ __forceinline__ float tb_pad_lo(float v) { return v - (1e-4f + 2e-6f * fabsf(v)); }   // tb_bvh.cpp pad_lo / pad_hi
__device__ __forceinline__ float tb_pad_hi(float v) { return v + (1e-4f + 2e-6f * fabsf(v)); }
__device__ __forceinline__ void tb_refit_child(const float4* nodes, const float4* tris, int ref, float* lo, float* hi) {
    if (ref < 0) {
        const float4* tp = tris + (size_t)(~ref) * 3;
        const float4 a = __ldcg(tp), b = __ldcg(tp + 1), c = __ldcg(tp + 2);
        const float xs[3][3] = { { a.x, b.x, c.x }, { a.y, b.y, c.y }, { a.z, b.z, c.z } };
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float l = xs[k][0] < xs[k][1] ? xs[k][0] : xs[k][1]; if (xs[k][2] < l) l = xs[k][2];
            float h = xs[k][0] > xs[k][1] ? xs[k][0] : xs[k][1]; if (xs[k][2] > h) h = xs[k][2];
            lo[k] = tb_pad_lo(l); hi[k] = tb_pad_hi(h);
        }
    } else {
        const float4* np = nodes + (size_t)ref * 4;
        const float4 q0 = __ldcg(np), q1 = __ldcg(np + 1), q2 = __ldcg(np + 2);
        lo[0] = fminf(q0.x, q1.z); lo[1] = fminf(q0.y, q1.w); lo[2] = fminf(q0.z, q2.x);
        hi[0] = fmaxf(q0.w, q2.y); hi[1] = fmaxf(q1.x, q2.z); hi[2] = fmaxf(q1.y, q2.w);
    }
}
__global__ void __launch_bounds__(256) k_refit_fit(float4* nodes, const float4* __restrict__ tris, const int* __restrict__ parent,
                                                   const int* __restrict__ need, int* arrive, int n_nodes) {
    const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (i0 >= n_nodes || need[i0] != 0) return;
    int i = i0;
    while (true) {
        const float4 q3 = __ldcg(nodes + (size_t)i * 4 + 3);
        float lo0[3], hi0[3], lo1[3], hi1[3];
        tb_refit_child(nodes, tris, __float_as_int(q3.x), lo0, hi0);
        tb_refit_child(nodes, tris, __float_as_int(q3.y), lo1, hi1);
        __stcg(nodes + (size_t)i * 4,     make_float4(lo0[0], lo0[1], lo0[2], hi0[0]));
        __stcg(nodes + (size_t)i * 4 + 1, make_float4(hi0[1], hi0[2], lo1[0], lo1[1]));
        __stcg(nodes + (size_t)i * 4 + 2, make_float4(lo1[2], hi1[0], hi1[1], hi1[2]));
        __threadfence();
        const int p = parent[i];
        if (p < 0) return;
        if (atomicAdd(&arrive[p], 1) + 1 < need[p]) return;        // the other inner child is still on its way: it continues
        __threadfence();
        i = p;
    }
}
// The 4-wide tree after a refit: every slot's box words are copied again from the binary node they came from.
__global__ void __launch_bounds__(256) k_refit_quad(const uint32_t* __restrict__ qnodes, const int* __restrict__ src, uint32_t* q4, int n_q4) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_q4 * 4; i += gridDim.x * blockDim.x) {
        const int s = src[i];
        if (s < 0) continue;
        const uint32_t* q = qnodes + (size_t)(s >> 1) * 8 + 3 * (s & 1);
        uint32_t* w = q4 + (size_t)(i >> 2) * 16 + 3 * (i & 3);
        w[0] = q[0]; w[1] = q[1]; w[2] = q[2];
    }
}

__global__ void __launch_bounds__(256) k_refit_quantise(const float4* __restrict__ nodes, uint4* qnodes, int n_nodes,
                                                        float l0, float l1, float l2, float c0, float c1, float c2) {
    const double gl[3] = { (double)l0, (double)l1, (double)l2 }, gc[3] = { (double)c0, (double)c1, (double)c2 };
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += gridDim.x * blockDim.x) {
        const float4 q0 = nodes[(size_t)i * 4], q1 = nodes[(size_t)i * 4 + 1], q2 = nodes[(size_t)i * 4 + 2], q3 = nodes[(size_t)i * 4 + 3];
        const float q[12] = { q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w };
        uint32_t g[12];
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double l = gl[k], cell = gc[k];
                long long a = (long long)floor((q[6 * c + k] - l) / cell);
                while (a > 0 && l + a * cell > q[6 * c + k]) --a;
                a -= 1;
                long long b = (long long)ceil((q[6 * c + 3 + k] - l) / cell);
                while (b < 65535 && l + b * cell < q[6 * c + 3 + k]) ++b;
                b += 1;
                if (a < 0) a = 0;
                if (b > 65535) b = 65535;
                g[6 * c + k] = (uint32_t)a; g[6 * c + 3 + k] = (uint32_t)b;
            }
        uint4 w0, w1;
        w0.x = g[0] | (g[3] << 16); w0.y = g[1] | (g[4] << 16); w0.z = g[2] | (g[5] << 16); w0.w = g[6] | (g[9] << 16);
        w1.x = g[7] | (g[10] << 16); w1.y = g[8] | (g[11] << 16); w1.z = __float_as_uint(q3.x); w1.w = __float_as_uint(q3.y);
        qnodes[(size_t)i * 2] = w0; qnodes[(size_t)i * 2 + 1] = w1;
    }
}

// raw_tris: 9 floats per triangle, objects concatenated; raw_props: 15 floats per triangle; obj_base[o] = first
// triangle of object o in those arrays. lights / light_area: the refreshed light table (same sizes as at commit).
extern "C" void tb_device_refit(TbDevice* d, const float* const* obj_tris, const float* const* obj_props, const uint32_t* obj_base, size_t n_objects,
                                const TbDevLight* lights, size_t n_lights, const float* light_area, size_t n_light_tris, float* ms_out) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = d->stream;
    const int n_tris = d->sc.n_tris, n_nodes = d->sc.n_nodes;
    cudaEvent_t e0, e1;
    TB_CUDA(cudaEventCreate(&e0)); TB_CUDA(cudaEventCreate(&e1));
    // the objects' own arrays go up as they are (object o at triangle offset obj_base[o]): no host-side gather
    float* d_raw_t = tb_alloc<float>((size_t)n_tris * 9);
    float* d_raw_p = tb_alloc<float>((size_t)n_tris * 15);
    for (size_t o = 0; o < n_objects; ++o) {
        const size_t cnt = (o + 1 < n_objects ? obj_base[o + 1] : (uint32_t)n_tris) - obj_base[o];
        TB_CUDA(cudaMemcpyAsync(d_raw_t + (size_t)obj_base[o] * 9, obj_tris[o], cnt * 36, cudaMemcpyHostToDevice, st));
        TB_CUDA(cudaMemcpyAsync(d_raw_p + (size_t)obj_base[o] * 15, obj_props[o], cnt * 60, cudaMemcpyHostToDevice, st));
        d->stats.h2d_bytes += cnt * 96;
    }
    uint32_t* d_base = tb_upload(obj_base, n_objects, d);
    int* d_parent = tb_alloc<int>((size_t)n_nodes);
    int* d_need = tb_alloc<int>((size_t)n_nodes);
    int* d_arrive = tb_alloc<int>((size_t)n_nodes);
    TB_CUDA(cudaEventRecord(e0, st));
    k_refit_gather<<<d->sm_count * 8, 256, 0, st>>>(d_raw_t, d_raw_p, d_base, d->d_tris, d->d_props, n_tris);
    k_refit_parents<<<d->sm_count * 8, 256, 0, st>>>(d->d_nodes, n_nodes, d_parent, d_need, d_arrive);
    k_refit_fit<<<(n_nodes + 255) / 256, 256, 0, st>>>(d->d_nodes, d->d_tris, d_parent, d_need, d_arrive, n_nodes);
    d->stats.kernel_launches += 3;
    if (d->d_qnodes) {
        // the grid of the quantised nodes covers the scene box, i.e. the root's two child boxes (tb_bvh_quantize)
        float root[12];
        TB_CUDA(cudaMemcpyAsync(root, d->d_nodes, sizeof root, cudaMemcpyDeviceToHost, st));
        TB_CUDA(cudaStreamSynchronize(st));
        for (int k = 0; k < 3; ++k) {
            const double lo = root[k] < root[6 + k] ? root[k] : root[6 + k];
            const double hi = root[3 + k] > root[9 + k] ? root[3 + k] : root[9 + k];
            double ext = hi - lo;
            if (!(ext > 0)) ext = 1.0;
            const double pad = ext * 1e-3;
            d->grid_lo[k] = (float)(lo - pad);
            d->grid_cell[k] = (float)((ext + 2 * pad) / 65535.0);
        }
        k_refit_quantise<<<d->sm_count * 8, 256, 0, st>>>(d->d_nodes, d->d_qnodes, n_nodes, d->grid_lo[0], d->grid_lo[1], d->grid_lo[2],
                                                          d->grid_cell[0], d->grid_cell[1], d->grid_cell[2]);
        d->stats.kernel_launches += 1;
        if (d->d_q4nodes) {
            k_refit_quad<<<d->sm_count * 8, 256, 0, st>>>((const uint32_t*)d->d_qnodes, d->d_q4src, (uint32_t*)d->d_q4nodes, (int)d->n_q4);
            d->stats.kernel_launches += 1;
        }
    }
    TB_CUDA(cudaEventRecord(e1, st));
    if (n_lights) {
        TB_CUDA(cudaMemcpyAsync(d->d_lights, lights, n_lights * sizeof(TbDevLight), cudaMemcpyHostToDevice, st));
        TB_CUDA(cudaMemcpyAsync(d->d_light_area, light_area, n_light_tris * sizeof(float), cudaMemcpyHostToDevice, st));
    }
    TB_CUDA(cudaStreamSynchronize(st));
    TB_CUDA(cudaGetLastError());
    if (ms_out) TB_CUDA(cudaEventElapsedTime(ms_out, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    tb_free(d_raw_t); tb_free(d_raw_p); tb_free(d_base); tb_free(d_parent); tb_free(d_need); tb_free(d_arrive);
    d->graph_epoch++;                                            // the grid constants are kernel arguments of captured graphs
}

extern "C" void tb_device_get_stats(TbDevice* d, TerraB200Stats* out) {
    TB_CUDA(cudaSetDevice(d->index));
    unsigned long long c[6] = { 0, 0, 0, 0, 0, 0 };          // rays, nodes, tris (incoherent launches); camera rays, nodes, tris
    TB_CUDA(cudaMemcpy(c, d->wave.counters + 8, sizeof c, cudaMemcpyDeviceToHost));
    d->stats.rays_traced = c[0];
    d->stats.nodes_visited = c[1] + c[4];
    d->stats.tris_tested = c[2] + c[5];
    d->stats.coherent_rays = c[3];
    d->stats.coherent_nodes_visited = c[4];
    d->stats.coherent_tris_tested = c[5];
    *out = d->stats;
}

extern "C" void tb_device_round_stats(TbDevice* d, uint64_t* out10) {
    TB_CUDA(cudaSetDevice(d->index));
    TB_CUDA(cudaMemcpy(out10, d->wave.counters + 32, 10 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
}
extern "C" void tb_device_set_counting(TbDevice* d, int on) { d->counting = on; d->graph_epoch++; }
// Name of the closest-hit kernel an incoherent launch (bounce / visibility / API rays) of this device runs right now.
extern "C" const char* tb_device_incoherent_kernel(TbDevice* d) {
    const int kind = tb_trace_kind(d, false);
    if (kind == TB_K_WALK || kind == TB_K_PACKET) return "k_trace";
    if (d->exact_boxes || kind == TB_K_V6) return "k_trace_v6";
    return kind == TB_K_V7 && d->d_q4nodes ? "k_trace_v7" : "k_trace_v8";
}
extern "C" void tb_device_set_graph_paths(TbDevice* d, size_t max_paths) { d->graph_max_paths = max_paths; }
// Builds TbDevDist2D over `f` (ny rows of nx, device) into freshly allocated device arrays.
static void tb_dist2d_build(TbDevice* d, cudaStream_t st, const float* d_f, int nx, int ny, float** cdf_out, float** rows_out,
                            float** marg_out, TbDevDist2D* out) {
    float* cdf = tb_alloc<float>((size_t)nx * ny);
    float* rows = tb_alloc<float>((size_t)ny);
    float* marg = tb_alloc<float>((size_t)ny);
    float* total = tb_alloc<float>(1);
    k_dist_rows<<<(ny + 127) / 128, 128, 0, st>>>(d_f, nx, ny, cdf, rows);            // conditionals, Terra.c:812-814
    k_dist_rows<<<1, 128, 0, st>>>(rows, ny, 1, marg, total);                          // marginal over the row totals, :816-832
    d->stats.kernel_launches += 2;
    float h_total = 0.f;
    TB_CUDA(cudaMemcpyAsync(&h_total, total, sizeof(float), cudaMemcpyDeviceToHost, st));
    TB_CUDA(cudaStreamSynchronize(st));
    tb_free(total);
    out->f = d_f; out->cdf = cdf; out->row_integral = rows; out->marginal_cdf = marg;
    out->integral = h_total; out->nx = nx; out->ny = ny;
    *cdf_out = cdf; *rows_out = rows; *marg_out = marg;
}

// on = 1: a miss adds throughput * environment at every bounce (Terra.c:1056 restored). on = 2: the same
// radiance, but the direct integrators sample a lat-long environment explicitly through the reference's
// distributions (tb_distribution.cuh) and misses count at the camera ray only; falls back to mode 1 when
// the environment is not a lat-long texture.
extern "C" void tb_device_set_env_lit(TbDevice* d, int on) {
    TB_CUDA(cudaSetDevice(d->index));
    d->graph_epoch++;
    d->env_mode = on;
    d->sc.env_lit = on ? 1 : 0;
    d->sc.env_nee = 0;
    tb_free(d->d_env_f); tb_free(d->d_env_cdf); tb_free(d->d_env_rows); tb_free(d->d_env_marg);
    d->d_env_f = d->d_env_cdf = d->d_env_rows = d->d_env_marg = NULL;
    if (on != 2 || (d->sc.env.kind_tex & 3) != TB_ATTR_LATLONG) return;
    const TbDevTexture t = d->h_texs[(size_t)(d->sc.env.kind_tex >> 2)];
    const int nx = t.width, ny = t.height;
    std::vector<float> w((size_t)ny);
    for (int y = 0; y < ny; ++y) w[(size_t)y] = sinf(TB_PI * (((float)y + 0.5f) / (float)ny));
    float* d_w = tb_upload(w.data(), w.size(), d);
    d->d_env_f = tb_alloc<float>((size_t)nx * ny);
    k_env_importance<<<d->sm_count * 4, 256, 0, d->stream>>>(t, d_w, d->d_env_f);
    d->stats.kernel_launches += 1;
    tb_dist2d_build(d, d->stream, d->d_env_f, nx, ny, &d->d_env_cdf, &d->d_env_rows, &d->d_env_marg, &d->sc.env_dist);
    tb_free(d_w);
    d->sc.env_nee = d->sc.env_dist.integral > 0.f ? 1 : 0;
}

// terra_distribution_2d_init + n x terra_distribution_2d_sample on the device (known-answer entry point):
// out[4*i] = {s1 (row axis), s2 (column axis), pdf, 0}.
extern "C" void tb_device_distribution_2d_sample(TbDevice* d, const float* f, size_t nx, size_t ny, size_t n,
                                                 const float* e1, const float* e2, float* out4) {
    TB_CUDA(cudaSetDevice(d->index));
    cudaStream_t st = d->stream;
    float* d_f = tb_upload(f, nx * ny, d);
    float *cdf, *rows, *marg;
    TbDevDist2D dist;
    tb_dist2d_build(d, st, d_f, (int)nx, (int)ny, &cdf, &rows, &marg, &dist);
    float* d_e1 = tb_upload(e1, n, d);
    float* d_e2 = tb_upload(e2, n, d);
    float4* d_out = tb_alloc<float4>(n);
    k_dist_sample<<<d->sm_count * 4, 256, 0, st>>>(dist, n, d_e1, d_e2, d_out);
    d->stats.kernel_launches += 1;
    TB_CUDA(cudaMemcpyAsync(out4, d_out, n * sizeof(float4), cudaMemcpyDeviceToHost, st));
    TB_CUDA(cudaStreamSynchronize(st));
    TB_CUDA(cudaGetLastError());
    tb_free(d_f); tb_free(cdf); tb_free(rows); tb_free(marg); tb_free(d_e1); tb_free(d_e2); tb_free(d_out);
}

extern "C" void tb_device_reset_stats(TbDevice* d) {
    TB_CUDA(cudaSetDevice(d->index));
    k_counters<<<1, 1, 0, d->stream>>>(d->wave.counters, 99, 0);
    TB_CUDA(cudaStreamSynchronize(d->stream));
    memset(&d->stats, 0, sizeof d->stats);
}
