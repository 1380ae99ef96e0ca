/*
 * terra_b200.h — C ABI of the B200-native Terra hot path.
 *
 * This header is the drop-in boundary.  Every type below has the same name, field order and
 * x86-64 layout as the corresponding type of the reference's public headers, and every
 * `terra_*` function has the reference's signature, so a caller compiled against the
 * reference's include/Terra.h + include/TerraPresets.h links against libterra_b200.so
 * unchanged.  Layouts are pinned by the _Static_asserts at the end of the file (sizes/offsets
 * measured on the reference).
 *
 * Reference interface each group replaces (paths relative to the reference checkout):
 *   math PODs ............ include/TerraMath.h:20-38
 *   shading types ........ include/Terra.h:36-99
 *   scene types .......... include/Terra.h:104-198
 *   scene API ............ include/Terra.h:205-211   (src/Terra.c:130-282)
 *   framebuffer API ...... include/Terra.h:213-215   (src/Terra.c:309-345)
 *   texture API .......... include/Terra.h:217-223   (src/Terra.c:350-507)
 *   attribute API ........ include/Terra.h:225-227   (src/Terra.c:287-304)
 *   terra_render ......... include/Terra.h:229       (src/Terra.c:512-635)  <- the hot path
 *   system API ........... include/Terra.h:237-245   (src/Terra.c:640-663)
 *   BSDF presets ......... include/TerraPresets.h:11-27 (src/TerraPresets.c:34-146)
 *
 * What differs from the reference, by design:
 *   - terra_render() runs on the GPU (sm_100a wavefront kernels).  There is no CPU fallback:
 *     without a CUDA device the call logs through terra_log() and abort()s.
 *   - Materials are host function-pointer tables in the reference.  A GPU cannot call them,
 *     so terra_scene_commit() classifies each TerraBSDF / TerraAttribute by POINTER IDENTITY
 *     against this library's own exported presets (terra_bsdf_*_init, terra_texture_sample,
 *     terra_texture_sample_latlong).  An unknown pointer is a hard error.
 *   - terra_bsdf_ggx_init / terra_bsdf_glass_init are additive: the reference only carries
 *     those two presets as dead code (src/TerraPresets.c:298-465).
 *   - Everything prefixed terra_b200_ is an extension (device selection, the traversal
 *     micro-benchmark entry points, device-pointer variants for multi-GPU, statistics).
 */
#ifndef TERRA_B200_H
#define TERRA_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ math PODs ---------- */
typedef struct TerraInt4     { int   x, y, z, w; } TerraInt4;
typedef struct TerraFloat2   { float x, y;       } TerraFloat2;
typedef struct TerraFloat3   { float x, y, z;    } TerraFloat3;
typedef struct TerraFloat4   { float x, y, z, w; } TerraFloat4;
typedef struct TerraFloat4x4 { TerraFloat4 rows[4]; } TerraFloat4x4;   /* row-major, M*v */

/* The reference's constants, verbatim (include/TerraMath.h:17-18).  terra_PI is NOT pi. */
#define terra_PI       3.1416926535f
#define terra_Epsilon  1e-4
#define terra_ior_air  1.f

/* ------------------------------------------------------------------ shading ------------ */
#ifndef TERRA_MATERIAL_MAX_ATTRIBUTES
#define TERRA_MATERIAL_MAX_ATTRIBUTES 8
#endif

/* Surface record handed to the BSDF callbacks (reference: 188 bytes). */
typedef struct {
    TerraFloat4x4 transform;   /* tangent | normal | bitangent as COLUMNS (TerraMath.inl:251-272) */
    TerraFloat3   normal;
    TerraFloat3   emissive;
    float         ior;         /* never written by the reference */
    TerraFloat3   attributes[TERRA_MATERIAL_MAX_ATTRIBUTES];
} TerraShadingSurface;

typedef TerraFloat3 (TerraBSDFSampleRoutine)(const TerraShadingSurface* surface,
                                             float e1, float e2, float e3, const TerraFloat3* wo);
typedef float       (TerraBSDFPdfRoutine)   (const TerraShadingSurface* surface,
                                             const TerraFloat3* wi, const TerraFloat3* wo);
typedef TerraFloat3 (TerraBSDFEvalRoutine)  (const TerraShadingSurface* surface,
                                             const TerraFloat3* wi, const TerraFloat3* wo);

typedef struct {
    TerraBSDFSampleRoutine* sample;
    TerraBSDFPdfRoutine*    pdf;
    TerraBSDFEvalRoutine*   eval;
} TerraBSDF;

typedef enum {
    kTerraFilterPoint,
    kTerraFilterBilinear,
    kTerraFilterTrilinear,     /* returns black in the reference (Terra.c:452-458) */
    kTerraFilterAnisotropic    /* idem */
} TerraFilter;

typedef enum {
    kTerraTextureAddressWrap,
    kTerraTextureAddressMirror,
    kTerraTextureAddressClamp
} TerraTextureAddressMode;

typedef struct {
    void*    pixels;           /* NULL = invalid texture */
    uint16_t width;
    uint16_t height;
    uint8_t  components;
    uint8_t  depth;            /* bytes per component: 1 (u8/255) or 4 (float) */
    uint8_t  filter;           /* TerraFilter; caller sets it */
    uint8_t  address_mode;     /* TerraTextureAddressMode; caller sets it */
} TerraTexture;

typedef void        (*TerraAttributeFinalize)(void* attribute);
typedef TerraFloat3 (*TerraAttributeEval)(void* attribute, const void* texcoord, const void* world_pos);

typedef struct {
    void*                  state;     /* NULL => constant `value` (Terra.c:1804-1810) */
    TerraAttributeFinalize finalize;
    TerraAttributeEval     eval;
    TerraFloat3            value;
} TerraAttribute;

typedef struct {
    TerraBSDF      bsdf;
    float          ior;
    TerraAttribute emissive;
    TerraAttribute attributes[TERRA_MATERIAL_MAX_ATTRIBUTES];
    size_t         attributes_count;
    bool           enable_bump_map_attr;
    bool           enable_normal_map_attr;
} TerraMaterial;

/* ------------------------------------------------------------------ scene -------------- */
typedef struct TerraAABB { TerraFloat3 min, max; } TerraAABB;
typedef struct { TerraFloat3 a, b, c; } TerraTriangle;

typedef struct {
    TerraFloat3 normal_a, normal_b, normal_c;
    TerraFloat2 texcoord_a, texcoord_b, texcoord_c;
} TerraTriangleProperties;

typedef struct {
    TerraTriangle*           triangles;        /* owned by the scene, filled in place by the caller */
    TerraTriangleProperties* properties;
    size_t                   triangles_count;
    TerraMaterial            material;
} TerraObject;

typedef enum {
    kTerraTonemappingOperatorNone,
    kTerraTonemappingOperatorLinear,
    kTerraTonemappingOperatorReinhard,
    kTerraTonemappingOperatorFilmic,
    kTerraTonemappingOperatorUncharted2
} TerraTonemappingOperator;

typedef enum { kTerraAcceleratorBVH } TerraAccelerator;

typedef enum {
    kTerraSamplingMethodRandom,
    kTerraSamplingMethodStratified,   /* only changes the spp rounding */
    kTerraSamplingMethodHalton
} TerraSamplingMethod;

typedef enum {
    kTerraIntegratorSimple,
    kTerraIntegratorDirect,
    kTerraIntegratorDirectMis,
    kTerraIntegratorDebugMono,
    kTerraIntegratorDebugDepth,
    kTerraIntegratorDebugNormals,
    kTerraIntegratorDebugMisWeights
} TerraIntegrator;

typedef struct {
    TerraAttribute           environment_map;
    TerraTonemappingOperator tonemapping_operator;
    TerraAccelerator         accelerator;
    TerraSamplingMethod      sampling_method;
    TerraIntegrator          integrator;
    float                    subpixel_jitter;
    size_t                   samples_per_pixel;
    size_t                   bounces;
    size_t                   strata;
    float                    manual_exposure;
    float                    gamma;
} TerraSceneOptions;

typedef struct {
    TerraFloat3 position, direction, up;
    float       fov;           /* degrees */
} TerraCamera;

typedef struct {
    TerraFloat3 acc;           /* running radiance sum */
    int         samples;       /* running sample count */
} TerraRawIntegrationResult;

typedef struct {
    TerraFloat3*               pixels;    /* tonemapped, row-major p = i*width + j */
    TerraRawIntegrationResult* results;   /* accumulates across terra_render calls */
    size_t                     width;
    size_t                     height;
} TerraFramebuffer;

typedef struct {
    uint32_t object_idx   : 8;            /* <= 256 objects          (TerraBVH.c:155,283-284) */
    uint32_t triangle_idx : 24;           /* <= 2^23 triangles/object (signed shift) */
} TerraPrimitiveRef;

typedef void* HTerraScene;

/* ------------------------------------------------------------------ reference API ------ */
HTerraScene        terra_scene_create(void);
TerraObject*       terra_scene_add_object(HTerraScene scene, size_t triangle_count);
size_t             terra_scene_count_objects(HTerraScene scene);
void               terra_scene_commit(HTerraScene scene);      /* BVH build + flatten + upload */
void               terra_scene_clear(HTerraScene scene);
TerraSceneOptions* terra_scene_get_options(HTerraScene scene); /* applied at the next commit */
void               terra_scene_destroy(HTerraScene scene);

bool terra_framebuffer_create(TerraFramebuffer* framebuffer, size_t width, size_t height);
void terra_framebuffer_clear(TerraFramebuffer* framebuffer);
void terra_framebuffer_destroy(TerraFramebuffer* framebuffer);

bool        terra_texture_init(TerraTexture* texture, size_t width, size_t height, size_t components, const void* data);
bool        terra_texture_init_hdr(TerraTexture* texture, size_t width, size_t height, size_t components, const float* data);
TerraFloat3 terra_texture_read(TerraTexture* texture, size_t x, size_t y);
TerraFloat3 terra_texture_sample(void* texture, const void* uv, const void* xyz);
TerraFloat3 terra_texture_sample_latlong(void* texture, const void* dir, const void* xyz);
void        terra_texture_destroy(TerraTexture* texture);
void        terra_texture_finalize(void* texture);

void terra_attribute_init_constant(TerraAttribute* attr, const TerraFloat3* value);
void terra_attribute_init_texture(TerraAttribute* attr, TerraTexture* texture);
void terra_attribute_init_cubemap(TerraAttribute* attr, TerraTexture* texture);

/* Renders region [x,x+width) x [y,y+height): adds samples_per_pixel samples to
 * framebuffer->results and rewrites framebuffer->pixels for the region.  Blocking; safe to call
 * concurrently on disjoint regions of one framebuffer with one committed scene. */
void terra_render(const TerraCamera* camera, HTerraScene scene, const TerraFramebuffer* framebuffer,
                  size_t x, size_t y, size_t width, size_t height);

void* terra_malloc(size_t size);
void* terra_realloc(void* ptr, size_t size);
void  terra_free(void* ptr);
void  terra_log(const char* fmt, ...);

/* ------------------------------------------------------------------ BSDF presets ------- */
/* attribute slots, reference include/TerraPresets.h:11-26 */
#define TERRA_DIFFUSE_ALBEDO           0
#define TERRA_DIFFUSE_END              1
#define TERRA_PHONG_SPECULAR_COLOR     0
#define TERRA_PHONG_ALBEDO             1
#define TERRA_PHONG_SPECULAR_INTENSITY 2
#define TERRA_PHONG_SAMPLE_PICK        3
#define TERRA_PHONG_END                4
void terra_bsdf_diffuse_init(TerraBSDF* bsdf);
void terra_bsdf_phong_init(TerraBSDF* bsdf);

/* Additive presets: live {sample,pdf,eval} restatements of the reference's dead code
 * (src/TerraPresets.c:298-394 rough dielectric GGX, :396-464 perfect glass).
 *   GGX   : [0]=albedo  [1]=(roughness, metalness, ior)
 *   GLASS : [0]=albedo  [1]=(ior, -, -)  [2].x = fresnel state written by sample() (init 0) */
#define TERRA_GGX_ALBEDO               0
#define TERRA_GGX_PARAMS               1
#define TERRA_GGX_END                  2
#define TERRA_GLASS_ALBEDO             0
#define TERRA_GLASS_PARAMS             1
#define TERRA_GLASS_STATE              2
#define TERRA_GLASS_END                3
void terra_bsdf_ggx_init(TerraBSDF* bsdf);
void terra_bsdf_glass_init(TerraBSDF* bsdf);

/* ------------------------------------------------------------------ extensions --------- */
/* Closest-hit record of the traversal entry points (16 bytes). */
typedef struct {
    float    t;          /* distance from the OFFSET origin (origin + 0.001*dir, Terra.c:1628-1630); <0 = miss */
    uint32_t prim;       /* object_idx | triangle_idx << 8   (0xFFFFFFFF = miss) */
    float    u, v;       /* U/det, V/det of TerraGeometry.c:248-249 */
} TerraB200Hit;

typedef struct {
    uint64_t kernel_launches;    /* CUDA kernels launched by this library since the last reset */
    uint64_t rays_traced;        /* terra_scene_raycast equivalents */
    uint64_t paths_traced;       /* (pixel,sample) pairs */
    double   trace_kernel_ms;    /* device time spent in the traversal kernel (CUDA events) */
    uint64_t trace_kernel_launches;
    double   render_ms;          /* device time of whole terra_render calls */
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t nodes_visited;      /* only while counting is on (terra_b200_set_counting): 64-byte nodes fetched */
    uint64_t tris_tested;        /*   and triangles tested by the traversal kernel */
    uint64_t graph_replays;      /* terra_render calls replayed from a CUDA graph (small calls of a shape seen before) */
    /* The camera-ray launches alone (they run the packet kernel; everything else — bounce, visibility and API rays —
     * is the difference to the totals above): one roofline per kernel. */
    uint64_t coherent_rays;
    double   coherent_kernel_ms;
    uint64_t coherent_kernel_launches;
    uint64_t coherent_nodes_visited, coherent_tris_tested;
    uint64_t p2p_bytes;          /* device-to-device copies (scene replicas over NVLink) */
} TerraB200Stats;

/* Where the lanes of the incoherent-ray kernel's rounds go (counting on, primary device; reset by terra_b200_reset_stats).
 * Lane-rounds: [0] all (32 x warp rounds)  [1] fetched a node  [2] popped an entry the closest hit had culled
 * [3] held a second leaf while the first waits for a triangle round  [4] parked with nothing else to do
 * [5] owned no ray  [6] lanes x triangle rounds (32 each)  [7] lanes that tested a triangle
 * [8] 32 x refills  [9] lanes refilled.  profiles/r2_summary.md reads the kernel's instruction budget from these. */
void terra_b200_round_stats(HTerraScene scene, uint64_t* out10);
/* Name of the closest-hit kernel that bounce / visibility / API rays of this (committed) scene run: "k_trace_v8" (binary
 * tree, scenes that fit L2), "k_trace_v7" (4-wide tree, larger scenes), "k_trace_v6" / "k_trace" under the A/B switches.
 * Static string. bench.py labels its roofline with it. */
const char* terra_b200_incoherent_kernel(HTerraScene scene);
int  terra_b200_device_count(void);
void terra_b200_set_device(int device);              /* before terra_scene_commit */
/* Multi-GPU behind terra_render (the splitter of SURVEY 8b / 8e; the reference's own fan-out is one terra_render per
 * tile from a thread pool, satellite/src/Renderer.cpp:316-349): a scene lives on one or more GPUs of one box —
 * terra_b200_scene_set_devices before its first commit, default the environment variable TERRA_B200_DEVICES
 * ("0,1,2,3" or "all"), else the terra_b200_set_device one. terra_scene_commit replicates the scene (geometry goes
 * up once and is copied device to device over NVLink). A terra_render call of at least min_paths (pixel, sample)
 * pairs (default 16M) is then split by sample range: GPU g renders range g of the whole region, and finalises and
 * returns row band g of it, summing the band over all GPUs' accumulators through peer-mapped memory inside the
 * accumulate / tonemap kernel. Smaller calls, and every other entry point, run on the first device. Sample ranges
 * use the same counter-based random numbers as one GPU; the per-pixel float sum re-associates.
 * Optional (off by default: it measured 1 % slower on the 1080p benchmark frame at 2 and at 8 GPUs): a split call of at
 * least 2 x chunk_paths (pixel, sample) pairs per GPU goes through as two row chunks, the first seven eighths of the
 * region and the rest, and the framebuffer traffic of the first chunk (PCIe in and out, NVLink reduce) runs beside the
 * rendering of the second. The image does not depend on the chunking, bit for bit.
 * One host thread per GPU issues that GPU's work. */
void terra_b200_scene_set_devices(HTerraScene scene, const int* device_ids, int count);
int  terra_b200_scene_device_count(HTerraScene scene);
void terra_b200_set_fanout_min_paths(HTerraScene scene, size_t min_paths);
void terra_b200_set_fanout_chunk_paths(HTerraScene scene, size_t chunk_paths);
void terra_b200_set_seed(HTerraScene scene, uint64_t seed);   /* default: time based, like the reference */
void terra_b200_get_stats(HTerraScene scene, TerraB200Stats* out);
void terra_b200_reset_stats(HTerraScene scene);
/* Switch the traversal kernel to its counting variant (slower; never on in timed runs). */
void terra_b200_set_counting(HTerraScene scene, int on);

/* Memory-system probe for the roofline denominators MEASURED_PEAKS.json lacks: GB/s of (mode 0) a
 * streaming 128-bit read or (mode 1) random 64-byte record gathers over a `bytes`-sized buffer
 * (<= ~64 MB: L2-resident; >= 1 GB: HBM). */
double terra_b200_probe_bandwidth(size_t bytes, int mode);

/* The host BVH exactly as the reference lays it out (64-byte nodes, TerraBVH.h:13-17). */
const void* terra_b200_bvh_nodes(HTerraScene scene, int* nodes_count);
int         terra_b200_bvh_stack_bound(HTerraScene scene);   /* worst-case ordered-traversal stack */
/* Host-only half of commit (no GPU needed): builds the BVH, optionally writes the flattened GPU
 * arrays (16 floats per node: two boxes + two child refs; 12 floats per triangle in rank order).
 * Returns the node count (= triangles - 1). */
int         terra_b200_build_host_bvh(HTerraScene scene, int mode, float* flat_nodes_out, float* flat_tris_out);
/* Acceleration structure built by the next terra_scene_commit, process-wide:
 *   TERRA_B200_BVH_REFERENCE  the reference's tree (x-only SAH, src/TerraBVH.c:79-244), byte-identical nodes;
 *   TERRA_B200_BVH_SAH3       binned SAH over x,y,z in the same 64-byte node format (default);
 *   TERRA_B200_BVH_LBVH       linear BVH built ON THE GPU in the same node format (replaces terra_bvh_create,
 *                             src/TerraBVH.c:128-244, where the host build dominates time-to-first-pixel).
 * Closest hits do not depend on the choice (same triangles, same intersection arithmetic);
 * traversal work does. Also settable with the environment variable TERRA_B200_BVH=reference|sah3|lbvh. */
#define TERRA_B200_BVH_REFERENCE 0
#define TERRA_B200_BVH_SAH3      1
#define TERRA_B200_BVH_LBVH      2
void        terra_b200_set_bvh_mode(int mode);
/* Extension mode "env-lit" (NOT the reference's behaviour): the reference evaluates the environment map on
 * a miss, multiplies it into the throughput and then discards it — `//Lo = terra_addf3 ( &Lo, &throughput );`
 * is commented out at src/Terra.c:1056 — so environment light contributes nothing. With on != 0 a path that
 * leaves the scene adds throughput * environment_map(direction) (constant, 2-D texture or lat-long texture,
 * src/Terra.c:468-477, 1804-1810). Off by default; call after terra_scene_commit. Checked against the
 * reference compiled with exactly that one line restored (oracle/Makefile, "patched oracle"). */
/* on = 2 ("env-lit, importance sampled"): the same radiance as on = 1, estimated with explicit environment
 * samples in the direct and direct-MIS integrators: the reference's written-but-never-called
 * TerraDistributon2D (src/Terra.c:760-846) over luminance x sin(theta) of the lat-long environment texture picks
 * a direction per bounce, a visibility ray tests it, power-heuristic weights combine it with the BSDF sample
 * under MIS, and rays that leave the scene count at the camera ray only. Needs a lat-long environment
 * (terra_attribute_init_cubemap); anything else behaves as on = 1. The simple integrator is unchanged. */
void        terra_b200_set_env_lit(HTerraScene scene, int on);

/* terra_distribution_2d_init + n x terra_distribution_2d_sample (src/Terra.c:811-846) on the device, for a
 * function f of ny rows x nx columns: out4[4*i] = { s1 (row axis), s2 (column axis), pdf, 0 } for (e1[i], e2[i]).
 * Bit-identical to the reference's exported functions (tests/test_distribution.py). */
void        terra_b200_distribution_2d_sample(HTerraScene scene, const float* f, size_t nx, size_t ny, size_t n,
                                              const float* e1, const float* e2, float* out4);
/* Wall time of the last terra_scene_commit's stages, ms: [0] tree build (terra_bvh_create's replacement),
 * [1] flatten to the GPU layout, [2] node quantisation, [3] host->device upload. */
void        terra_b200_commit_ms(HTerraScene scene, float* out4);
/* Device time (kernels + sort) and time including the triangle upload of the last LBVH build, ms. */
void        terra_b200_lbvh_build_ms(float* device_ms, float* with_upload_ms);

/* Asynchronous tile jobs: what satellite's TerraRenderer does with its worker threads
 * (satellite/src/Renderer.cpp:70-98 terra_render_launcher, :316-349 tile grid + job push) as a library call.
 * The region is cut into tile_size x tile_size tiles in row-major order (tile_size 0 = one call for the whole
 * region) and a worker thread calls terra_render on each; the call returns at once. on_tile_begin / on_tile_end
 * (either may be NULL) run on the worker thread around each tile. The camera is copied; scene and framebuffer
 * must outlive the job. Results are those of the same terra_render calls made in that order by the caller. */
typedef struct TerraB200Job* HTerraB200Job;
typedef void (*TerraB200TileEvent)(size_t x, size_t y, size_t width, size_t height, void* user);
HTerraB200Job terra_b200_render_async(const TerraCamera* camera, HTerraScene scene, const TerraFramebuffer* framebuffer,
                                      size_t x, size_t y, size_t width, size_t height, size_t tile_size,
                                      TerraB200TileEvent on_tile_begin, TerraB200TileEvent on_tile_end, void* user);
int         terra_b200_job_poll(HTerraB200Job job);      /* tiles left (the reference's _tile_counter); 0 = finished */
void        terra_b200_job_wait(HTerraB200Job job);      /* block until every tile is done */
void        terra_b200_job_cancel(HTerraB200Job job);    /* skip tiles not started yet; returns once the running one ends */
void        terra_b200_job_release(HTerraB200Job job);   /* wait, then free the handle */

/* Re-fit instead of rebuild (the "refit" half of replacing terra_bvh_create, src/TerraBVH.c:128-244): after editing
 * vertex positions / normals / texcoords of the committed objects IN PLACE (same objects, same triangle counts),
 * every box of the tree is recomputed bottom-up on the GPU and the nodes are re-quantised; light triangle areas are
 * refreshed. Closest hits afterwards are those of a fresh commit of the edited geometry (t, u, v bit-identical).
 * Not for the reference tree (TERRA_B200_BVH_REFERENCE). Returns the device time of the refit kernels, ms. */
float       terra_b200_refit(HTerraScene scene);

/* Render calls of at most max_paths (pixel, sample) pairs whose shape (width, height, spp) repeats are replayed
 * from a CUDA graph instead of being launched kernel by kernel (same results; default 4M paths, environment
 * TERRA_B200_GRAPH_PATHS; 0 = never). Per-kernel trace timings in TerraB200Stats are not taken on replayed calls. */
void        terra_b200_set_graph_paths(HTerraScene scene, size_t max_paths);

/* Host-only (tests): the 4-wide tree terra_scene_commit derives for the incoherent-ray kernel from the tree of `mode`
 * (16 words per node: 4 child boxes x 3 words lo | hi << 16, 4 refs), where each slot's box comes from
 * (binary node * 2 + child, -1 = empty) and the quantised binary nodes themselves (8 words each). Returns the wide
 * node count; writes at most `cap` wide nodes. */
int terra_b200_build_host_bvh4(HTerraScene scene, int mode, uint32_t* q4nodes_out, int32_t* src_out, uint32_t* qnodes_out,
                               int cap, int* stack_bound_out);
/* Page-locked host memory for large ray batches (and anything else handed to the entry points below):
 * copies from / to it run at PCIe speed and overlap the kernels. NULL when no CUDA device is visible.
 * Framebuffers made by terra_framebuffer_create are page-locked already. */
void*       terra_b200_host_alloc(size_t bytes);
void        terra_b200_host_free(void* p);

/* Traversal micro-benchmark / parity entry points (= terra_scene_raycast without shading, src/Terra.c:276-279,
 * for a whole batch). Host buffers: origins/dirs are n x 3 floats; the H2D/D2H copies are part of the call
 * and are pipelined against the traversal in chunks (copy in | trace | copy out on three streams) when the
 * buffers are page-locked; pageable buffers give the same hits, serially. */
void terra_b200_trace_rays(HTerraScene scene, size_t n, const float* origins, const float* directions,
                           TerraB200Hit* hits_out);
/* Device buffers: float4 per ray (xyz used), caller's stream (cudaStream_t as void*). */
void terra_b200_trace_rays_device(HTerraScene scene, size_t n, const void* d_origins, const void* d_directions,
                                  void* d_hits_out, void* stream);
/* No counterpart in the reference: added for callers that cannot allocate device
 * memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). Closest hits of a DEVICE
 * batch of packed rays — n x 3 floats each for origins and directions, the layout terra_b200_trace_rays takes from the
 * host — into n TerraB200Hit on the device, on the caller's stream (NULL: the library's); returns when the hits are
 * final, like terra_b200_trace_rays_device, and gives the same hits bit for bit. The call allocates nothing: the
 * float4 ray records and the fetch cursor the kernels need live in a file-scope device array of
 * 256 + 32 x TERRA_B200_NOALLOC_MAX_RAYS bytes (8 MB) that every GPU holds from the moment the library is loaded.
 * Returns 0, or -1 (and logs, writes nothing) when n exceeds TERRA_B200_NOALLOC_MAX_RAYS. Runs on the scene's first
 * device. All calls of a process share the one array per GPU: it keeps the previous call's contents (each call clears
 * what must start at zero), and two calls that overlap on the same GPU — other scenes, other streams — race. */
#define TERRA_B200_NOALLOC_MAX_RAYS 262144
int  terra_b200_trace_rays_packed_noalloc(HTerraScene scene, size_t n, const float* d_origins_xyz, const float* d_directions_xyz,
                                          void* d_hits_out, void* stream);
/* Primary camera rays at subpixel_jitter = 0, one per pixel, row-major. */
void terra_b200_primary_hits(const TerraCamera* camera, HTerraScene scene, size_t width, size_t height,
                             TerraB200Hit* hits_out);

/* Device-resident rendering for multi-GPU drivers: adds `spp` samples per pixel, sample
 * indices [sample_base, sample_base+spp), to d_results (TerraRawIntegrationResult[width*height]
 * in device memory) for the given region; no tonemap, no host copy. */
void terra_b200_render_device(const TerraCamera* camera, HTerraScene scene, size_t fb_width, size_t fb_height,
                              size_t x, size_t y, size_t width, size_t height,
                              size_t sample_base, size_t spp, void* d_results, void* stream);
/* acc/samples * exposure + tonemap (Terra.c:574-630) from d_results into d_pixels (float3). */
void terra_b200_resolve_device(HTerraScene scene, size_t fb_width, size_t fb_height,
                               const void* d_results, void* d_pixels, void* stream);

/* BSDF known-answer entry point: runs sample/pdf/eval of the object's preset on the GPU for n
 * (e1,e2,e3,wo,normal) tuples. out: n x 8 floats = wi.xyz, pdf, f.xyz, state. */
void terra_b200_bsdf_kat(HTerraScene scene, size_t object_idx, size_t n,
                         const float* e123, const float* wo, const float* normal, float* out);

#ifdef __cplusplus
}
#endif

/* ------------------------------------------------------------------ ABI pins ----------- */
#if defined(__cplusplus)
#define TERRA_B200_SA(c, m) static_assert(c, m)
#else
#define TERRA_B200_SA(c, m) _Static_assert(c, m)
#endif
#if defined(__x86_64__) || defined(__aarch64__)
TERRA_B200_SA(sizeof(TerraShadingSurface) == 188, "TerraShadingSurface ABI");
TERRA_B200_SA(sizeof(TerraTexture) == 16, "TerraTexture ABI");
TERRA_B200_SA(sizeof(TerraAttribute) == 40, "TerraAttribute ABI");
TERRA_B200_SA(sizeof(TerraMaterial) == 408, "TerraMaterial ABI");
TERRA_B200_SA(offsetof(TerraMaterial, emissive) == 32, "TerraMaterial.emissive ABI");
TERRA_B200_SA(offsetof(TerraMaterial, attributes) == 72, "TerraMaterial.attributes ABI");
TERRA_B200_SA(offsetof(TerraMaterial, attributes_count) == 392, "TerraMaterial.attributes_count ABI");
TERRA_B200_SA(sizeof(TerraTriangle) == 36, "TerraTriangle ABI");
TERRA_B200_SA(sizeof(TerraTriangleProperties) == 60, "TerraTriangleProperties ABI");
TERRA_B200_SA(sizeof(TerraObject) == 432, "TerraObject ABI");
TERRA_B200_SA(offsetof(TerraObject, material) == 24, "TerraObject.material ABI");
TERRA_B200_SA(sizeof(TerraSceneOptions) == 96, "TerraSceneOptions ABI");
TERRA_B200_SA(offsetof(TerraSceneOptions, subpixel_jitter) == 56, "TerraSceneOptions.subpixel_jitter ABI");
TERRA_B200_SA(offsetof(TerraSceneOptions, samples_per_pixel) == 64, "TerraSceneOptions.spp ABI");
TERRA_B200_SA(offsetof(TerraSceneOptions, manual_exposure) == 88, "TerraSceneOptions.exposure ABI");
TERRA_B200_SA(sizeof(TerraCamera) == 40, "TerraCamera ABI");
TERRA_B200_SA(sizeof(TerraRawIntegrationResult) == 16, "TerraRawIntegrationResult ABI");
TERRA_B200_SA(sizeof(TerraFramebuffer) == 32, "TerraFramebuffer ABI");
TERRA_B200_SA(sizeof(TerraB200Hit) == 16, "TerraB200Hit ABI");
#endif

#endif /* TERRA_B200_H */
