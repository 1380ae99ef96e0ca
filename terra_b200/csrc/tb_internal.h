// tb_internal.h — types shared by the host layer (tb_host.cpp) and the CUDA layer (tb_render.cu).
// Not part of the ABI (include/terra_b200.h is).
#pragma once

#include <stddef.h>
#include <stdint.h>

#include <vector_types.h>

#include "../../include/terra_b200.h"

// ---------------------------------------------------------------- host BVH, reference layout
// Same 64-byte node the reference builds (src/TerraBVH.h:13-17): two child boxes, two child
// indices, two child types (-1 inner, 1 leaf holding object | triangle << 8).
struct TbHostNode {
    TerraAABB aabb[2];
    int32_t   index[2];
    int32_t   type[2];
};
static_assert(sizeof(TbHostNode) == 64, "host node must stay 64 bytes");

// ---------------------------------------------------------------- device layouts (HBM)
// Node, 64 B = 4 x 128-bit loads:
//   q0 = (min0.x, min0.y, min0.z, max0.x)   q1 = (max0.y, max0.z, min1.x, min1.y)
//   q2 = (min1.z, max1.x, max1.y, max1.z)   q3 = (ref0, ref1, -, -) as int bits
// ref >= 0: inner node index; ref < 0: leaf, ~ref = triangle RANK.
//
// Triangles are re-ordered into the reference traversal's leaf visiting order ("rank"), so an
// exact-t tie resolves to the smaller rank exactly as the reference's first-found-wins rule
// (src/TerraBVH.c:290) does. Per rank, 48 B = 3 x 128-bit:
//   (a.xyz, prim bits)  (b.xyz, material/preset bits)  (c.xyz, -)
// Shading properties per rank, 64 B = 4 x 128-bit:
//   (na.xyz, nb.x) (nb.yz, nc.xy) (nc.z, ta.xy, tb.x) (tb.y, tc.xy, -)
enum { TB_PRESET_DIFFUSE = 0, TB_PRESET_PHONG = 1, TB_PRESET_GGX = 2, TB_PRESET_GLASS = 3, TB_PRESET_COUNT = 4 };
enum { TB_ATTR_CONST = 0, TB_ATTR_TEX2D = 1, TB_ATTR_LATLONG = 2 };

struct TbDevAttr {            // 16 B
    float   vx, vy, vz;       // constant value
    int32_t kind_tex;         // kind | texture index << 2
};

struct TbDevTexture {         // 24 B
    const void* pixels;       // device copy
    uint16_t    width, height;
    uint8_t     components, depth, filter, address_mode;
    uint32_t    pad_;
};

struct TbDevMaterial {        // per object
    int32_t   preset;
    int32_t   attr_count;
    float     ior;
    int32_t   light_index;    // index into lights[] or -1
    TbDevAttr emissive;
    TbDevAttr attr[TERRA_MATERIAL_MAX_ATTRIBUTES];
};

struct TbDevLight {
    int32_t  object;          // object index
    uint32_t tri_count;
    uint32_t tri_base;        // first entry of this light in light_tri_area[] / light_tri_rank[]
    float    area;
};

struct TbDevDist2D {          // TerraDistributon2D (src/TerraPrivate.h:85-96) as flat device arrays
    const float* f;              // ny rows of nx: the function
    const float* cdf;            // ny rows of nx: conditional cdfs
    const float* row_integral;   // ny: marginal function (row totals)
    const float* marginal_cdf;   // ny
    float        integral;
    int          nx, ny;
};

struct TbDevScene {
    const float4*        nodes;
    const float4*        tris;
    const float4*        props;
    const TbDevMaterial* materials;
    const TbDevTexture*  textures;
    const TbDevLight*    lights;
    const float*         light_tri_area;   // terra_triangle_area per light triangle (Terra.c:208-215)
    const uint32_t*      light_tri_rank;   // rank of that triangle
    int32_t              n_nodes, n_tris, n_objects, n_lights;
    uint64_t             lights_triangles_count;  // never reset across commits in the reference
    TbDevAttr            env;
    int32_t              integrator, bounces, tonemap, preset_mask;
    float                jitter, exposure, gamma, inv_gamma;
    int32_t              env_lit;          // extension mode: a miss adds throughput * environment (Terra.c:1056 restored)
    int32_t              env_nee;          // env_lit mode 2: the direct integrators sample the lat-long environment explicitly (tb_distribution.cuh)
    TbDevDist2D          env_dist;
};

// ---------------------------------------------------------------- host scene
struct TbTextureUpload { const TerraTexture* host; void* dev_pixels; size_t bytes; };

struct TbDevice;              // opaque to the host layer: all CUDA state (tb_render.cu)

struct TbCameraFrame {        // computed on the host with the reference's float ops
    float rot[3][3];          // rows of terra_camera_to_world_frame (Terra.c:1770-1781)
    float pos[3];
    float tan_half_fov;       // (float) tan(double(fov * 0.0174533f) / 2), Terra.c:1794
};

// CUDA layer entry points used by the host layer ------------------------------------------
struct TbFlatScene {          // host staging of everything the device needs
    const float*         nodes;      size_t n_nodes;      // 16 floats per node
    const float*         tris;       size_t n_tris;       // 12 floats per rank
    const float*         props;                           // 16 floats per rank
    const TbDevMaterial* materials;  size_t n_objects;
    const TbDevLight*    lights;     size_t n_lights;
    const float*         light_tri_area;
    const uint32_t*      light_tri_rank;  size_t n_light_tris;
    const TerraTexture* const* textures;  size_t n_textures;   // distinct host textures, index = TbDevAttr tex id
    uint64_t             lights_triangles_count;
    TbDevAttr            env;
    int                  stack_bound;
    const uint32_t*      qnodes;          // 8 words per node, or NULL (reference tree: float boxes only)
    const uint32_t*      q4nodes;         // 4-wide tree (tb_bvh4_build): 16 words per node, or NULL
    const int32_t*       q4src;           // 4 per wide node
    size_t               n_q4;
    int                  stack_bound4;
    float                grid_lo[3], grid_cell[3];
    int                  exact_boxes;     // 1: reference tree -> reference box arithmetic in k_trace (tb_trace.cuh)
};

#ifdef __cplusplus
extern "C" {
#endif
TbDevice* tb_device_create(int device_index);
void      tb_device_destroy(TbDevice* dev);
void      tb_device_upload_geometry(TbDevice* dev, const TbFlatScene* flat);
void      tb_device_upload_tables(TbDevice* dev, const TbFlatScene* flat);
void      tb_device_download_nodes(TbDevice* dev, float* out16, size_t n_nodes);
void      tb_device_set_options(TbDevice* dev, const TerraSceneOptions* opts, const TbDevAttr* env);
void      tb_device_render(TbDevice* dev, const TbCameraFrame* cam, const TerraFramebuffer* fb,
                           size_t x, size_t y, size_t w, size_t h, size_t spp, uint64_t seed);
void      tb_device_render_resident(TbDevice* dev, const TbCameraFrame* cam, size_t fbw, size_t fbh,
                                    size_t x, size_t y, size_t w, size_t h, size_t sample_base, size_t spp,
                                    uint64_t seed, void* d_results, void* stream);
void      tb_device_partial_begin(TbDevice* dev, const TbCameraFrame* cam, const TerraFramebuffer* fb, size_t x, size_t y, size_t w, size_t h,
                                  size_t chunk, size_t chunk_row0, size_t chunk_rows, size_t slice_row0, size_t slice_rows,
                                  size_t stage_px, size_t sample_base, size_t spp, uint64_t seed);
void      tb_device_partial_finish(TbDevice* dev, TbDevice* const* devs, int n_dev, const TerraFramebuffer* fb, size_t x, size_t y, size_t w, size_t h,
                                   size_t slice_row0, size_t slice_rows, size_t spp_total, int last);
void      tb_device_partial_wait(TbDevice* dev, TbDevice* const* devs, int n_dev);
int       tb_device_enable_peers(TbDevice* const* devs, int n_dev);
void      tb_device_clone_geometry(TbDevice* dst, const TbDevice* src);
void      tb_device_resolve(TbDevice* dev, size_t fbw, size_t fbh, const void* d_results, void* d_pixels, void* stream);
void      tb_device_trace_host(TbDevice* dev, size_t n, const float* origins, const float* dirs, TerraB200Hit* out);
void      tb_device_trace_resident(TbDevice* dev, size_t n, const void* d_o, const void* d_d, void* d_hits, void* stream);
void      tb_device_primary_hits(TbDevice* dev, const TbCameraFrame* cam, size_t w, size_t h, TerraB200Hit* out);
void      tb_device_bsdf_kat(TbDevice* dev, size_t object_idx, size_t n, const float* e123, const float* wo,
                             const float* normal, float* out);
int       tb_device_trace_packed_noalloc(TbDevice* dev, size_t n, const float* d_o3, const float* d_d3, void* d_hits, void* stream);
void      tb_device_get_stats(TbDevice* dev, TerraB200Stats* out);
void      tb_device_round_stats(TbDevice* dev, uint64_t* out10);
const char* tb_device_incoherent_kernel(TbDevice* dev);
void      tb_device_reset_stats(TbDevice* dev);
void      tb_device_set_counting(TbDevice* dev, int on);
void      tb_device_set_env_lit(TbDevice* dev, int on);
void      tb_device_set_graph_paths(TbDevice* dev, size_t max_paths);
void      tb_device_refit(TbDevice* dev, const float* const* obj_tris, const float* const* obj_props, const uint32_t* obj_base, size_t n_objects,
                          const TbDevLight* lights, size_t n_lights, const float* light_area, size_t n_light_tris, float* ms_out);
void      tb_device_distribution_2d_sample(TbDevice* dev, const float* f, size_t nx, size_t ny, size_t n,
                                           const float* e1, const float* e2, float* out4);
int       tb_cuda_device_count(void);
void*     tb_host_alloc_pinned(size_t bytes);
void      tb_host_free_pinned(void* p);
double    tb_device_probe_bandwidth(int device_index, size_t bytes, int mode);
#ifdef __cplusplus
}
#endif
