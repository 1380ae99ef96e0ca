/*
 * terra_oracle.h — TEST INFRASTRUCTURE ONLY (nothing under oracle/ is on the product path).
 *
 * Plain-C CPU restatement of the reference's hot path (terra-render/Terra @ fe2404b7), written
 * from the algorithm, every function citing the reference file:line it follows. It is pinned
 * against the UNMODIFIED reference (oracle/_ref/libterra_ref.so) by tests/test_oracle_pin.py:
 *   - BVH node array byte-identical to terra_bvh_create;
 *   - closest hits (id, t, u, v) bit-identical to terra_bvh_traverse on seeded ray sets;
 *   - whole images bit-identical to terra_render at subpixel_jitter = 0 under the same srand()
 *     (the reference draws BSDF/RR numbers from libc rand(); the oracle draws them in the same order);
 *   - its Diffuse/Phong callbacks, plugged into the reference's terra_render, reproduce the
 *     reference's own presets bit for bit.
 * GGX and Glass exist in the reference only as dead code (src/TerraPresets.c:298-465): the
 * oracle restates them for the live {sample,pdf,eval} signature and they run inside the
 * unmodified reference terra_render through the function-pointer table. No reference test or
 * vector covers those two; they are pinned against the dead code itself, compiled through the
 * adapter of the retired API in oracle/ref_dead_presets.c (tests/golden/bsdf_dead.npz): sampled
 * directions and glass weights bit-identical, GGX weight / shade within 1e-4 relative.
 *
 * Types come from include/terra_b200.h (layout-identical to the reference's public headers).
 */
#ifndef TERRA_ORACLE_H
#define TERRA_ORACLE_H

#include "../include/terra_b200.h"

typedef struct OracleScene OracleScene;

typedef struct { float t; uint32_t prim; float u, v; } OracleHit;     /* == TerraB200Hit */

/* Builds the BVH (reference builder restated) and the light table over caller-owned objects. */
OracleScene* oracle_scene_create(const TerraObject* objects, int n_objects, const TerraSceneOptions* opts);
void         oracle_scene_destroy(OracleScene* s);
/* Extension mode "env-lit" (SURVEY §8a row 12): a path that leaves the scene adds throughput * environment,
 * i.e. the reference with the commented-out line Terra.c:1056 restored. Off by default (parity mode).
 * Pinned against oracle/_ref/libterra_ref_envlit.so, the reference compiled with exactly that line restored. */
void         oracle_set_env_lit(OracleScene* s, int on);
const void*  oracle_bvh_nodes(const OracleScene* s, int* n_nodes);              /* 64-byte reference-layout nodes */

/* terra_scene_raycast without shading, reference (un-culled, LIFO) walk. */
void oracle_trace_rays(const OracleScene* s, size_t n, const float* origins, const float* dirs, OracleHit* out);
/* Canonical work counter of SURVEY §8(d): near-first, t-culled walk of an arbitrary 64-byte-node
 * tree (`nodes` may be the reference tree or any other tree in the same format). Returns the same
 * hits; counters[0] += nodes fetched, counters[1] += triangles tested. */
void oracle_trace_counted(const OracleScene* s, const void* nodes, size_t n, const float* origins,
                          const float* dirs, OracleHit* out, uint64_t counters[2]);
void oracle_primary_hits(const OracleScene* s, const TerraCamera* cam, size_t w, size_t h, OracleHit* out, float* dirs_out);

/* terra_render (src/Terra.c:512-635). Jitter randoms come from a PCG32 seeded with `pcg_seed`
 * (the reference seeds from time()); BSDF / roulette randoms from libc rand(), as the reference. */
void oracle_render(const OracleScene* s, const TerraCamera* cam, const TerraFramebuffer* fb,
                   size_t x, size_t y, size_t w, size_t h, uint32_t pcg_seed);

/* terra_surface_init at the closest hit of one ray; returns object index or -1. */
int  oracle_raycast_surface(const OracleScene* s, const float* origin, const float* dir,
                            TerraShadingSurface* surf, float* point, int* tri);

/* KAT helper: surface record (frame from `normal`, 8 x float3 attribute values) for BSDF callbacks. */
void oracle_surface_from_normal(const float* normal, const float* attributes24, TerraShadingSurface* sf);

/* Presets as TerraBSDF callbacks (usable inside the reference's terra_render). */
/* Discrete distributions (src/Terra.c:760-846, src/TerraPrivate.h:85-102): the reference's written-but-unused
 * building block for an importance-sampled environment. Flat API over caller arrays: cdf/func_out have n
 * (nx*ny) floats. Pinned bit-exact against the functions libterra_ref.so exports. */
float oracle_distribution_1d_init(const float* f, size_t n, float* cdf_out);                      /* returns the integral */
float oracle_distribution_1d_sample(const float* f, const float* cdf, float integral, size_t n, float e,
                                    float* pdf_out, size_t* idx_out);
typedef struct {
    size_t nx, ny;
    float* f;            /* ny rows of nx */
    float* cdf;          /* conditional cdfs, ny rows of nx */
    float* row_integral; /* = marginal f, ny */
    float* marginal_cdf; /* ny */
    float  integral;
} OracleDistribution2D;
OracleDistribution2D* oracle_distribution_2d_create(const float* f, size_t nx, size_t ny);
void  oracle_distribution_2d_destroy(OracleDistribution2D* d);
/* n samples: out[4*i] = {s1 (row axis), s2 (column axis), pdf, 0} */
void  oracle_distribution_2d_sample(const OracleDistribution2D* d, size_t n, const float* e1, const float* e2, float* out);
/* Environment importance function of the env-NEE extension: luminance (Terra.c:1036 weights) of the point-sampled
 * lat-long texel times sin(pi_ref * (y + 0.5) / ny). rgb = ny*nx*3 floats. */
void  oracle_env_importance(const float* rgb, size_t nx, size_t ny, float* f_out);

void terra_oracle_bsdf_diffuse_init(TerraBSDF* b);
void terra_oracle_bsdf_phong_init(TerraBSDF* b);
void terra_oracle_bsdf_ggx_init(TerraBSDF* b);
void terra_oracle_bsdf_glass_init(TerraBSDF* b);

#endif
