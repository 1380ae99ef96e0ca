/* TEST INFRASTRUCTURE ONLY — nothing under oracle/ is on the product path.
 *
 * ref_dead_presets.c — compiles the reference's DEAD GGX and glass presets so that the restatements in
 * oracle/terra_oracle.c (and, through them, terra_b200/csrc/tb_bsdf.h) can be checked against the reference's own
 * arithmetic instead of against each other only.
 *
 * The reference keeps its rough-dielectric (GGX) and perfect-glass presets inside `#if 0` blocks written against a
 * RETIRED plug-in API — {sample, weight, shade} callbacks over (material, shading state, shading context) —
 * src/TerraPresets.c:13-28 (Schlick fresnel, R0) and :298-465 (the two presets). oracle/Makefile streams exactly
 * those function bodies (lines 14-27, 302-388, 399-458: everything but the two init functions, which assign
 * callback fields the live TerraBSDF no longer has) through sed into a temporary file outside the repository and
 * this translation unit #includes it (TB_DEAD_PRESETS_INC). No reference source enters the tree.
 *
 * What is OURS here, and therefore not evidence: the declarations of the retired API that the dead text needs in
 * order to compile (the three structs below: only the members the text names, with the types its expressions
 * require), `terra_eval_attribute` for a constant attribute, the retired diffuse callbacks expressed through the
 * reference's LIVE diffuse preset (terra_bsdf_diffuse_init, TerraPresets.c:34-61), the two names the text calls but
 * the file defines under other names (terra_F_0 -> terra_bsdf_R0, terra_fresnel -> terra_bsdf_fresnel_schlick: the
 * only candidates with those signatures), and the adapters at the bottom that present the retired callbacks as live
 * {sample, pdf, eval} ones the way the restatements do: pdf := weight, eval := shade, the state written by sample()
 * kept for the pdf() / eval() calls that follow it on the same surface (terra_trace's order, Terra.c:1068-1075).
 * What is the REFERENCE'S: every arithmetic operation of the two presets. */
#include <math.h>
#include <string.h>
#include "Terra.h"

/* ---- the retired API, as far as the dead text uses it */
typedef struct { TerraFloat3 value; } DeadAttribute;
typedef struct {
    DeadAttribute albedo, roughness, metalness;
    float         ior;
} DeadMaterial;
typedef struct {
    float       roughness, metalness;
    TerraFloat3 half_vector;
    float       fresnel;
} TerraShadingState;
typedef struct {
    TerraFloat2   texcoord;
    TerraFloat3   view, normal;
    TerraFloat4x4 rot;
} TerraShadingContext;

static TerraFloat3 terra_eval_attribute ( const DeadAttribute* attr, const TerraFloat2* uv ) { ( void ) uv; return attr->value; }

/* the retired diffuse callbacks = the live diffuse preset on the surface record the context describes */
void terra_bsdf_diffuse_init ( TerraBSDF* bsdf );          /* live, TerraPresets.c:57-61 (TerraPresets.h is not included: the
                                                              dead text reuses the preset names with the retired signatures) */
static void dead_surface ( const DeadMaterial* m, const TerraShadingContext* ctx, TerraShadingSurface* s ) {
    memset ( s, 0, sizeof *s );
    s->normal = ctx->normal;
    s->transform = ctx->rot;
    s->attributes[0] = m->albedo.value;
}
static TerraFloat3 dead_diffuse_sample ( const DeadMaterial* m, TerraShadingState* state, const TerraShadingContext* ctx, float e1, float e2, float e3 ) {
    TerraBSDF b; TerraShadingSurface s;
    ( void ) state;
    terra_bsdf_diffuse_init ( &b ); dead_surface ( m, ctx, &s );
    return b.sample ( &s, e1, e2, e3, &ctx->view );
}
static float dead_diffuse_weight ( const DeadMaterial* m, TerraShadingState* state, const TerraFloat3* light, const TerraShadingContext* ctx ) {
    TerraBSDF b; TerraShadingSurface s;
    ( void ) state;
    terra_bsdf_diffuse_init ( &b ); dead_surface ( m, ctx, &s );
    return b.pdf ( &s, light, &ctx->view );
}
static TerraFloat3 dead_diffuse_shade ( const DeadMaterial* m, TerraShadingState* state, const TerraFloat3* light, const TerraShadingContext* ctx ) {
    TerraBSDF b; TerraShadingSurface s;
    ( void ) state;
    terra_bsdf_diffuse_init ( &b ); dead_surface ( m, ctx, &s );
    return b.eval ( &s, light, &ctx->view );
}

#define TerraMaterial             DeadMaterial
#define terra_bsdf_diffuse_sample dead_diffuse_sample
#define terra_bsdf_diffuse_weight dead_diffuse_weight
#define terra_bsdf_diffuse_shade  dead_diffuse_shade
#define terra_F_0                 terra_bsdf_R0
#define terra_fresnel             terra_bsdf_fresnel_schlick

/* ---- the reference's dead text (see the header) */
#include TB_DEAD_PRESETS_INC

#undef TerraMaterial
#undef terra_bsdf_diffuse_sample
#undef terra_bsdf_diffuse_weight
#undef terra_bsdf_diffuse_shade

/* ---- live-API adapters. Attribute slots as include/terra_b200.h documents them for the restatements:
 *      GGX [0] albedo [1] (roughness, metalness, ior);  glass [0] albedo [1].x ior. */
static __thread TerraShadingState g_state;

static void ggx_inputs ( const TerraShadingSurface* s, const TerraFloat3* wo, DeadMaterial* m, TerraShadingContext* ctx ) {
    memset ( m, 0, sizeof *m ); memset ( ctx, 0, sizeof *ctx );
    m->albedo.value = s->attributes[0];
    m->roughness.value = terra_f3_set1 ( s->attributes[1].x );
    m->metalness.value = terra_f3_set1 ( s->attributes[1].y );
    m->ior = s->attributes[1].z;
    ctx->view = *wo; ctx->normal = s->normal; ctx->rot = s->transform;
}
static TerraFloat3 a_ggx_sample ( const TerraShadingSurface* s, float e1, float e2, float e3, const TerraFloat3* wo ) {
    DeadMaterial m; TerraShadingContext ctx;
    ggx_inputs ( s, wo, &m, &ctx );
    return terra_bsdf_rough_dielectric_sample ( &m, &g_state, &ctx, e1, e2, e3 );
}
static float a_ggx_pdf ( const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo ) {
    DeadMaterial m; TerraShadingContext ctx;
    ggx_inputs ( s, wo, &m, &ctx );
    return terra_bsdf_rough_dielectric_weight ( &m, &g_state, wi, &ctx );
}
static TerraFloat3 a_ggx_eval ( const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo ) {
    DeadMaterial m; TerraShadingContext ctx;
    ggx_inputs ( s, wo, &m, &ctx );
    return terra_bsdf_rough_dielectric_shade ( &m, &g_state, wi, &ctx );
}
static void glass_inputs ( const TerraShadingSurface* s, const TerraFloat3* wo, DeadMaterial* m, TerraShadingContext* ctx ) {
    memset ( m, 0, sizeof *m ); memset ( ctx, 0, sizeof *ctx );
    m->albedo.value = s->attributes[0];
    m->ior = s->attributes[1].x;
    ctx->view = *wo; ctx->normal = s->normal; ctx->rot = s->transform;
}
static TerraFloat3 a_glass_sample ( const TerraShadingSurface* s, float e1, float e2, float e3, const TerraFloat3* wo ) {
    DeadMaterial m; TerraShadingContext ctx;
    glass_inputs ( s, wo, &m, &ctx );
    return terra_bsdf_glass_sample ( &m, &g_state, &ctx, e1, e2, e3 );
}
static float a_glass_pdf ( const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo ) {
    DeadMaterial m; TerraShadingContext ctx;
    glass_inputs ( s, wo, &m, &ctx );
    return terra_bsdf_glass_weight ( &m, &g_state, wi, &ctx );
}
static TerraFloat3 a_glass_eval ( const TerraShadingSurface* s, const TerraFloat3* wi, const TerraFloat3* wo ) {
    DeadMaterial m; TerraShadingContext ctx;
    glass_inputs ( s, wo, &m, &ctx );
    return terra_bsdf_glass_shade ( &m, &g_state, wi, &ctx );
}

/* Retired {sample, weight, shade} presented as {sample, pdf, eval}: what run_bsdf_callbacks (oracle/binding.py) drives. */
void ref_dead_bsdf_ggx_init ( TerraBSDF* b )   { b->sample = a_ggx_sample;   b->pdf = a_ggx_pdf;   b->eval = a_ggx_eval; }
void ref_dead_bsdf_glass_init ( TerraBSDF* b ) { b->sample = a_glass_sample; b->pdf = a_glass_pdf; b->eval = a_glass_eval; }
