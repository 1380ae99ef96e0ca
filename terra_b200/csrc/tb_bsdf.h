// tb_bsdf.h — the four material presets as {sample, pdf, eval}, host+device.
//   Diffuse, Phong : live reference code, src/TerraPresets.c:34-60 and :66-146.
//   GGX, Glass     : the reference carries these only as dead code written against an older
//                    {sample, weight, shade} API (src/TerraPresets.c:298-394, :396-464). They are
//                    re-expressed here for the live signature: pdf := weight, eval := shade
//                    without its trailing *NoL (terra_trace multiplies NoL itself, Terra.c:1078).
//                    No reference TEST or live code path covers these two; the pin is the dead
//                    code itself, compiled through an adapter of the retired API (oracle/ref_dead_presets.c):
//                    these functions reproduce its sampled directions and the glass weights bit for bit and
//                    the GGX weight / shade to 1e-4 (tests/test_presets_textures.py, tests/golden/bsdf_dead.npz).
// On the device the functions are dispatched by enum after the material sort; on the host they
// back the function pointers that terra_bsdf_*_init installs (classification is by pointer
// identity, include/terra_b200.h) — terra_render never calls the host versions.
#pragma once

#include "tb_math.h"

struct TbSurface {
    M3    xf;          // terra_f4x4_basis(normal)
    V3    n;           // interpolated, normalised shading normal
    V3    emissive;
    V3    a0, a1, a2;  // attribute slots 0..2 (slot meaning depends on the preset)
    float state;       // hidden sample->pdf/eval state: Phong lobe pick (+1/-1), Glass fresnel weight
};

#define TB_INV_PI_F ((float)(1. / TB_PI))      /* `1. / terra_PI` is a double expression, TerraPresets.c:53 */
#define TB_TWO_PI_F (2 * TB_PI)

// ------------------------------------------------------------------------------------ Diffuse
TB_HD V3 tb_diffuse_sample(const TbSurface& s, float e1, float e2) {          // TerraPresets.c:34-45
    float r = sqrtf(e1);
    float theta = 2 * TB_PI * e2;
    float x = r * cosf(theta);
    float z = r * sinf(theta);
    V3 wi = v3(x, sqrtf(tmaxf(0.f, 1 - e1)), z);
    wi = mtransform(s.xf, wi);
    return vnorm(wi);
}
TB_HD float tb_diffuse_pdf(const TbSurface& s, V3 wi) {                       // :47-50
    float NoL = tmaxf(0.f, vdot(s.n, wi));
    return NoL / TB_PI;
}
TB_HD V3 tb_diffuse_eval(V3 albedo) { return vmul(albedo, TB_INV_PI_F); }     // :52-54

// ------------------------------------------------------------------------------------ Phong
// slots: a0 = specular colour, a1 = albedo, a2.x = exponent, state = lobe pick (TerraPresets.h:20-24)
TB_HD void tb_phong_kd_ks(const TbSurface& s, float* kd, float* ks) {         // :66-82
    float diffuse = tmaxf(s.a1.x + s.a1.y + s.a1.z, TB_EPS_F);
    float specular = s.a0.x + s.a0.y + s.a0.z;
    if (specular > diffuse) {
        *kd = 0.5f * diffuse / specular;
        *ks = 1.f - *kd;
    } else {
        *ks = 0.5f * specular / diffuse;
        *kd = 1.f - *ks;
    }
}
TB_HD V3 tb_reflect_about_normal(const TbSurface& s, V3 wo) {                 // :96-97
    V3 wr = vmul(s.n, 2.f * vdot(wo, s.n));
    return vsub(wr, wo);
}
TB_HD V3 tb_phong_sample(TbSurface& s, float e1, float e2, float e3, V3 wo) { // :84-106
    float kd, ks;
    tb_phong_kd_ks(s, &kd, &ks);
    if (e3 < kd) {
        s.state = 1.f;
        return tb_diffuse_sample(s, e1, e2);
    }
    s.state = -1.f;
    V3 wr = tb_reflect_about_normal(s, wo);
    M3 wr_xf = basis_from_normal(wr);
    float phi = 2 * TB_PI * e1;
    float theta = acosf(powf(1.f - e2, 1.f / (s.a2.x + 1)));
    float sin_theta = sinf(theta);
    V3 wi = v3(sin_theta * cosf(phi), cosf(theta), sin_theta * sinf(phi));
    wi = mtransform(wr_xf, wi);
    return vnorm(wi);
}
TB_HD float tb_phong_pdf(const TbSurface& s, V3 wi, V3 wo) {                  // :108-123
    if (s.state == 1.f) return tb_diffuse_pdf(s, wi);
    if (s.state == -1.f) {
        V3 wr = tb_reflect_about_normal(s, wo);
        float cos_alpha = vdot(wi, wr);
        float n = s.a2.x;
        return (n + 1) / (2 * TB_PI) * powf(cos_alpha, n);
    }
    return 0.f;   // reference: assert(false) / falls off the end; unreachable on the live paths
}
TB_HD V3 tb_phong_eval(const TbSurface& s, V3 wi, V3 wo) {                    // :125-140
    float kd, ks;
    tb_phong_kd_ks(s, &kd, &ks);
    float n = s.a2.x;
    V3 diffuse_term = vmul(s.a1, kd * 1.f / TB_PI);
    V3 wr = tb_reflect_about_normal(s, wo);
    float cos_alpha = vdot(wi, wr);
    float cos_n_alpha = powf(cos_alpha, n);
    V3 specular_term = vmul(s.a0, ks * cos_n_alpha * (n + 2) / (2 * TB_PI));
    return vadd(diffuse_term, specular_term);
}

// ------------------------------------------------------------------------------------ GGX (restated)
// slots: a0 = albedo, a1 = (roughness, metalness, ior). Stateless: the old code cached the half
// vector from sample(); for both lobes it equals normalize(wi + wo), which pdf/eval recompute.
TB_HD float tb_ggx_chi(float v) { return v <= 0.f ? 0.f : 1.f; }              // :303-305
TB_HD float tb_ggx_G1(V3 v, V3 n, V3 h, float alpha2) {                       // :307-314
    float VoH = vdot(v, h);
    float VoN = vdot(v, n);
    float chi = tb_ggx_chi(VoH / VoN);
    float VoH2 = VoH * VoH;
    if (VoH2 == 0.f || VoN == 0.f) return 0.f;      // restatement guard: 0/0 and 0*inf would be NaN
    float tan2 = (1.f - VoH2) / VoH2;
    return (chi * 2) / (sqrtf(1 + alpha2 * tan2) + 1);
}
TB_HD float tb_ggx_D(float NoH, float alpha2) {                               // :316-320
    float NoH2 = NoH * NoH;
    float den = NoH2 * alpha2 + (1 - NoH2);
    return (tb_ggx_chi(NoH) * alpha2) / (TB_PI * den * den);
}
TB_HD V3 tb_ggx_sample(const TbSurface& s, float e1, float e2, float e3, V3 wo) {   // :322-345
    const float pd = 1.f - s.a1.y;
    if (e3 <= pd) return tb_diffuse_sample(s, e1, e2);
    float alpha = s.a1.x;
    float theta = atanf((alpha * sqrtf(e1)) / sqrtf(1 - e1));
    float phi = 2 * TB_PI * e2;
    float sin_theta = sinf(theta);
    V3 h = v3(sin_theta * cosf(phi), cosf(theta), sin_theta * sinf(phi));
    h = mtransform(s.xf, h);
    h = vnorm(h);
    float HoV = tmaxf(0.f, vdot(h, wo));
    V3 r = vmul(h, 2 * HoV);
    return vsub(r, wo);
}
// Guard added by the restatement: sample() returns wi = -wo when the sampled half vector faces away
// from wo (HoV clamps to 0, :341-343); wi + wo is then the zero vector and normalising it would
// poison the path with NaN. Such a direction carries no energy: pdf and eval return 0.
TB_HD bool tb_ggx_half(V3 wi, V3 wo, V3* h) {
    V3 hs = vadd(wi, wo);
    if (!(vsqlen(hs) > 0.f)) return false;
    *h = vnorm(hs);
    return true;
}
TB_HD float tb_ggx_pdf(const TbSurface& s, V3 wi, V3 wo) {                    // :347-356
    V3 h;
    if (!tb_ggx_half(wi, wo, &h)) return 0.f;
    float alpha = s.a1.x;
    float alpha2 = alpha * alpha;
    float NoH = vdot(s.n, h);
    float weight_specular = tb_ggx_D(NoH, alpha2) * NoH;
    float weight_diffuse = tb_diffuse_pdf(s, wi);
    const float pd = 1.f - s.a1.y;
    const float ps = 1.f - pd;
    return weight_diffuse * pd + weight_specular * ps;
}
TB_HD V3 tb_ggx_eval(const TbSurface& s, V3 wi, V3 wo) {                      // :358-388 (minus the final *NoL)
    V3 h;
    if (!tb_ggx_half(wi, wo, &h)) return v3s(0.f);
    float metal = s.a1.y;
    float f0 = (1.f - s.a1.z) / (1.f + s.a1.z);                               // :21-27 terra_bsdf_R0
    f0 *= f0;
    V3 F0 = v3(tlerp(f0, s.a0.x, metal), tlerp(f0, s.a0.y, metal), tlerp(f0, s.a0.z, metal));
    float cos_theta = tmaxf(0.f, vdot(wo, h));                                // :14-19 fresnel_schlick
    float p5 = powf(1 - cos_theta, 5);
    V3 ks = vadd(vmul(v3(1 - F0.x, 1 - F0.y, 1 - F0.z), p5), F0);
    float NoL = tmaxf(vdot(s.n, wi), 0.f);
    float NoV = tmaxf(vdot(s.n, wo), 0.f);
    float NoH = tmaxf(vdot(s.n, h), 0.f);
    float alpha = s.a1.x;
    float alpha2 = alpha * alpha;
    float D = tb_ggx_D(NoH, alpha2);
    float G = tb_ggx_G1(wo, s.n, h, alpha2) * tb_ggx_G1(wi, s.n, h, alpha2);
    float den_CT = tminf((4 * NoL * NoV) + 0.05f, 1.f);
    V3 specular_term = vmul(ks, G * D / den_CT);
    V3 diffuse_term = tb_diffuse_eval(s.a0);
    const float pd = 1.f - metal;
    const float ps = 1.f - pd;
    V3 diffuse_factor = vmul(v3(1.f - ks.x, 1.f - ks.y, 1.f - ks.z), (1.f - metal) * pd);
    diffuse_term = vpoint(diffuse_term, diffuse_factor);
    specular_term = vmul(specular_term, ps);
    return vadd(diffuse_term, specular_term);
}

// ------------------------------------------------------------------------------------ Glass (restated)
// slots: a0 = albedo, a1.x = ior, state = fresnel weight left by sample() (0 before any sample on
// this hit, so a delta lobe contributes nothing to explicit light sampling).
TB_HD V3 tb_glass_sample(TbSurface& s, float e3, V3 wo) {                     // :399-449
    V3 normal = s.n;
    V3 incident = vneg(wo);
    float n1, n2;
    float cos_i = vdot(normal, incident);
    if (cos_i > 0.f) {
        n1 = s.a1.x; n2 = 1.f;
        normal = vneg(normal);
    } else {
        n1 = 1.f; n2 = s.a1.x;
        cos_i = -cos_i;
    }
    V3 refl = vmul(normal, 2 * vdot(normal, incident));
    refl = vsub(incident, refl);
    float nni = n1 / n2;
    float cos_t = 1.f - nni * nni * (1.f - cos_i * cos_i);
    if (cos_t < 0.f) { s.state = 1.f; return refl; }
    cos_t = sqrtf(cos_t);
    float t = 1.f - (n1 <= n2 ? cos_i : cos_t);
    float R0 = (n1 - n2) / (n1 + n2);
    R0 *= R0;
    float R = R0 + (1 - R0) * (t * t * t * t * t);
    if (e3 < R) { s.state = R; return refl; }
    V3 trans_v = vmul(normal, (n1 / n2) * cos_i - cos_t);
    V3 trans_n = vmul(incident, n1 / n2);
    V3 trans = vnorm(vadd(trans_v, trans_n));
    s.state = 1 - R;
    return trans;
}
TB_HD float tb_glass_pdf(const TbSurface& s) { return s.state; }              // :451-453
TB_HD V3 tb_glass_eval(const TbSurface& s, V3 wi) {                           // :455-458, divided by the signed NoL
    float NoL = vdot(s.n, wi);                                                //  that terra_trace re-applies
    if (NoL == 0.f) return v3s(0.f);
    return vmul(s.a0, s.state / NoL);
}

// ------------------------------------------------------------------------------------ dispatch
TB_HD V3 tb_bsdf_sample(int preset, TbSurface& s, float e1, float e2, float e3, V3 wo) {
    switch (preset) {
        case 1:  return tb_phong_sample(s, e1, e2, e3, wo);
        case 2:  return tb_ggx_sample(s, e1, e2, e3, wo);
        case 3:  return tb_glass_sample(s, e3, wo);
        default: return tb_diffuse_sample(s, e1, e2);
    }
}
TB_HD float tb_bsdf_pdf(int preset, const TbSurface& s, V3 wi, V3 wo) {
    switch (preset) {
        case 1:  return tb_phong_pdf(s, wi, wo);
        case 2:  return tb_ggx_pdf(s, wi, wo);
        case 3:  return tb_glass_pdf(s);
        default: return tb_diffuse_pdf(s, wi);
    }
}
TB_HD V3 tb_bsdf_eval(int preset, const TbSurface& s, V3 wi, V3 wo) {
    switch (preset) {
        case 1:  return tb_phong_eval(s, wi, wo);
        case 2:  return tb_ggx_eval(s, wi, wo);
        case 3:  return tb_glass_eval(s, wi);
        default: return tb_diffuse_eval(s.a0);
    }
}
