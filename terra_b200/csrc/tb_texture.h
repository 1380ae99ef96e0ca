// tb_texture.h — texture read / point+bilinear sample / lat-long lookup, host+device.
// Follows reference src/Terra.c:368-408 (read), :410-466 (sample), :468-477 (lat-long).
// Texel coordinates are NOT normalised in the reference: uv is used directly as texel index.
// One deliberate difference: the reference's mirror mode can index one past the edge
// (x = width - x % width with x % width == 0, Terra.c:385-386); here that index is clamped.
#pragma once

#include "tb_internal.h"
#include "tb_math.h"

TB_HD size_t tb_mini(size_t a, size_t b) { return a < b ? a : b; }
// x % m and x / m for m <= 65535 (TerraTexture dims are uint16_t): 32-bit arithmetic whenever x fits — the
// 64-bit forms are ~100-instruction subroutines on the GPU and took 12 % of k_shade (ncu n5). Same values.
TB_HD size_t tb_modz(size_t x, uint32_t m) { return x <= 0xffffffffull ? (size_t)((uint32_t)x % m) : x % m; }
TB_HD size_t tb_divz(size_t x, uint32_t m) { return x <= 0xffffffffull ? (size_t)((uint32_t)x / m) : x / m; }

TB_HD V3 tb_texture_read(const TbDevTexture& t, size_t x, size_t y) {
    switch (t.address_mode) {
        case kTerraTextureAddressClamp:
            x = tb_mini(x, (size_t)t.width - 1);
            y = tb_mini(y, (size_t)t.height - 1);
            break;
        case kTerraTextureAddressMirror:
            if (tb_divz(x, t.width) % 2 == 0) {
                x = tb_modz(x, t.width);
                y = tb_modz(y, t.height);
            } else {
                x = tb_mini(t.width - tb_modz(x, t.width), (size_t)t.width - 1);
                y = tb_mini(t.height - tb_modz(y, t.height), (size_t)t.height - 1);
            }
            break;
        default:   // kTerraTextureAddressWrap
            x = tb_modz(x, t.width);
            y = tb_modz(y, t.height);
            break;
    }
    size_t at = (y * t.width + x) * t.components;
    if (t.depth == 1) {
        const uint8_t* p = (const uint8_t*)t.pixels + at;
        return v3(p[0] / 255.f, p[1] / 255.f, p[2] / 255.f);
    }
    const float* p = (const float*)t.pixels + at;
    return v3(p[0], p[1], p[2]);
}

TB_HD V3 tb_texture_sample(const TbDevTexture& t, float u, float v) {
    size_t ix = (size_t)(long long)u;
    size_t iy = (size_t)(long long)v;
    if (t.filter == kTerraFilterPoint) return tb_texture_read(t, ix, iy);
    if (t.filter != kTerraFilterBilinear) return v3s(0.f);     // trilinear/anisotropic: TODO in the reference -> black
    size_t x2 = tb_mini(ix + 1, (size_t)t.width - 1);
    size_t y3 = tb_mini(iy + 1, (size_t)t.height - 1);
    V3 n1 = tb_texture_read(t, ix, iy);
    V3 n2 = tb_texture_read(t, x2, iy);
    V3 n3 = tb_texture_read(t, ix, y3);
    V3 n4 = tb_texture_read(t, x2, y3);
    float w_u = u - ix;
    float w_v = v - iy;
    float w_ou = 1.f - w_u;
    float w_ov = 1.f - w_v;
    return v3((n1.x * w_ou + n2.x * w_u) * w_ov + (n3.x * w_ou + n4.x * w_u) * w_v,
              (n1.y * w_ou + n2.y * w_u) * w_ov + (n3.y * w_ou + n4.y * w_u) * w_v,
              (n1.z * w_ou + n2.z * w_u) * w_ov + (n3.z * w_ou + n4.z * w_u) * w_v);
}

TB_HD V3 tb_texture_sample_latlong(const TbDevTexture& t, V3 dir) {
    V3 d = vnorm(dir);
    float theta = acosf(d.y);
    float phi = atan2f(d.z, d.x) + TB_PI;
    size_t u = (size_t)((phi / (2 * TB_PI)) * t.width);
    size_t v = (size_t)((theta / (TB_PI)) * t.height);
    return tb_texture_read(t, u, v);
}

// terra_attribute_eval (Terra.c:1804-1810): constant unless a texture is bound.
TB_HD V3 tb_attr_eval(const TbDevAttr& a, const TbDevTexture* textures, V2 uv, V3 dir_or_pos) {
    int kind = a.kind_tex & 3;
    if (kind == TB_ATTR_CONST) return v3(a.vx, a.vy, a.vz);
    const TbDevTexture& t = textures[a.kind_tex >> 2];
    if (kind == TB_ATTR_TEX2D) return tb_texture_sample(t, uv.x, uv.y);
    return tb_texture_sample_latlong(t, dir_or_pos);
}
