// tb_math.h — vector math with the reference's operation ORDER (include/TerraMath.inl), usable from
// host and device. Bit-exact agreement with the reference's x86-64 build (which emits no FMA)
// needs this file compiled with -fmad=false on the device side and IEEE div/sqrt (nvcc defaults).
#pragma once

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define TB_HD __host__ __device__ __forceinline__
#else
#define TB_HD static inline
#endif

#define TB_PI       3.1416926535f          /* reference terra_PI, include/TerraMath.h:17 (not pi) */
#define TB_EPS_D    1e-4                   /* reference terra_Epsilon: a DOUBLE literal, :18 */
#define TB_EPS_F    ((float)1e-4)

struct V3 { float x, y, z; };
struct V2 { float x, y; };

TB_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
TB_HD V3 v3s(float s) { return v3(s, s, s); }
TB_HD V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
TB_HD V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
TB_HD V3 vmul(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
TB_HD V3 vdiv(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
TB_HD V3 vpoint(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
TB_HD V3 vneg(V3 a) { return v3(-a.x, -a.y, -a.z); }
TB_HD float vdot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }            // TerraMath.inl:108
TB_HD V3 vcross(V3 a, V3 b) {                                                            // :112
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
TB_HD float vlen(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }            // :126
TB_HD float vsqlen(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
TB_HD V3 vnorm(V3 a) { float l = vlen(a); return v3(a.x / l, a.y / l, a.z / l); }     // :150 (divides, no rsqrt)
TB_HD float vsqdist(V3 a, V3 b) { V3 d = vsub(b, a); return vdot(d, d); }
// terra_minf / terra_maxf are ternaries, NOT fminf/fmaxf: NaN propagates through the second
// operand (TerraMath.inl:171-177). Keep the operand order of every call site.
TB_HD float tmaxf(float a, float b) { return a > b ? a : b; }
TB_HD float tminf(float a, float b) { return a < b ? a : b; }
TB_HD float tclamp(float v, float lo, float hi) { return v < lo ? lo : v > hi ? hi : v; }
TB_HD float tlerp(float a, float b, float t) { return a + (b - a) * t; }
TB_HD uint32_t fbits(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
TB_HD float bitsf(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

// 3x3 part of the reference's row-major 4x4: r[i] = rows[i].xyz; transform = (dot(r0,v), dot(r1,v), dot(r2,v)).
struct M3 { V3 r0, r1, r2; };
TB_HD V3 mtransform(const M3& m, V3 v) { return v3(vdot(m.r0, v), vdot(m.r1, v), vdot(m.r2, v)); }   // :234

// terra_f4x4_basis (TerraMath.inl:251-272): Hughes-Moeller frame whose tangent is MULTIPLIED by
// sqrt(..) instead of divided — reproduced, not fixed. Columns = tangent|normal|bitangent.
TB_HD M3 basis_from_normal(V3 n) {
    V3 t;
    if (fabsf(n.x) > fabsf(n.y)) {
        t = v3(n.z, 0.f, -n.x);
        t = vmul(t, sqrtf(n.x * n.x + n.z * n.z));
    } else {
        t = v3(0.f, -n.z, n.y);
        t = vmul(t, sqrtf(n.y * n.y + n.z * n.z));
    }
    V3 b = vcross(n, t);
    M3 m;
    m.r0 = v3(t.x, n.x, b.x);
    m.r1 = v3(t.y, n.y, b.y);
    m.r2 = v3(t.z, n.z, b.z);
    return m;
}
