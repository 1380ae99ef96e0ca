// tb_distribution.cuh — the reference's discrete distributions (src/Terra.c:760-846, src/TerraPrivate.h:85-102)
// on the device, and the environment importance map built from them.
//
// The reference writes TerraDistribution1D / TerraDistributon2D and calls them from nowhere. Here they
// drive next-event estimation of a lat-long environment (terra_b200_set_env_lit(scene, 2)).
//
// Same numbers as the reference's functions, by construction:
//   init:   a row's cdf is a RUNNING float sum in index order, then every entry is divided by the total. A
//           parallel scan would round differently, so one thread owns one row (rows are independent: the
//           grid is the ny rows); the marginal over the row totals is one more such row.
//   sample: the reference scans for the first bucket with e < cdf[i]; the cdf is non-decreasing (a sum of
//           non-negative terms divided by one positive number), so a binary search finds the same bucket.
#pragma once

// terra_distribution_1d_sample, Terra.c:781-809. FLT_MAX when no bucket holds e (the reference asserts).
__device__ __forceinline__ float tb_dist1d_sample(const float* __restrict__ f, const float* __restrict__ cdf, float integral,
                                                  int n, float e, float* pdf, int* idx) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (e < __ldg(cdf + mid)) hi = mid; else lo = mid + 1;
    }
    if (lo >= n) { *pdf = 0.f; *idx = n - 1; return FLT_MAX; }
    const float prev = lo ? __ldg(cdf + lo - 1) : 0.f;
    const float curr = __ldg(cdf + lo);
    *pdf = __ldg(f + lo) / integral;
    *idx = lo;
    float d = e - prev;
    d = d / (curr - prev);
    return ((float)lo + d) / (float)n;
}

// terra_distribution_2d_sample, Terra.c:835-846: returns (s1 = row axis, s2 = column axis), pdf = product of
// the two bucket probabilities.
__device__ __forceinline__ float2 tb_dist2d_sample(const TbDevDist2D& d, float e1, float e2, float* pdf, int* row, int* col) {
    float p0, p1;
    const float s1 = tb_dist1d_sample(d.row_integral, d.marginal_cdf, d.integral, d.ny, e1, &p0, row);
    const size_t at = (size_t)(*row) * d.nx;
    const float s2 = tb_dist1d_sample(d.f + at, d.cdf + at, __ldg(d.row_integral + *row), d.nx, e2, &p1, col);
    *pdf = p0 * p1;
    return make_float2(s1, s2);
}

// terra_distribution_1d_init (Terra.c:760-779) for `rows` independent rows of n: thread r owns row r.
// total_out[r] = the row's integral.
__global__ void __launch_bounds__(128) k_dist_rows(const float* __restrict__ f, int n, int rows, float* __restrict__ cdf, float* __restrict__ total_out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const float* fr = f + (size_t)r * n;
    float* cr = cdf + (size_t)r * n;
    float integral = 0.f;
    for (int i = 0; i < n; ++i) {
        integral = integral + fr[i];
        cr[i] = integral;
    }
    for (int i = 0; i < n; ++i) cr[i] = cr[i] / integral;
    total_out[r] = integral;
}

__global__ void __launch_bounds__(256) k_dist_sample(TbDevDist2D d, size_t n, const float* __restrict__ e1, const float* __restrict__ e2, float4* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float pdf; int row, col;
        const float2 s = tb_dist2d_sample(d, e1[i], e2[i], &pdf, &row, &col);
        out[i] = make_float4(s.x, s.y, pdf, 0.f);
    }
}

// Importance of the lat-long texels: luminance (Terra.c:1036 weights) of the texel terra_texture_sample_latlong
// would read, times the row's solid-angle weight (sin(theta) at the row centre, computed on the host with the
// same libm as the oracle and handed in as row_weight).
__global__ void __launch_bounds__(256) k_env_importance(TbDevTexture t, const float* __restrict__ row_weight, float* __restrict__ f) {
    const size_t n = (size_t)t.width * t.height;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t y = i / t.width, x = i - y * t.width;
        const V3 c = tb_texture_read(t, x, y);
        const float lum = 0.212671f * c.x + 0.715160f * c.y + 0.072169f * c.z;
        f[i] = lum * row_weight[y];
    }
}

// ---- lat-long parametrisation, the inverse of terra_texture_sample_latlong (Terra.c:468-477):
//      theta = acos(d.y) = v * pi, phi = atan2(d.z, d.x) + pi = u * 2 pi   (pi = terra_PI as the reference has it)
__device__ __forceinline__ V3 tb_latlong_dir(float u, float v) {
    const float theta = v * TB_PI, phi = u * (2 * TB_PI) - TB_PI;
    const float st = sinf(theta);
    return v3(st * cosf(phi), cosf(theta), st * sinf(phi));
}
// Solid-angle density of picking direction d from the map: bucket probability * nx * ny / (2 pi^2 sin(theta)).
__device__ __forceinline__ float tb_env_pdf(const TbDevDist2D& d, V3 dir) {
    const V3 n = vnorm(dir);
    const float theta = acosf(n.y);
    const float phi = atan2f(n.z, n.x) + TB_PI;
    size_t u = (size_t)((phi / (2 * TB_PI)) * d.nx);
    size_t v = (size_t)((theta / (TB_PI)) * d.ny);
    if (u > (size_t)d.nx - 1) u = (size_t)d.nx - 1;
    if (v > (size_t)d.ny - 1) v = (size_t)d.ny - 1;
    const float st = sinf(theta);
    if (!(st > 0.f)) return 0.f;
    const float p_uv = __ldg(d.f + v * d.nx + u) / d.integral * (float)d.nx * (float)d.ny;
    return p_uv / (2.f * TB_PI * TB_PI * st);
}

// One explicit environment sample: direction + solid-angle density. false when the map has nothing to offer
// for these numbers (a zero bucket, or the pole where the parametrisation collapses).
__device__ __forceinline__ bool tb_env_sample(const TbDevDist2D& d, float e1, float e2, V3* dir, float* pdf) {
    float p; int row, col;
    const float2 s = tb_dist2d_sample(d, e1, e2, &p, &row, &col);
    if (!(p > 0.f) || s.x == FLT_MAX || s.y == FLT_MAX) return false;
    const float v = fminf(s.x, 0.99999994f), u = fminf(s.y, 0.99999994f);
    const float st = sinf(v * TB_PI);
    if (!(st > 0.f)) return false;
    *dir = tb_latlong_dir(u, v);
    *pdf = p * (float)d.nx * (float)d.ny / (2.f * TB_PI * TB_PI * st);
    return true;
}
