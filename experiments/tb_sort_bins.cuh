// experiments/tb_sort_bins.cuh — NOT part of the build (see experiments/README.md).
#define TB_SORT_BINS 32
// ---- direction binning of the next bounce's queue (TERRA_B200_SORT=1|2; A/B option, off by default)
// The rays a shade pass emits sit in the queue in slot order: a warp of the next trace launch gets 32 rays
// leaving (nearly) one point in 32 unrelated directions. Reordering the queue ids — the rays stay where they
// are — so that a warp's rays share their direction octant (mode 1) or octant and dominant axis (mode 2)
// makes the lanes agree on near / far children. Two passes over the ids: histogram, then scatter at
// warp-aggregated cursors. Results do not depend on the order of the queue.
__device__ __forceinline__ int tb_dir_bin(float4 d, int mode) {
    int key = (d.x < 0.f ? 1 : 0) | (d.y < 0.f ? 2 : 0) | (d.z < 0.f ? 4 : 0);
    if (mode >= 2) {
        const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
        const int major = ax >= ay ? (ax >= az ? 0 : 2) : (ay >= az ? 1 : 2);
        key |= major << 3;
    }
    return key;
}
__global__ void __launch_bounds__(256) k_bin_count(const int* __restrict__ q, const int* __restrict__ q_count,
                                                   const float4* __restrict__ ray_d, int mode, int* __restrict__ bins) {
    __shared__ int s_bins[TB_SORT_BINS];
    if (threadIdx.x < TB_SORT_BINS) s_bins[threadIdx.x] = 0;
    __syncthreads();
    const int count = *q_count;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
        atomicAdd(&s_bins[tb_dir_bin(ray_d[q[i]], mode)], 1);
    __syncthreads();
    if (threadIdx.x < TB_SORT_BINS && s_bins[threadIdx.x]) atomicAdd(&bins[threadIdx.x], s_bins[threadIdx.x]);
}
__global__ void __launch_bounds__(256) k_bin_scatter(const int* __restrict__ q, const int* __restrict__ q_count,
                                                     const float4* __restrict__ ray_d, int mode, int* __restrict__ bins,
                                                     int* __restrict__ out) {
    __shared__ int s_base[TB_SORT_BINS];
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int b = 0; b < TB_SORT_BINS; ++b) { s_base[b] = acc; acc += bins[b]; }
    }
    __syncthreads();
    int* cursors = bins + TB_SORT_BINS;
    const int count = *q_count;
    const int stride = gridDim.x * blockDim.x;
    const int rounds = (count + stride - 1) / stride;
    const unsigned lane = threadIdx.x & 31u;
    for (int it = 0; it < rounds; ++it) {
        const int i = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        const bool valid = i < count;
        int slot = 0, key = TB_SORT_BINS;                 // invalid lanes share a dummy key
        if (valid) { slot = q[i]; key = tb_dir_bin(ray_d[slot], mode); }
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        const int leader = __ffs(peers) - 1;
        int base = 0;
        if (valid && (int)lane == leader) base = atomicAdd(&cursors[key], __popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (valid) out[s_base[key] + base + __popc(peers & ((1u << lane) - 1u))] = slot;
    }
}

