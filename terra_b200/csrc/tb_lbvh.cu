// tb_lbvh.cu — BVH build ON THE GPU, replacing terra_bvh_create (reference
// src/TerraBVH.c:128-244) for scenes where the host build dominates the time to the first pixel
// (reference builder: 31 s at 10 M triangles, single-threaded, O(N log^2 N)).
//
// Linear BVH (Karras 2012): Morton-order the triangle centroids, derive the N-1 inner nodes of the
// binary radix tree over the sorted keys in parallel, then fit the boxes bottom-up. The output is the
// SAME 64-byte node the reference builds (TbHostNode == TerraBVHNode: two child boxes, two child
// indices, two child types), root at index 0, one triangle per leaf — so the flattener, the quantiser
// and every trace kernel take it unchanged, and the closest hit is the reference's (the hit depends on
// the triangles and the intersection arithmetic, not on the tree; tests/test_gpu_lbvh.py).
// Tree quality is below the binned-SAH host build (more nodes per ray): this is the fast-build mode,
// TERRA_B200_BVH=lbvh / terra_b200_set_bvh_mode(2), not the default.
//
// Kernels: k_lbvh_boxes (padded leaf boxes + scene bounds), k_lbvh_keys (63-bit Morton codes),
// cub radix sort (library plumbing, not the hot path), k_lbvh_hierarchy, k_lbvh_fit.
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>

#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "tb_internal.h"

#define LB_CUDA(call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            terra_log("terra_b200: CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            abort();                                                                               \
        }                                                                                          \
    } while (0)

struct LbBox { float lo[3], hi[3]; };

// Same conservative padding as the host sah3 builder (tb_bvh.cpp: the reference pads leaves by 1e-4,
// Terra.c:982-996; the relative term keeps the pad above one ulp for large coordinates).
__device__ __forceinline__ float lb_pad_lo(float v) { return v - (1e-4f + 2e-6f * fabsf(v)); }
__device__ __forceinline__ float lb_pad_hi(float v) { return v + (1e-4f + 2e-6f * fabsf(v)); }

// floats ordered as unsigned ints (for atomicMin/atomicMax on the scene bounds)
__device__ __forceinline__ unsigned lb_order(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ __forceinline__ float lb_unorder(unsigned u) {
    unsigned v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f; memcpy(&f, &v, 4); return f;
}

__global__ void __launch_bounds__(256) k_lbvh_boxes(const float* tri9, int n, LbBox* boxes, float* centroid3, unsigned* bounds6) {
    __shared__ unsigned s_lo[3], s_hi[3];
    if (threadIdx.x < 3) { s_lo[threadIdx.x] = 0xffffffffu; s_hi[threadIdx.x] = 0u; }
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float* t = tri9 + (size_t)i * 9;
        LbBox b;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float a = t[k], bb = t[3 + k], c = t[6 + k];
            const float lo = fminf(a, fminf(bb, c)), hi = fmaxf(a, fmaxf(bb, c));
            b.lo[k] = lb_pad_lo(lo); b.hi[k] = lb_pad_hi(hi);
            const float mid = 0.5f * lo + 0.5f * hi;
            centroid3[(size_t)i * 3 + k] = mid;
            atomicMin(&s_lo[k], lb_order(mid));
            atomicMax(&s_hi[k], lb_order(mid));
        }
        boxes[i] = b;
    }
    __syncthreads();
    if (threadIdx.x < 3) { atomicMin(&bounds6[threadIdx.x], s_lo[threadIdx.x]); atomicMax(&bounds6[3 + threadIdx.x], s_hi[threadIdx.x]); }
}

__device__ __forceinline__ uint64_t lb_spread21(uint64_t v) {      // 21 bits -> every third bit
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8)  & 0x100f00f00f00f00full;
    v = (v | v << 4)  & 0x10c30c30c30c30c3ull;
    v = (v | v << 2)  & 0x1249249249249249ull;
    return v;
}

__global__ void __launch_bounds__(256) k_lbvh_keys(const float* centroid3, int n, const unsigned* bounds6, uint64_t* keys, uint32_t* vals) {
    float lo[3], inv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        lo[k] = lb_unorder(bounds6[k]);
        const float ext = lb_unorder(bounds6[3 + k]) - lo[k];
        inv[k] = ext > 0.f ? 2097151.f / ext : 0.f;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint64_t code = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float q = (centroid3[(size_t)i * 3 + k] - lo[k]) * inv[k];
            q = fminf(fmaxf(q, 0.f), 2097151.f);
            code |= lb_spread21((uint64_t)q) << (2 - k);
        }
        keys[i] = code;
        vals[i] = (uint32_t)i;
    }
}

// Karras 2012, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees": inner node
// i of the binary radix tree over sorted keys. delta(i,j) = length of the common prefix of keys i and j,
// with the index as tie-break for equal keys (so duplicate centroids still form a valid tree).
__device__ __forceinline__ int lb_delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

// child refs: >= 0 inner node index, < 0 leaf holding ~(sorted position)
__global__ void __launch_bounds__(256) k_lbvh_hierarchy(const uint64_t* keys, int n, int2* children, int* parent_inner, int* parent_leaf) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x) {
        const int d = lb_delta(keys, n, i, i + 1) - lb_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
        const int dmin = lb_delta(keys, n, i, i - d);
        int lmax = 2;
        while (lb_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
        int l = 0;
        for (int t = lmax / 2; t >= 1; t /= 2)
            if (lb_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
        const int j = i + l * d;
        const int dnode = lb_delta(keys, n, i, j);
        int s = 0;
        for (int t = (l + 1) / 2; ; t = (t + 1) / 2) {          // ceil(l / 2^k)
            if (lb_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
            if (t == 1) break;
        }
        const int gamma = i + s * d + min(d, 0);
        const int lo = min(i, j), hi = max(i, j);
        int2 c;
        if (lo == gamma)     { c.x = ~gamma;       parent_leaf[gamma] = i * 2; }
        else                 { c.x = gamma;        parent_inner[gamma] = i * 2; }
        if (hi == gamma + 1) { c.y = ~(gamma + 1); parent_leaf[gamma + 1] = i * 2 + 1; }
        else                 { c.y = gamma + 1;    parent_inner[gamma + 1] = i * 2 + 1; }
        children[i] = c;
    }
}

__device__ __forceinline__ TerraAABB lb_to_aabb(const LbBox& b) {
    TerraAABB a;
    a.min.x = b.lo[0]; a.min.y = b.lo[1]; a.min.z = b.lo[2];
    a.max.x = b.hi[0]; a.max.y = b.hi[1]; a.max.z = b.hi[2];
    return a;
}

// Bottom-up fit: one thread per leaf climbs; the second thread to arrive at a node owns both child boxes,
// merges them and carries the union up. Writes the reference's node: the boxes live in the PARENT.
__global__ void __launch_bounds__(256) k_lbvh_fit(const LbBox* boxes, const uint32_t* sorted_vals, const uint32_t* prim, int n,
                                                  const int2* children, const int* parent_inner, const int* parent_leaf,
                                                  int* arrived, TbHostNode* nodes) {
    for (int leaf = blockIdx.x * blockDim.x + threadIdx.x; leaf < n; leaf += gridDim.x * blockDim.x) {
        const uint32_t tri = sorted_vals[leaf];
        LbBox b = boxes[tri];
        int link = parent_leaf[leaf];                       // parent * 2 + child slot
        int32_t index = (int32_t)prim[tri], type = 1;
        while (true) {
            const int p = link >> 1, c = link & 1;
            nodes[p].aabb[c] = lb_to_aabb(b);
            nodes[p].index[c] = index;
            nodes[p].type[c] = type;
            __threadfence();
            if (atomicAdd(&arrived[p], 1) == 0) break;      // first to arrive: the sibling's thread goes on
            __threadfence();
            const float* sib = (const float*)&nodes[p].aabb[c ^ 1];     // written by another SM: read through L2
            TerraAABB o;
            o.min.x = __ldcg(sib); o.min.y = __ldcg(sib + 1); o.min.z = __ldcg(sib + 2);
            o.max.x = __ldcg(sib + 3); o.max.y = __ldcg(sib + 4); o.max.z = __ldcg(sib + 5);
            b.lo[0] = fminf(b.lo[0], o.min.x); b.lo[1] = fminf(b.lo[1], o.min.y); b.lo[2] = fminf(b.lo[2], o.min.z);
            b.hi[0] = fmaxf(b.hi[0], o.max.x); b.hi[1] = fmaxf(b.hi[1], o.max.y); b.hi[2] = fmaxf(b.hi[2], o.max.z);
            if (p == 0) break;                              // root
            link = parent_inner[p];
            index = p; type = -1;
        }
    }
}

template <typename T> static T* lb_alloc(size_t n) { T* p = NULL; LB_CUDA(cudaMalloc((void**)&p, (n ? n : 1) * sizeof(T))); return p; }

// Builds the tree for all triangles of `objects` on device `device_index` and returns the reference-layout
// node array (n - 1 nodes, root first). Timings (ms) of the device part and of the whole call go to `ms_out[2]`.
extern "C" void tb_lbvh_build(int device_index, const TerraObject* objects, int objects_count, std::vector<TbHostNode>* nodes_out, float* ms_out) {
    LB_CUDA(cudaSetDevice(device_index));
    size_t total = 0;
    for (int j = 0; j < objects_count; ++j) total += objects[j].triangles_count;
    const int n = (int)total;
    nodes_out->assign(total > 1 ? total - 1 : 0, TbHostNode());
    if (n < 2) return;
    cudaEvent_t e0, e1, e2;
    LB_CUDA(cudaEventCreate(&e0)); LB_CUDA(cudaEventCreate(&e1)); LB_CUDA(cudaEventCreate(&e2));
    LB_CUDA(cudaEventRecord(e0));
    float* d_tri = lb_alloc<float>(total * 9);
    uint32_t* d_prim = lb_alloc<uint32_t>(total);
    {
        std::vector<uint32_t> prim(total);
        size_t at = 0;
        for (int j = 0; j < objects_count; ++j) {
            const size_t m = objects[j].triangles_count;
            static_assert(sizeof(TerraTriangle) == 36, "TerraTriangle is 9 floats");
            LB_CUDA(cudaMemcpy(d_tri + at * 9, objects[j].triangles, m * 36, cudaMemcpyHostToDevice));
            for (size_t i = 0; i < m; ++i) prim[at + i] = (uint32_t)j | ((uint32_t)i << 8);      // TerraBVH.c:155
            at += m;
        }
        LB_CUDA(cudaMemcpy(d_prim, prim.data(), total * 4, cudaMemcpyHostToDevice));
    }
    LB_CUDA(cudaEventRecord(e1));
    cudaDeviceProp prop;
    LB_CUDA(cudaGetDeviceProperties(&prop, device_index));
    const int grid = prop.multiProcessorCount * 8;
    LbBox* d_boxes = lb_alloc<LbBox>(total);
    float* d_centroid = lb_alloc<float>(total * 3);
    unsigned* d_bounds = lb_alloc<unsigned>(6);
    const unsigned init[6] = { 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u };
    LB_CUDA(cudaMemcpy(d_bounds, init, sizeof init, cudaMemcpyHostToDevice));
    uint64_t* d_keys = lb_alloc<uint64_t>(total); uint64_t* d_keys2 = lb_alloc<uint64_t>(total);
    uint32_t* d_vals = lb_alloc<uint32_t>(total); uint32_t* d_vals2 = lb_alloc<uint32_t>(total);
    k_lbvh_boxes<<<grid, 256>>>(d_tri, n, d_boxes, d_centroid, d_bounds);
    k_lbvh_keys<<<grid, 256>>>(d_centroid, n, d_bounds, d_keys, d_vals);
    size_t tmp_bytes = 0;
    LB_CUDA(cub::DeviceRadixSort::SortPairs(NULL, tmp_bytes, d_keys, d_keys2, d_vals, d_vals2, n, 0, 63));
    void* d_tmp = NULL;
    LB_CUDA(cudaMalloc(&d_tmp, tmp_bytes ? tmp_bytes : 1));
    LB_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_keys, d_keys2, d_vals, d_vals2, n, 0, 63));
    int2* d_children = lb_alloc<int2>(total);
    int* d_parent_inner = lb_alloc<int>(total);
    int* d_parent_leaf = lb_alloc<int>(total);
    int* d_arrived = lb_alloc<int>(total);
    TbHostNode* d_nodes = lb_alloc<TbHostNode>(total);
    LB_CUDA(cudaMemset(d_arrived, 0, total * sizeof(int)));
    k_lbvh_hierarchy<<<grid, 256>>>(d_keys2, n, d_children, d_parent_inner, d_parent_leaf);
    k_lbvh_fit<<<grid, 256>>>(d_boxes, d_vals2, d_prim, n, d_children, d_parent_inner, d_parent_leaf, d_arrived, d_nodes);
    LB_CUDA(cudaGetLastError());
    LB_CUDA(cudaEventRecord(e2));
    LB_CUDA(cudaMemcpy(nodes_out->data(), d_nodes, (total - 1) * sizeof(TbHostNode), cudaMemcpyDeviceToHost));
    if (ms_out) {
        LB_CUDA(cudaEventElapsedTime(&ms_out[0], e1, e2));      // kernels + sort
        LB_CUDA(cudaEventElapsedTime(&ms_out[1], e0, e2));      // + triangle upload
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    cudaFree(d_tri); cudaFree(d_prim); cudaFree(d_boxes); cudaFree(d_centroid); cudaFree(d_bounds);
    cudaFree(d_keys); cudaFree(d_keys2); cudaFree(d_vals); cudaFree(d_vals2); cudaFree(d_tmp);
    cudaFree(d_children); cudaFree(d_parent_inner); cudaFree(d_parent_leaf); cudaFree(d_arrived); cudaFree(d_nodes);
}
