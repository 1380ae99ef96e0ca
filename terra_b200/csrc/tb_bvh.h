// tb_bvh.h — host BVH build (reference routine restated) and flatten to the GPU layout.
#pragma once

#include <vector>

#include "tb_internal.h"

struct TbFlatBvh {
    std::vector<float>    nodes;              // 16 floats per node
    std::vector<float>    tris;               // 12 floats per rank
    std::vector<float>    props;              // 16 floats per rank
    std::vector<uint32_t> rank_of_triangle;   // [object_base[obj] + tri] -> rank
    std::vector<size_t>   object_base;
    int                   stack_bound;        // inner nodes on the deepest root->leaf path
};

struct TbQuantBvh {
    std::vector<uint32_t> nodes;              // 8 words per node (tb_trace.cuh)
    float grid_lo[3], grid_cell[3];
};
void tb_bvh_quantize(const TbFlatBvh& flat, TbQuantBvh* out);

// 4-wide tree of the incoherent-ray kernel k_trace_v7 (tb_trace_v7.cuh), derived from the binary tree: same child boxes
// (the quantised words are COPIED from the binary nodes, so every box a ray is tested against is one it would also meet
// in the binary walk), same leaves, same ranks. 16 words per node: word 3 c + k = lo | hi << 16 of child c on axis k
// (c = 0..3), words 12..15 = child refs (>= 0: wide node index, < 0: leaf ~rank; empty slot: an inverted box).
struct TbQuadBvh {
    std::vector<uint32_t> nodes;              // 16 words per node
    std::vector<int32_t>  src;                // 4 per node: binary node * 2 + child the slot's box comes from, or -1 (refit re-copies them)
    int                   stack_bound;        // worst-case stack entries of a near-first walk: sum over a path of (children - 1)
};
void tb_bvh4_build(const TbFlatBvh& flat, const TbQuantBvh& quant, TbQuadBvh* out);

void tb_bvh_build(const TerraObject* objects, int objects_count, std::vector<TbHostNode>& nodes);
// Conventional binned-SAH tree over x, y and z in the same node format (see tb_bvh.cpp).
void tb_bvh_build_sah3(const TerraObject* objects, int objects_count, std::vector<TbHostNode>& nodes);
// leaf_rank: NULL = rank by the reference traversal's leaf visiting order; else the rank of child c of node n at
// [n * 2 + c] (any fixed order gives the same hits outside exact-t ties).
void tb_bvh_flatten(const std::vector<TbHostNode>& nodes, const TerraObject* objects, int objects_count, TbFlatBvh* out,
                    const std::vector<int32_t>* leaf_rank = NULL);

