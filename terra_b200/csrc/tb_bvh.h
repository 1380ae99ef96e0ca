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

void tb_bvh_build(const TerraObject* objects, int objects_count, std::vector<TbHostNode>& nodes);
// Conventional binned-SAH tree over x, y and z in the same node format (see tb_bvh.cpp).
void tb_bvh_build_sah3(const TerraObject* objects, int objects_count, std::vector<TbHostNode>& nodes);
// leaf_rank: NULL = rank by the reference traversal's leaf visiting order; else the rank of child c of node n at
// [n * 2 + c] (any fixed order gives the same hits outside exact-t ties).
void tb_bvh_flatten(const std::vector<TbHostNode>& nodes, const TerraObject* objects, int objects_count, TbFlatBvh* out,
                    const std::vector<int32_t>* leaf_rank = NULL);

