// tb_bvh8.cpp — the 8-wide tree of the incoherent-ray kernel (tb_trace_w8.cuh), derived on the host from the binary
// tree commit has just built ("may collapse to a wider BVH derived from the reference tree as long as hit
// results match"). Same triangles, same leaf boxes, one triangle per leaf: only the number of dependent node fetches
// a ray makes changes (ncu r2d: 35 % of all stall samples of the binary kernel sit on the first use of the node load,
// ~22 dependent fetches per ray; here ~8).
//
// Collapse: a wide node starts as the two children of a binary node and, while it has fewer than eight, replaces its
// inner child of largest surface area by that child's two children (the greedy surface-area collapse of wide-BVH
// builders). Children are then placed in the eight slots so that slot bit k says on which side of the node's centre
// the child lies along axis k (greedy assignment on centroid . slot direction): a ray visits hit children in the
// order slot ^ octant, which is near-to-far without computing or sorting distances (Ylitie et al. 2017).
//
// Node, 96 bytes = 3 x 256-bit loads (24 words):
//   w0-2   origin of the node's local grid (float3: the node box's min corner)
//   w3-5   cell size per axis, an exact power of two (float3)
//   w6     index of the first inner child (inner children are contiguous, in slot order)
//   w7     rank of the first leaf child's triangle (leaf children's triangles are contiguous in rank, in slot order)
//   w8     imask | lmask << 8      bit s: slot s holds an inner node / a leaf
//   w9-11  unused
//   w12-23 child boxes on the local grid, one byte per slot: lo.x[8] lo.y[8] lo.z[8] hi.x[8] hi.y[8] hi.z[8]
// Boxes round outwards and then move one more cell outwards: the kernel folds the decode into one FMA per plane
// (plane = fma(2^23 + q, cell * inv_dir, ...)), whose constant term rounds by up to half a cell.
//
// Ranks: the triangle / property arrays are ordered by RANK and an exact-t tie goes to the smaller rank. With the
// reference tree the rank is the reference traversal's leaf visiting order (tb_bvh_flatten). For every other tree any
// fixed order will do, and here it is the wide tree's: leaf children of one wide node get consecutive ranks, so a
// leaf's triangle address is tri_base + popcount — no per-child reference in the node.
#include <float.h>
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "tb_bvh.h"    // terra_b200/csrc

// 8-wide tree derived from the binary one (tb_bvh8.cpp): 24 words per node.
struct TbWideBvh {
    std::vector<uint32_t> nodes;
    std::vector<int32_t>  leaf_rank;     // [binary node * 2 + child] -> rank, for tb_bvh_flatten
    int                   depth;         // wide nodes on the deepest root->leaf path
};
void tb_bvh8_build(const std::vector<TbHostNode>& nodes, TbWideBvh* out);

namespace {

struct Entry {            // one child of a wide node: child `c` of binary node `node`
    int   node, c;
    float lo[3], hi[3];
    bool  inner;
    int   bin_child;      // inner: the binary node it stands for
};

Entry entry_of(const std::vector<TbHostNode>& bin, int node, int c) {
    const TbHostNode& n = bin[(size_t)node];
    Entry e;
    e.node = node; e.c = c;
    e.lo[0] = n.aabb[c].min.x; e.lo[1] = n.aabb[c].min.y; e.lo[2] = n.aabb[c].min.z;
    e.hi[0] = n.aabb[c].max.x; e.hi[1] = n.aabb[c].max.y; e.hi[2] = n.aabb[c].max.z;
    e.inner = n.type[c] == -1;
    e.bin_child = e.inner ? n.index[c] : -1;
    return e;
}

double area(const Entry& e) {
    const double dx = (double)e.hi[0] - e.lo[0], dy = (double)e.hi[1] - e.lo[1], dz = (double)e.hi[2] - e.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

}  // namespace

void tb_bvh8_build(const std::vector<TbHostNode>& bin, TbWideBvh* out) {
    out->nodes.clear();
    out->leaf_rank.assign(bin.size() * 2, -1);
    out->depth = 0;
    struct Todo { int bin_node, depth; };
    std::vector<Todo> todo;                 // wide node i stands for binary node todo[i].bin_node (BFS: children contiguous)
    todo.push_back(Todo{ 0, 1 });
    int32_t next_rank = 0;
    for (size_t w = 0; w < todo.size(); ++w) {
        const int src = todo[w].bin_node, depth = todo[w].depth;
        if (depth > out->depth) out->depth = depth;
        // ---- collapse
        Entry ch[8];
        int cnt = 2;
        ch[0] = entry_of(bin, src, 0); ch[1] = entry_of(bin, src, 1);
        while (cnt < 8) {
            int pick = -1; double best = -1.0;
            for (int i = 0; i < cnt; ++i) if (ch[i].inner) { const double a = area(ch[i]); if (a > best) { best = a; pick = i; } }
            if (pick < 0) break;
            const int b = ch[pick].bin_child;
            ch[pick] = entry_of(bin, b, 0);
            ch[cnt++] = entry_of(bin, b, 1);
        }
        // ---- node box, grid
        float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
        for (int i = 0; i < cnt; ++i)
            for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], ch[i].lo[k]); hi[k] = std::max(hi[k], ch[i].hi[k]); }
        float cell[3];
        for (int k = 0; k < 3; ++k) {
            // boxes use cells 1..254 of the axis: cell 0 and 255 are the extra cell each side moves outwards
            const double ext = (double)hi[k] - lo[k];
            int e = ext > 0 ? (int)ceil(log2(ext / 253.0)) : -100;
            if (e < -100) e = -100;
            if (e > 100) e = 100;
            cell[k] = (float)ldexp(1.0, e);
            while ((double)cell[k] * 253.0 < ext) cell[k] *= 2.f;          // log2 rounding
            lo[k] = (float)((double)lo[k] - (double)cell[k]);              // the grid starts one cell below the box
        }
        // ---- slots: greedy assignment maximising centroid . direction of the slot
        double cx[3];
        for (int k = 0; k < 3; ++k) cx[k] = 0.5 * ((double)lo[k] + hi[k]);
        int slot_of[8], in_slot[8];
        for (int i = 0; i < 8; ++i) { slot_of[i] = -1; in_slot[i] = -1; }
        for (int round = 0; round < cnt; ++round) {
            int bi = -1, bs = -1; double bv = -DBL_MAX;
            for (int i = 0; i < cnt; ++i) {
                if (slot_of[i] >= 0) continue;
                for (int s = 0; s < 8; ++s) {
                    if (in_slot[s] >= 0) continue;
                    double v = 0;
                    for (int k = 0; k < 3; ++k) {
                        const double c = 0.5 * ((double)ch[i].lo[k] + ch[i].hi[k]) - cx[k];
                        v += ((s >> k) & 1) ? c : -c;
                    }
                    if (v > bv) { bv = v; bi = i; bs = s; }
                }
            }
            slot_of[bi] = bs; in_slot[bs] = bi;
        }
        // ---- encode
        uint32_t wd[24];
        memset(wd, 0, sizeof wd);
        memcpy(&wd[0], lo, 12);
        memcpy(&wd[3], cell, 12);
        wd[6] = (uint32_t)todo.size();            // first inner child: appended to the BFS queue below, in slot order
        wd[7] = (uint32_t)next_rank;
        uint32_t imask = 0, lmask = 0;
        uint8_t q[6][8];
        memset(q, 0, sizeof q);
        for (int s = 0; s < 8; ++s) {
            const int i = in_slot[s];
            if (i < 0) {                           // empty slot: an inverted box no ray enters
                for (int k = 0; k < 3; ++k) { q[k][s] = 255; q[3 + k][s] = 0; }
                continue;
            }
            for (int k = 0; k < 3; ++k) {
                const double l = lo[k], c = cell[k];
                long a = (long)floor(((double)ch[i].lo[k] - l) / c);
                while (a > 0 && l + a * c > (double)ch[i].lo[k]) --a;
                a -= 1;
                long b = (long)ceil(((double)ch[i].hi[k] - l) / c);
                while (b < 255 && l + b * c < (double)ch[i].hi[k]) ++b;
                b += 1;
                if (a < 0) a = 0;
                if (b > 255) b = 255;
                q[k][s] = (uint8_t)a; q[3 + k][s] = (uint8_t)b;
            }
            if (ch[i].inner) {
                imask |= 1u << s;
                todo.push_back(Todo{ ch[i].bin_child, depth + 1 });
            } else {
                lmask |= 1u << s;
                out->leaf_rank[(size_t)ch[i].node * 2 + ch[i].c] = next_rank++;
            }
        }
        wd[8] = imask | (lmask << 8);
        memcpy(&wd[12], q, 48);
        out->nodes.insert(out->nodes.end(), wd, wd + 24);
    }
}
