// tb_bvh.cpp — host BVH: the reference's builder restated (src/TerraBVH.c:24-244 and
// src/Terra.c:972-997), plus the flatten into the GPU layout (tb_internal.h).
//
// The tree must be THE reference's tree (north_star: "the reference BVH build from TerraBVH.c,
// flattened into a GPU node/triangle layout"), because which triangles a ray may report depends on
// which boxes it passes. What decides the tree shape is reproduced operation by
// operation: the leaf padding (double 1e-4 subtracted/added then rounded to float), the +1e-4
// on every max-merge (Q11), the x-only sort with a boolean "comparator" handed to libc qsort
// (Q12: the order then depends on glibc's algorithm, so the same qsort is called on the same
// 32-byte elements), the float SAH cost expression, and the LIFO task order that numbers nodes.
// tests/test_abi_host.py::test_host_bvh_reference_mode_is_byte_identical checks the node array byte-for-byte against terra_bvh_create.
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>

#include "tb_bvh.h"
#include "tb_math.h"

// ------------------------------------------------------------------------------------ host threads
// The commit stages below (sah3 build, flatten, quantise) are loops over millions of independent
// primitives / nodes: they are split over the host cores in contiguous chunks. Results never depend on the
// thread count (exact min/max merges, disjoint writes).
namespace {

int g_par_min = 1 << 16;        // ranges below this are handled by one thread (TERRA_B200_BUILD_PAR_MIN: tests lower it)

template <typename F> void parallel_chunks(int begin, int end, int threads, F fn) {     // fn(lo, hi, tid)
    const int n = end - begin;
    if (threads <= 1 || n < g_par_min) { fn(begin, end, 0); return; }
    std::vector<std::thread> pool;
    const int per = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        const int lo = begin + t * per, hi = std::min(end, lo + per);
        if (lo >= hi) break;
        pool.emplace_back([=]() { fn(lo, hi, t); });
    }
    for (std::thread& th : pool) th.join();
}

int host_threads() {
    int threads = (int)std::thread::hardware_concurrency();
    if (const char* e = getenv("TERRA_B200_BUILD_THREADS")) threads = atoi(e);
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    g_par_min = 1 << 16;
    if (const char* e = getenv("TERRA_B200_BUILD_PAR_MIN")) g_par_min = std::max(2, atoi(e));
    return threads;
}

}  // namespace


namespace {

struct Volume {            // 32 bytes, like the reference's TerraBVHVolume (TerraBVH.c:11-15)
    TerraAABB box;
    uint32_t  index;
    int32_t   type;
};
static_assert(sizeof(Volume) == 32, "qsort element size must match the reference");

inline TerraAABB empty_box() {
    TerraAABB b;
    b.min.x = b.min.y = b.min.z = FLT_MAX;
    b.max.x = b.max.y = b.max.z = -FLT_MAX;
    return b;
}

// terra_aabb_fit_triangle, Terra.c:972-997. `-= 1e-4` is evaluated in double, as in the reference.
inline void grow_by_triangle(TerraAABB* b, const TerraTriangle* t) {
    b->min.x = tminf(tminf(tminf(b->min.x, t->a.x), t->b.x), t->c.x);
    b->min.y = tminf(tminf(tminf(b->min.y, t->a.y), t->b.y), t->c.y);
    b->min.z = tminf(tminf(tminf(b->min.z, t->a.z), t->b.z), t->c.z);
    b->min.x -= TB_EPS_D;
    b->min.y -= TB_EPS_D;
    b->min.z -= TB_EPS_D;
    b->max.x = tmaxf(tmaxf(tmaxf(b->max.x, t->a.x), t->b.x), t->c.x);
    b->max.y = tmaxf(tmaxf(tmaxf(b->max.y, t->a.y), t->b.y), t->c.y);
    b->max.z = tmaxf(tmaxf(tmaxf(b->max.z, t->a.z), t->b.z), t->c.z);
    b->max.x += TB_EPS_D;
    b->max.y += TB_EPS_D;
    b->max.z += TB_EPS_D;
}

// terra_aabb_fit_aabb, TerraBVH.c:70-77: the epsilon lands on max only, on every merge.
inline void grow_by_box(TerraAABB* b, const TerraAABB* o) {
    b->min.x = tminf(b->min.x, o->min.x);
    b->min.y = tminf(b->min.y, o->min.y);
    b->min.z = tminf(b->min.z, o->min.z);
    b->max.x = tmaxf(b->max.x, o->max.x) + TB_EPS_D;
    b->max.y = tmaxf(b->max.y, o->max.y) + TB_EPS_D;
    b->max.z = tmaxf(b->max.z, o->max.z) + TB_EPS_D;
}

inline float box_area(const TerraAABB* b) {                // TerraBVH.c:24-29
    float w = b->max.x - b->min.x;
    float h = b->max.y - b->min.y;
    float d = b->max.z - b->min.z;
    return 2 * (w * d + w * h + d * h);
}

inline float centre_x(const TerraAABB* b) {                // TerraBVH.c:31-35 (.x of (min+max)/2)
    return (b->min.x + b->max.x) / 2;
}

// TerraBVH.c:37-46: returns true/false, not <0/0/>0. Kept as is; see the file header.
int order_by_x(const void* l, const void* r) {
    const TerraAABB* lb = &((const Volume*)l)->box;
    const TerraAABB* rb = &((const Volume*)r)->box;
    return centre_x(lb) < centre_x(rb) ? 1 : 0;
}

// terra_bvh_sah_split_volumes, TerraBVH.c:79-126. Returns the index (within the range) of the
// last volume that goes left. Scratch arrays are reused across calls (the reference mallocs).
int sah_split(Volume* v, int n, const TerraAABB* container, std::vector<float>& left_area, std::vector<float>& right_area) {
    float container_area = container != NULL ? box_area(container) : FLT_MAX;
    float min_cost = FLT_MAX;
    int   min_cost_idx = -1;
    qsort(v, (size_t)n, sizeof(Volume), order_by_x);
    TerraAABB acc = empty_box();
    for (int i = 0; i < n; ++i) {
        grow_by_box(&acc, &v[i].box);
        left_area[i] = box_area(&acc);
    }
    acc = empty_box();
    for (int i = n - 1; i >= 0; --i) {
        grow_by_box(&acc, &v[i].box);
        right_area[i] = box_area(&acc);
    }
    for (int i = 0; i < n; ++i) {
        int left_count = i + 1;
        int right_count = n - left_count;
        float cost = left_count * left_area[i] / container_area + right_count * right_area[i] / container_area;
        if (cost < min_cost) {
            min_cost = cost;
            min_cost_idx = i;
        }
    }
    return min_cost_idx;
}

struct Task { int begin, end, node; const TerraAABB* box; };

}  // namespace

void tb_bvh_build(const TerraObject* objects, int objects_count, std::vector<TbHostNode>& nodes) {   // TerraBVH.c:128-244
    TerraAABB scene_box = empty_box();
    size_t total = 0;
    for (int j = 0; j < objects_count; ++j) total += objects[j].triangles_count;
    std::vector<Volume> volumes(total);
    size_t p = 0;
    for (int j = 0; j < objects_count; ++j)
        for (size_t i = 0; i < objects[j].triangles_count; ++i, ++p) {
            grow_by_triangle(&scene_box, &objects[j].triangles[i]);
            volumes[p].box = empty_box();
            grow_by_triangle(&volumes[p].box, &objects[j].triangles[i]);
            volumes[p].type = 1;
            volumes[p].index = (uint32_t)(j | ((int)i << 8));
        }
    nodes.assign(total * 2, TbHostNode());     // boxes are referenced by address while building: never reallocate
    int nodes_count = 1;
    std::vector<float> left_area(total), right_area(total);
    std::vector<Task> stack;
    stack.reserve(256);
    stack.push_back(Task{ 0, (int)total, 0, &scene_box });
    while (!stack.empty()) {
        Task t = stack.back();
        stack.pop_back();
        int split = sah_split(volumes.data() + t.begin, t.end - t.begin, t.box, left_area, right_area) + t.begin;
        TbHostNode* node = &nodes[t.node];
        if (split == t.begin) {                 // a single volume on the left: leaf
            node->type[0] = volumes[t.begin].type;
            node->aabb[0] = volumes[t.begin].box;
            node->index[0] = (int32_t)volumes[t.begin].index;
        } else {
            node->type[0] = -1;
            TerraAABB b = empty_box();
            for (int i = t.begin; i < split + 1; ++i) grow_by_box(&b, &volumes[i].box);
            node->aabb[0] = b;
            node->index[0] = nodes_count;
            stack.push_back(Task{ t.begin, split + 1, nodes_count, &node->aabb[0] });
            ++nodes_count;
        }
        if (split == t.end - 2) {               // a single volume on the right: leaf
            node->type[1] = volumes[t.end - 1].type;
            node->aabb[1] = volumes[t.end - 1].box;
            node->index[1] = (int32_t)volumes[t.end - 1].index;
        } else {
            node->type[1] = -1;
            TerraAABB b = empty_box();
            for (int i = split + 1; i < t.end; ++i) grow_by_box(&b, &volumes[i].box);
            node->aabb[1] = b;
            node->index[1] = nodes_count;
            stack.push_back(Task{ split + 1, t.end, nodes_count, &node->aabb[1] });
            ++nodes_count;
        }
    }
    nodes.resize((size_t)nodes_count);
}

// ------------------------------------------------------------------------------------ flatten
// Rank = position of a leaf in the reference traversal's visiting order when every box is hit
// (TerraBVH.c:266-306: pop a node, handle child 0 then child 1 — leaves are tested on the spot,
// inner children pushed — so the later-pushed child 1 subtree is walked before child 0's).
// Pruned traversals visit a subsequence of this order, hence "first found on an exact t tie"
// == "smallest rank".
void tb_bvh_flatten(const std::vector<TbHostNode>& nodes, const TerraObject* objects, int objects_count,
                    TbFlatBvh* out, const std::vector<int32_t>* rank_override) {
    size_t total = 0;
    std::vector<size_t> object_base((size_t)objects_count + 1, 0);
    for (int j = 0; j < objects_count; ++j) { object_base[j] = total; total += objects[j].triangles_count; }
    object_base[objects_count] = total;
    out->nodes.assign(nodes.size() * 16, 0.f);
    out->tris.assign(total * 12, 0.f);
    out->props.assign(total * 16, 0.f);
    out->rank_of_triangle.assign(total, 0u);
    std::vector<int32_t> leaf_rank(nodes.size() * 2, -1);
    // pass 1: ranks + depth bound, walking exactly like the reference's stack
    std::vector<std::pair<int, int>> stack;    // node, depth (inner nodes above it)
    stack.push_back({ 0, 0 });
    uint32_t next_rank = 0;
    int max_depth = 0;
    while (!stack.empty()) {
        std::pair<int, int> e = stack.back();
        stack.pop_back();
        const TbHostNode& n = nodes[(size_t)e.first];
        if (e.second + 1 > max_depth) max_depth = e.second + 1;
        for (int c = 0; c < 2; ++c) {
            if (n.type[c] == -1) stack.push_back({ n.index[c], e.second + 1 });
            else leaf_rank[(size_t)e.first * 2 + c] = (int32_t)next_rank++;
        }
    }
    if (rank_override) leaf_rank = *rank_override;
    out->stack_bound = max_depth;
    // pass 2: write nodes and rank-ordered triangle/property records
    parallel_chunks(0, (int)nodes.size(), host_threads(), [&](int lo_, int hi_, int) {
    for (size_t i = (size_t)lo_; i < (size_t)hi_; ++i) {
        const TbHostNode& n = nodes[i];
        float* q = &out->nodes[i * 16];
        memcpy(q, &n.aabb[0], 48);             // min0 max0 min1 max1, same order as the reference node
        int32_t ref[4] = { 0, 0, 0, 0 };
        for (int c = 0; c < 2; ++c) {
            if (n.type[c] == -1) { ref[c] = n.index[c]; continue; }
            int32_t rank = leaf_rank[i * 2 + c];
            ref[c] = ~rank;
            uint32_t prim = (uint32_t)n.index[c];
            uint32_t obj = prim & 0xffu, tri = prim >> 8;
            const TerraTriangle* t = &objects[obj].triangles[tri];
            const TerraTriangleProperties* pr = &objects[obj].properties[tri];
            float* d = &out->tris[(size_t)rank * 12];
            d[0] = t->a.x; d[1] = t->a.y; d[2] = t->a.z; memcpy(&d[3], &prim, 4);
            d[4] = t->b.x; d[5] = t->b.y; d[6] = t->b.z; memcpy(&d[7], &obj, 4);
            d[8] = t->c.x; d[9] = t->c.y; d[10] = t->c.z;
            float* s = &out->props[(size_t)rank * 16];
            memcpy(s, pr, sizeof *pr);         // 15 floats: na nb nc ta tb tc
            out->rank_of_triangle[object_base[obj] + tri] = (uint32_t)rank;
        }
        memcpy(q + 12, ref, 16);
    }
    });
    out->object_base = object_base;
}

// ------------------------------------------------------------------------------------ "sah3" builder
// The reference's tree sorts on x only ("TODO for other axis", TerraBVH.c:92), so its subtrees are
// x-slabs spanning the whole scene in y and z: a ray visits O(N * dx / X) nodes, not O(log N)
// (measured: ~5e4 node visits per camera ray on the 100k-triangle sphere field). The closest hit
// does not depend on the tree, only the work does, so the default acceleration structure of the
// GPU path is a conventional binned-SAH tree over all three axes in the SAME node format (two
// children per node, one triangle per leaf, N-1 nodes). TERRA_B200_BVH=reference selects the
// reference tree instead; tests/ pin both against the reference's results.
namespace {

struct BuildPrim { float lo[3], hi[3], c[3]; uint32_t index; };

struct Bin { float lo[3], hi[3]; int count; };

inline void bin_reset(Bin* b) { for (int k = 0; k < 3; ++k) { b->lo[k] = FLT_MAX; b->hi[k] = -FLT_MAX; } b->count = 0; }
inline void bin_grow(Bin* b, const float* lo, const float* hi) {
    for (int k = 0; k < 3; ++k) { if (lo[k] < b->lo[k]) b->lo[k] = lo[k]; if (hi[k] > b->hi[k]) b->hi[k] = hi[k]; }
}
inline float bin_area(const Bin* b) {
    float w = b->hi[0] - b->lo[0], h = b->hi[1] - b->lo[1], d = b->hi[2] - b->lo[2];
    return 2.f * (w * d + w * h + d * h);
}
// Conservative padding: the reference pads leaf boxes by 1e-4 (Terra.c:982-996); add a relative
// term so the slab test stays conservative w.r.t. the watertight triangle test at large coordinates.
inline float pad_lo(float v) { return v - (1e-4f + 2e-6f * fabsf(v)); }
inline float pad_hi(float v) { return v + (1e-4f + 2e-6f * fabsf(v)); }

}  // namespace

// The build is parallel on the host: a subtree over n primitives has exactly n-1 nodes, so node ids can be
// assigned in PREORDER before anything is built (left child = id+1, right child = id + n_left) and disjoint
// ranges can be built by different threads without talking to each other. The top levels (ranges above
// `grain`) run one node at a time with the binning / bounding passes split over the threads; the ranges
// below run as independent serial tasks. A last serial pass renumbers the nodes into the allocation order
// of the original serial builder (a node's two inner children are adjacent), so the node array — and with
// it every committed measurement — is byte-identical to what the serial code produced.
namespace {

const int SAH_NB = 16;
inline int bin_of(const BuildPrim& q, int k, float clo, float scale) {
    int b = (int)((q.c[k] - clo) * scale);
    if (b >= SAH_NB) b = SAH_NB - 1;
    if (b < 0) b = 0;
    return b;
}

// std::partition over a big range, on all threads, with the SAME resulting permutation as the serial
// call: the two-pointer (Hoare) scheme swaps the k-th misplaced element counted from the left end with the
// k-th misplaced element counted from the right end and touches nothing else, so once the split point is
// known (a count) the misplaced elements of both sides can be listed per chunk and swapped pairwise.
template <typename P> int hoare_partition_parallel(BuildPrim* prims, int begin, int end, int T, P left) {
    std::vector<int> cnt((size_t)T, 0);
    std::vector<int> lo_of((size_t)T, end), hi_of((size_t)T, end);
    parallel_chunks(begin, end, T, [&](int lo, int hi, int tid) {
        int c = 0;
        for (int i = lo; i < hi; ++i) c += left(prims[i]) ? 1 : 0;
        cnt[(size_t)tid] = c; lo_of[(size_t)tid] = lo; hi_of[(size_t)tid] = hi;
    });
    int n_left = 0;
    for (int t = 0; t < T; ++t) n_left += cnt[(size_t)t];
    const int mid = begin + n_left;
    // misplaced on the left side (ascending) / on the right side (listed ascending per chunk, read descending)
    std::vector<std::vector<int>> wrong_l((size_t)T), wrong_r((size_t)T);
    parallel_chunks(begin, end, T, [&](int lo, int hi, int tid) {
        std::vector<int>& wl = wrong_l[(size_t)tid];
        std::vector<int>& wr = wrong_r[(size_t)tid];
        for (int i = lo; i < hi; ++i) {
            const bool l = left(prims[i]);
            if (i < mid) { if (!l) wl.push_back(i); } else { if (l) wr.push_back(i); }
        }
    });
    std::vector<int> all_l, all_r;
    for (int t = 0; t < T; ++t) all_l.insert(all_l.end(), wrong_l[(size_t)t].begin(), wrong_l[(size_t)t].end());
    for (int t = T - 1; t >= 0; --t) all_r.insert(all_r.end(), wrong_r[(size_t)t].rbegin(), wrong_r[(size_t)t].rend());
    const int pairs = (int)all_l.size();                   // == all_r.size()
    parallel_chunks(0, pairs, pairs >= g_par_min ? T : 1, [&](int lo, int hi, int) {
        for (int i = lo; i < hi; ++i) std::swap(prims[all_l[(size_t)i]], prims[all_r[(size_t)i]]);
    });
    return mid;
}

// Split [begin,end) (n > 2): 16-bin SAH over the three axes; falls back to the median. Returns mid.
int sah3_split(BuildPrim* prims, int begin, int end, int threads) {
    const int n = end - begin;
    const int T = (threads > 1 && n >= g_par_min) ? threads : 1;
    float part_local[6];                                  // serial nodes (the vast majority) stay off the heap
    std::vector<float> part_heap;
    float* part = part_local;
    if (T > 1) { part_heap.resize((size_t)T * 6); part = part_heap.data(); }
    for (int t = 0; t < T; ++t) { part[t * 6] = FLT_MAX; part[t * 6 + 3] = -FLT_MAX; }
    parallel_chunks(begin, end, T, [&](int lo, int hi, int tid) {
        float* c = &part[(size_t)tid * 6];
        for (int k = 0; k < 3; ++k) { c[k] = FLT_MAX; c[3 + k] = -FLT_MAX; }
        for (int i = lo; i < hi; ++i)
            for (int k = 0; k < 3; ++k) { if (prims[i].c[k] < c[k]) c[k] = prims[i].c[k]; if (prims[i].c[k] > c[3 + k]) c[3 + k] = prims[i].c[k]; }
    });
    float clo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, chi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (int t = 0; t < T; ++t) {
        const float* c = &part[(size_t)t * 6];
        if (c[0] == FLT_MAX && c[3] == -FLT_MAX) continue;           // chunk that got no work
        for (int k = 0; k < 3; ++k) { if (c[k] < clo[k]) clo[k] = c[k]; if (c[3 + k] > chi[k]) chi[k] = c[3 + k]; }
    }
    float scale[3]; bool usable[3];
    for (int k = 0; k < 3; ++k) { const float ext = chi[k] - clo[k]; usable[k] = ext > 0.f; scale[k] = usable[k] ? SAH_NB / ext : 0.f; }
    Bin tbins_local[3 * SAH_NB];
    std::vector<Bin> tbins_heap;
    Bin* tbins = tbins_local;
    if (T > 1) { tbins_heap.resize((size_t)T * 3 * SAH_NB); tbins = tbins_heap.data(); }
    for (int i = 0; i < T * 3 * SAH_NB; ++i) bin_reset(&tbins[i]);
    parallel_chunks(begin, end, T, [&](int lo, int hi, int tid) {
        Bin* bins = &tbins[(size_t)tid * 3 * SAH_NB];
        for (int i = lo; i < hi; ++i)
            for (int k = 0; k < 3; ++k) {
                if (!usable[k]) continue;
                Bin* b = &bins[k * SAH_NB + bin_of(prims[i], k, clo[k], scale[k])];
                bin_grow(b, prims[i].lo, prims[i].hi);
                b->count++;
            }
    });
    float best_cost = FLT_MAX; int best_axis = -1, best_bin = -1;
    for (int k = 0; k < 3; ++k) {
        if (!usable[k]) continue;
        Bin bins[SAH_NB];
        for (int b = 0; b < SAH_NB; ++b) {
            bin_reset(&bins[b]);
            for (int t = 0; t < T; ++t) {
                const Bin* src = &tbins[((size_t)t * 3 + k) * SAH_NB + b];
                if (src->count) { bin_grow(&bins[b], src->lo, src->hi); bins[b].count += src->count; }
            }
        }
        float right_area[SAH_NB]; int right_count[SAH_NB];
        Bin acc; bin_reset(&acc); int cnt = 0;
        for (int b = SAH_NB - 1; b > 0; --b) {
            if (bins[b].count) bin_grow(&acc, bins[b].lo, bins[b].hi);
            cnt += bins[b].count;
            right_area[b] = cnt ? bin_area(&acc) : 0.f; right_count[b] = cnt;
        }
        bin_reset(&acc); cnt = 0;
        for (int b = 0; b < SAH_NB - 1; ++b) {
            if (bins[b].count) bin_grow(&acc, bins[b].lo, bins[b].hi);
            cnt += bins[b].count;
            if (cnt == 0 || right_count[b + 1] == 0) continue;
            const float cost = cnt * bin_area(&acc) + right_count[b + 1] * right_area[b + 1];
            if (cost < best_cost) { best_cost = cost; best_axis = k; best_bin = b; }
        }
    }
    int mid = begin + n / 2;
    if (best_axis >= 0) {
        const int k = best_axis;
        auto left = [&](const BuildPrim& q) { return bin_of(q, k, clo[k], scale[k]) <= best_bin; };
        if (T == 1) {
            BuildPrim* m = std::partition(prims + begin, prims + end, left);
            mid = (int)(m - prims);
        } else {
            mid = hoare_partition_parallel(prims, begin, end, T, left);
        }
    }
    if (best_axis < 0 || mid == begin || mid == end) mid = begin + n / 2;     // coincident centroids: median split
    return mid;
}

void range_box(const BuildPrim* prims, int b0, int b1, int threads, Bin* out) {
    const int T = (threads > 1 && b1 - b0 >= g_par_min) ? threads : 1;
    bin_reset(out);
    if (T == 1) { for (int i = b0; i < b1; ++i) bin_grow(out, prims[i].lo, prims[i].hi); return; }
    std::vector<Bin> part((size_t)T);
    for (Bin& b : part) bin_reset(&b);
    parallel_chunks(b0, b1, T, [&](int lo, int hi, int tid) { for (int i = lo; i < hi; ++i) bin_grow(&part[(size_t)tid], prims[i].lo, prims[i].hi); });
    for (const Bin& b : part) if (b.lo[0] <= b.hi[0]) bin_grow(out, b.lo, b.hi);
}

struct SahRange { int begin, end, node; };

// One node: split its range, write the two child slots with PREORDER child ids, hand back the inner children.
int sah3_node(BuildPrim* prims, TbHostNode* nodes, const SahRange& t, int threads, SahRange out[2]) {
    const int n = t.end - t.begin;
    const int mid = n > 2 ? sah3_split(prims, t.begin, t.end, threads) : t.begin + 1;
    TbHostNode* node = &nodes[(size_t)t.node];
    const int ranges[2][2] = { { t.begin, mid }, { mid, t.end } };
    const int child_id[2] = { t.node + 1, t.node + (mid - t.begin) };
    int produced = 0;
    for (int c = 0; c < 2; ++c) {
        const int b0 = ranges[c][0], b1 = ranges[c][1];
        Bin box;
        range_box(prims, b0, b1, threads, &box);
        node->aabb[c].min.x = box.lo[0]; node->aabb[c].min.y = box.lo[1]; node->aabb[c].min.z = box.lo[2];
        node->aabb[c].max.x = box.hi[0]; node->aabb[c].max.y = box.hi[1]; node->aabb[c].max.z = box.hi[2];
        if (b1 - b0 == 1) {
            node->type[c] = 1;
            node->index[c] = (int32_t)prims[b0].index;
        } else {
            node->type[c] = -1;
            node->index[c] = child_id[c];
            out[produced++] = SahRange{ b0, b1, child_id[c] };
        }
    }
    return produced;
}

}  // namespace

void tb_bvh_build_sah3(const TerraObject* objects, int objects_count, std::vector<TbHostNode>& nodes) {
    size_t total = 0;
    std::vector<size_t> base((size_t)objects_count + 1, 0);
    for (int j = 0; j < objects_count; ++j) { base[j] = total; total += objects[j].triangles_count; }
    if (total < 2) { tb_bvh_build(objects, objects_count, nodes); return; }   // nothing to split: same tree as the reference mode
    const int threads = host_threads();
    std::vector<BuildPrim> prims(total);
    for (int j = 0; j < objects_count; ++j) {
        const TerraTriangle* tris = objects[j].triangles;
        BuildPrim* dst = prims.data() + base[j];
        parallel_chunks(0, (int)objects[j].triangles_count, threads, [=](int lo_, int hi_, int) {
            for (int i = lo_; i < hi_; ++i) {
                const TerraTriangle* t = &tris[i];
                const float xs[3][3] = { { t->a.x, t->b.x, t->c.x }, { t->a.y, t->b.y, t->c.y }, { t->a.z, t->b.z, t->c.z } };
                for (int k = 0; k < 3; ++k) {
                    float lo = xs[k][0] < xs[k][1] ? xs[k][0] : xs[k][1]; if (xs[k][2] < lo) lo = xs[k][2];
                    float hi = xs[k][0] > xs[k][1] ? xs[k][0] : xs[k][1]; if (xs[k][2] > hi) hi = xs[k][2];
                    dst[i].lo[k] = pad_lo(lo); dst[i].hi[k] = pad_hi(hi);
                    dst[i].c[k] = 0.5f * (lo + hi);
                }
                dst[i].index = (uint32_t)(j | (i << 8));
            }
        });
    }
    std::vector<TbHostNode> pre(total > 1 ? total - 1 : 1, TbHostNode());
    // ---- top levels: big ranges one at a time, passes split over the threads
    const int grain = std::max(std::min(4096, g_par_min), (int)(total / ((size_t)threads * 8)));
    std::vector<SahRange> top, tasks;
    top.push_back(SahRange{ 0, (int)total, 0 });
    while (!top.empty()) {
        const SahRange t = top.back();
        top.pop_back();
        if (threads > 1 && t.end - t.begin <= grain) { tasks.push_back(t); continue; }
        SahRange ch[2];
        const int k = sah3_node(prims.data(), pre.data(), t, threads, ch);
        for (int c = 0; c < k; ++c) top.push_back(ch[c]);
    }
    // ---- subtrees: independent serial builds
    std::atomic<size_t> next(0);
    auto worker = [&]() {
        std::vector<SahRange> stack;
        for (size_t i = next.fetch_add(1); i < tasks.size(); i = next.fetch_add(1)) {
            stack.push_back(tasks[i]);
            while (!stack.empty()) {
                const SahRange t = stack.back();
                stack.pop_back();
                SahRange ch[2];
                const int k = sah3_node(prims.data(), pre.data(), t, 1, ch);
                for (int c = 0; c < k; ++c) stack.push_back(ch[c]);
            }
        }
    };
    if (!tasks.empty()) {
        std::vector<std::thread> pool;
        for (int t = 1; t < threads; ++t) pool.emplace_back(worker);
        worker();
        for (std::thread& th : pool) th.join();
    }
    // ---- node order in memory. "preorder" (TERRA_B200_NODE_ORDER=preorder): keep the build's ids — a node's first
    //      child follows it, so a ray that descends into it usually finds it in the cache line its parent came in
    //      (four 32-byte quantised nodes per line); "alloc" (round 1's order): pop a node, number its inner children
    //      consecutively, push them — siblings adjacent, parents far.
    const size_t n_nodes = total > 1 ? total - 1 : 1;
    const char* order_env = getenv("TERRA_B200_NODE_ORDER");
    if (order_env && strcmp(order_env, "preorder") == 0) {
        pre.resize(n_nodes);
        nodes.swap(pre);
        return;
    }
    std::vector<int32_t> new_id(n_nodes, -1);
    std::vector<int32_t> order;                        // order[new] = old
    order.reserve(n_nodes);
    std::vector<int32_t> stack;
    new_id[0] = 0; order.push_back(0); stack.push_back(0);
    while (!stack.empty()) {
        const int32_t old = stack.back();
        stack.pop_back();
        const TbHostNode& nd = pre[(size_t)old];
        for (int c = 0; c < 2; ++c)
            if (nd.type[c] == -1) { new_id[(size_t)nd.index[c]] = (int32_t)order.size(); order.push_back(nd.index[c]); stack.push_back(nd.index[c]); }
    }
    nodes.assign(order.size(), TbHostNode());
    parallel_chunks(0, (int)order.size(), threads, [&](int lo_, int hi_, int) {
        for (int i = lo_; i < hi_; ++i) {
            TbHostNode nd = pre[(size_t)order[(size_t)i]];
            for (int c = 0; c < 2; ++c) if (nd.type[c] == -1) nd.index[c] = new_id[(size_t)nd.index[c]];
            nodes[(size_t)i] = nd;
        }
    });
}

// ------------------------------------------------------------------------------------ quantised nodes
// 32-byte nodes for the traversal kernel (layout and rationale: tb_trace.cuh). Child boxes go on a 16-bit
// grid over the scene box padded by 0.1 %; mins round down, maxes round up (checked in double against
// the decode expression lo + q * cell), then both move one more cell outwards to cover the half-cell
// rounding of the kernel's folded decode constant. The decoded box therefore contains the float box.
void tb_bvh_quantize(const TbFlatBvh& flat, TbQuantBvh* out) {
    const size_t n = flat.nodes.size() / 16;
    const int threads = host_threads();
    std::vector<double> part((size_t)threads * 6);
    for (int t = 0; t < threads; ++t) for (int k = 0; k < 3; ++k) { part[(size_t)t * 6 + k] = DBL_MAX; part[(size_t)t * 6 + 3 + k] = -DBL_MAX; }
    parallel_chunks(0, (int)n, threads, [&](int lo_, int hi_, int tid) {
        double* plo = &part[(size_t)tid * 6];
        double* phi = plo + 3;
        for (size_t i = (size_t)lo_; i < (size_t)hi_; ++i) {
            const float* q = &flat.nodes[i * 16];
            for (int c = 0; c < 2; ++c)
                for (int k = 0; k < 3; ++k) {
                    if (q[6 * c + k] < plo[k]) plo[k] = q[6 * c + k];
                    if (q[6 * c + 3 + k] > phi[k]) phi[k] = q[6 * c + 3 + k];
                }
        }
    });
    double lo[3] = { DBL_MAX, DBL_MAX, DBL_MAX }, hi[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };
    for (int t = 0; t < threads; ++t)
        for (int k = 0; k < 3; ++k) {
            if (part[(size_t)t * 6 + k] < lo[k]) lo[k] = part[(size_t)t * 6 + k];
            if (part[(size_t)t * 6 + 3 + k] > hi[k]) hi[k] = part[(size_t)t * 6 + 3 + k];
        }
    for (int k = 0; k < 3; ++k) {
        double ext = hi[k] - lo[k];
        if (!(ext > 0)) ext = 1.0;
        const double pad = ext * 1e-3;
        out->grid_lo[k] = (float)(lo[k] - pad);
        out->grid_cell[k] = (float)((ext + 2 * pad) / 65535.0);
    }
    out->nodes.assign(n * 8, 0u);
    parallel_chunks(0, (int)n, threads, [&](int lo_, int hi_, int) {
    for (size_t i = (size_t)lo_; i < (size_t)hi_; ++i) {
        const float* q = &flat.nodes[i * 16];
        uint16_t g[12];
        for (int c = 0; c < 2; ++c)
            for (int k = 0; k < 3; ++k) {
                const double l = out->grid_lo[k], cell = out->grid_cell[k];
                long a = (long)floor((q[6 * c + k] - l) / cell);
                while (a > 0 && l + a * cell > q[6 * c + k]) --a;
                a -= 1;
                long b = (long)ceil((q[6 * c + 3 + k] - l) / cell);
                while (b < 65535 && l + b * cell < q[6 * c + 3 + k]) ++b;
                b += 1;
                if (a < 0) a = 0;
                if (b > 65535) b = 65535;
                g[6 * c + k] = (uint16_t)a;
                g[6 * c + 3 + k] = (uint16_t)b;
            }
        uint32_t* w = &out->nodes[i * 8];
        for (int c = 0; c < 2; ++c)          // the two planes of an axis share a word (tb_trace.cuh)
            for (int k = 0; k < 3; ++k) w[3 * c + k] = (uint32_t)g[6 * c + k] | ((uint32_t)g[6 * c + 3 + k] << 16);
        memcpy(&w[6], &q[12], 8);              // the two child references, as in the float node
    }
    });
}

// ---- 4-wide collapse (k_trace_v7). A wide node starts as the two children of a binary node and, while it has fewer
// than four, replaces its inner child of largest surface area by that child's two children (greedy surface-area collapse).
// Wide nodes are numbered breadth-first (a child's index is larger than its parent's).
void tb_bvh4_build(const TbFlatBvh& flat, const TbQuantBvh& quant, TbQuadBvh* out) {
    const size_t n_bin = flat.nodes.size() / 16;
    out->nodes.clear(); out->src.clear(); out->stack_bound = 0;
    if (n_bin == 0) return;
    out->nodes.reserve(n_bin / 2 * 16);
    out->src.reserve(n_bin / 2 * 4);
    std::vector<int32_t> todo;                 // wide node i stands for binary node todo[i]
    todo.reserve(n_bin / 2);
    todo.push_back(0);
    auto ref_of = [&](int32_t e) { int32_t r; memcpy(&r, &flat.nodes[(size_t)(e >> 1) * 16 + 12 + (e & 1)], 4); return r; };
    auto area_of = [&](int32_t e) {
        const float* q = &flat.nodes[(size_t)(e >> 1) * 16 + 6 * (e & 1)];
        const double dx = (double)q[3] - q[0], dy = (double)q[4] - q[1], dz = (double)q[5] - q[2];
        return dx * dy + dy * dz + dz * dx;
    };
    for (size_t w = 0; w < todo.size(); ++w) {
        int32_t ch[4];                         // binary node * 2 + child
        int cnt = 2;
        ch[0] = todo[w] * 2; ch[1] = todo[w] * 2 + 1;
        while (cnt < 4) {
            int pick = -1; double best = -1.0;
            for (int i = 0; i < cnt; ++i) if (ref_of(ch[i]) >= 0) { const double a = area_of(ch[i]); if (a > best) { best = a; pick = i; } }
            if (pick < 0) break;
            const int32_t b = ref_of(ch[pick]);
            ch[pick] = b * 2;
            ch[cnt++] = b * 2 + 1;
        }
        uint32_t wd[16];
        int32_t sr[4];
        for (int j = 0; j < 4; ++j) {
            if (j >= cnt) {                    // empty slot: lo = 0xffff, hi = 0 on every axis -> no ray enters; the ref is never read
                wd[3 * j] = wd[3 * j + 1] = wd[3 * j + 2] = 0x0000ffffu;
                wd[12 + j] = 0x80000001u;      // TB_REF_POP
                sr[j] = -1;
                continue;
            }
            const uint32_t* q = &quant.nodes[(size_t)(ch[j] >> 1) * 8 + 3 * (ch[j] & 1)];
            wd[3 * j] = q[0]; wd[3 * j + 1] = q[1]; wd[3 * j + 2] = q[2];
            const int32_t r = ref_of(ch[j]);
            if (r >= 0) { wd[12 + j] = (uint32_t)todo.size(); todo.push_back(r); }
            else wd[12 + j] = (uint32_t)r;
            sr[j] = ch[j];
        }
        out->nodes.insert(out->nodes.end(), wd, wd + 16);
        out->src.insert(out->src.end(), sr, sr + 4);
    }
    // stack bound, children before parents
    const size_t n = todo.size();
    std::vector<int32_t> bound(n, 0);
    for (size_t i = n; i-- > 0;) {
        int cnt = 0, deepest = 0;
        for (int j = 0; j < 4; ++j) {
            if (out->src[i * 4 + j] < 0) continue;
            ++cnt;
            const int32_t r = (int32_t)out->nodes[i * 16 + 12 + j];
            if (r >= 0 && bound[(size_t)r] > deepest) deepest = bound[(size_t)r];
        }
        bound[i] = cnt - 1 + deepest;
    }
    out->stack_bound = bound[0] + 1;
}
