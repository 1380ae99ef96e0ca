// tb_trace.cuh — K2, the closest-hit kernel (included by tb_render.cu after TbRay / tb_box / tb_triangle).
//
// What it computes: terra_scene_raycast without the surface record (Terra.c:1623-1657 ->
// terra_bvh_traverse TerraBVH.c:250-310 -> box Terra.c:851-878 -> triangle TerraGeometry.c:159-259).
//
// Shape: persistent CTAs (SMs x TB_TRACE_CTAS_PER_SM x 128 threads). Each warp pulls rays from a
// global cursor and refills its idle lanes by ballot when fewer than TB_REFETCH_BELOW are busy.
// Traversal is "while-while": every lane first descends inner nodes until its next LEAF (or the end
// of its walk), the warp re-converges, and only then do the lanes that hold a leaf run the
// ray-triangle test together. The first kernel (tb_trace_v1.cuh) tested leaves inside the node loop,
// so nearly every warp iteration paid for the ~150-instruction watertight test because some lane had
// a leaf child (ncu r1a: 17.9 of 32 lanes active on coherent camera rays, 8.6 on bounce rays, 76 %
// issue-slot utilisation — instruction-bound, not memory-bound).
//
// Stack entries are child references with their entry distance: ref >= 0 inner node, ref < 0 leaf
// holding ~rank. Near-first order, culled by the closest hit so far; an exact-t tie goes to the
// smaller rank, which is the reference's first-found-wins rule when the reference tree is used.
//
// Box arithmetic, template parameter EXACT:
//   true   the reference's expression, (plane - origin) * inv_dir with its min/max operand order
//          (NaN behaviour); used with the reference tree so that even the set of boxes a
//          ray enters is the reference's.
//   false  one FMA per plane: plane * inv_dir + (-origin * inv_dir). One rounding instead of two and
//          half the instructions. Zero direction components get a huge finite inverse (no inf - inf).
//          Only which boxes are ENTERED can differ, by an ulp at a box face; the boxes of the 3-axis
//          tree are padded by 1e-4 + 2e-6*|x| (tb_bvh.cpp), four orders of magnitude more than that,
//          and the triangle test — what decides id, t, u, v — is the exact one in both modes.
#pragma once

struct TbFastRay {             // EXACT == false: plane * inv + oinv
    float ix, iy, iz;          // inverse direction (finite)
    float ox, oy, oz;          // -origin * inverse direction
};

__device__ __forceinline__ float tb_finite_inverse(float d) {
    const float lim = 1e-30f;                     // |d| below this -> +-1e30 (still orders the slab correctly)
    const float a = fabsf(d) < lim ? copysignf(lim, d) : d;
    return 1.f / a;
}

__device__ __forceinline__ TbFastRay tb_fast_ray(const TbRay& r, V3 d) {
    TbFastRay f;
    f.ix = tb_finite_inverse(d.x); f.iy = tb_finite_inverse(d.y); f.iz = tb_finite_inverse(d.z);
    f.ox = -r.o.x * f.ix; f.oy = -r.o.y * f.iy; f.oz = -r.o.z * f.iz;
    return f;
}

__device__ __forceinline__ bool tb_box_fast(const TbFastRay& f, float mnx, float mny, float mnz,
                                            float mxx, float mxy, float mxz, float* tmin_out) {
    const float ax = __fmaf_rn(mnx, f.ix, f.ox), bx = __fmaf_rn(mxx, f.ix, f.ox);
    const float ay = __fmaf_rn(mny, f.iy, f.oy), by = __fmaf_rn(mxy, f.iy, f.oy);
    const float az = __fmaf_rn(mnz, f.iz, f.oz), bz = __fmaf_rn(mxz, f.iz, f.oz);
    const float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));
    const float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    *tmin_out = tmin;
    return tmax > tmin;                            // == tmax > max(tmin, 0), Terra.c:877
}

// One 64-byte node as two 256-bit loads (LDG.E.256, new with sm_100): for incoherent rays k_trace is bound
// by L1 tag-stage wavefronts (ncu n1: l1tex 81 % on C3), and a divergent warp costs one wavefront per
// lane per load instruction — two instructions per node instead of four.
__device__ __forceinline__ void tb_load_node256(const float4* np, float4& q0, float4& q1, float4& q2, float4& q3) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(q0.x), "=f"(q0.y), "=f"(q0.z), "=f"(q0.w), "=f"(q1.x), "=f"(q1.y), "=f"(q1.z), "=f"(q1.w) : "l"(np));
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(q2.x), "=f"(q2.y), "=f"(q2.z), "=f"(q2.w), "=f"(q3.x), "=f"(q3.y), "=f"(q3.z), "=f"(q3.w) : "l"(np + 2));
}

// ---- quantised nodes: 32 bytes = ONE 256-bit request per node visit (a divergent warp pays one L1 wavefront
// per lane per load instruction), used by the rounds kernels for every tree but the reference's.
// Layout (tb_bvh.cpp, tb_bvh_quantize): child boxes on a 16-bit grid over the padded scene box, the two planes
// of one axis share a word — so that a per-ray byte-permute selector picks the NEAR or the FAR plane of the
// axis straight out of the word (tb_trace_v6.cuh):
//   word0 = min0.x | max0.x << 16   word1 = min0.y | max0.y << 16   word2 = min0.z | max0.z << 16
//   word3 = min1.x | max1.x << 16   word4 = min1.y | max1.y << 16   word5 = min1.z | max1.z << 16
//   word6 = ref0                    word7 = ref1
// mins rounded down and maxes up, then widened by one more cell, so the decoded box always CONTAINS the
// float box (the triangle test, which decides id/t/u/v, is untouched). Decode is folded into the slab
// test: a grid value q becomes the float 8388608 + q with one PRMT (0x4B00_0000 | q), and
//   t = fma(8388608 + q, cell * inv_dir, (grid_lo - origin) * inv_dir - 8388608 * cell * inv_dir).
// The rounding of that last constant is worth at most half a cell: the extra cell of padding covers it.
__device__ __forceinline__ TbFastRay tb_quant_ray(const TbRay& r, V3 d, const float* glo, const float* gcell) {
    TbFastRay f;
    const float ix = tb_finite_inverse(d.x), iy = tb_finite_inverse(d.y), iz = tb_finite_inverse(d.z);
    f.ix = gcell[0] * ix; f.iy = gcell[1] * iy; f.iz = gcell[2] * iz;
    f.ox = __fmaf_rn(-8388608.f, f.ix, (glo[0] - r.o.x) * ix);
    f.oy = __fmaf_rn(-8388608.f, f.iy, (glo[1] - r.o.y) * iy);
    f.oz = __fmaf_rn(-8388608.f, f.iz, (glo[2] - r.o.z) * iz);
    return f;
}
__device__ __forceinline__ float tb_qlo(unsigned w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7610)); }
__device__ __forceinline__ float tb_qhi(unsigned w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7632)); }

#define TB_REF_DONE ((int)0x80000000)             /* not a valid ~rank (rank < 2^31 - 1) nor a node index */
#define TB_REF_POP  ((int)0x80000001)             /* rounds kernels: the lane must pop its stack next round */

template <bool COUNT, bool EXACT>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_TRACE_CTAS_PER_SM) k_trace(TbTraceArgs a) {
    __shared__ int   s_ref[TB_STACK_SMEM][TB_TRACE_THREADS];
    __shared__ float s_tmin[TB_STACK_SMEM][TB_TRACE_THREADS];
    int   l_ref[TB_STACK_LOCAL];
    float l_tmin[TB_STACK_LOCAL];
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned FULL = 0xffffffffu;
    const int total = a.count_ptr ? *a.count_ptr : a.count;

    int   slot = -1, cur = TB_REF_DONE, sp = 0;
    int   best_rank = -1;
    float best_t = FLT_MAX, cull_t = FLT_MAX, best_u = 0.f, best_v = 0.f;
    TbRay r;
    r.o = r.inv = v3s(0.f); r.sx = r.sy = r.sz = 0.f; r.kx = r.ky = r.kz = 0;
    TbFastRay fr;
    fr.ix = fr.iy = fr.iz = fr.ox = fr.oy = fr.oz = 0.f;
    bool exhausted = false;
    unsigned n_nodes = 0, n_tris = 0;              // COUNT: the work counters of SURVEY §8(d)
    // (Handing rays out in per-warp chunks — one cursor atomic per ~800 rays instead of per refill — was
    // measured and changed nothing: the cursor is not a bottleneck.)

#define TB_PUSH(REF, TMIN)                                                                         \
    do {                                                                                           \
        if (sp < TB_STACK_SMEM) { s_ref[sp][tid] = (REF); s_tmin[sp][tid] = (TMIN); }              \
        else { l_ref[sp - TB_STACK_SMEM] = (REF); l_tmin[sp - TB_STACK_SMEM] = (TMIN); }           \
        ++sp;                                                                                      \
    } while (0)
#define TB_POP()                                                                                   \
    do {                                                                                           \
        cur = TB_REF_DONE;                                                                         \
        while (sp > 0) {                                                                           \
            --sp;                                                                                  \
            int n_; float tm_;                                                                     \
            if (sp < TB_STACK_SMEM) { n_ = s_ref[sp][tid]; tm_ = s_tmin[sp][tid]; }                \
            else { n_ = l_ref[sp - TB_STACK_SMEM]; tm_ = l_tmin[sp - TB_STACK_SMEM]; }             \
            if (!(tm_ > cull_t)) { cur = n_; break; }                                              \
        }                                                                                          \
    } while (0)

    while (true) {
        // ---- refill idle lanes (warp converged here)
        const bool idle = cur == TB_REF_DONE;
        const unsigned idle_mask = __ballot_sync(FULL, idle);
        if (idle_mask && !exhausted) {
            const int want = __popc(idle_mask);
            const int leader = __ffs(idle_mask) - 1;
            int base = 0;
            if ((int)lane == leader) base = atomicAdd(a.cursor, want);
            base = __shfl_sync(FULL, base, leader);
            if (base + want >= total) exhausted = true;
            if (idle) {
                const int q = base + __popc(idle_mask & ((1u << lane) - 1u));
                if (q < total) {
                    slot = a.queue ? a.queue[q] : q;
                    const float4 o4 = a.ray_o[slot], d4 = a.ray_d[slot];
                    const V3 d = v3(d4.x, d4.y, d4.z);
                    r = tb_ray_setup(v3(o4.x, o4.y, o4.z), d);
                    if (!EXACT) fr = tb_fast_ray(r, d);
                    cur = 0; sp = 0;
                    best_rank = -1; best_t = FLT_MAX; cull_t = FLT_MAX;
                }
            }
        }
        if (__ballot_sync(FULL, cur != TB_REF_DONE) == 0u) break;

        // ---- traverse: warp-synchronous rounds of (descend to the next leaf) + (test it)
        while (true) {
            while (cur >= 0) {                     // inner nodes
                if (COUNT) ++n_nodes;
                int ref0, ref1;
                float tmin0, tmin1;
                bool hit0, hit1;
                {
                    const float4* np = a.nodes + (size_t)cur * 4;
                    float4 q0, q1, q2, q3;
                    tb_load_node256(np, q0, q1, q2, q3);
                    ref0 = __float_as_int(q3.x); ref1 = __float_as_int(q3.y);
                    if (EXACT) {
                        hit0 = tb_box(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
                        hit1 = tb_box(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
                    } else {
                        hit0 = tb_box_fast(fr, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
                        hit1 = tb_box_fast(fr, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
                    }
                }
                hit0 = hit0 && !(tmin0 > cull_t);
                hit1 = hit1 && !(tmin1 > cull_t);
                if (hit0 && hit1) {
                    const bool near1 = tmin1 < tmin0;
                    const int   far_ref = near1 ? ref0 : ref1;
                    const float far_tmin = near1 ? tmin0 : tmin1;
                    cur = near1 ? ref1 : ref0;
                    TB_PUSH(far_ref, far_tmin);
                } else if (hit0) {
                    cur = ref0;
                } else if (hit1) {
                    cur = ref1;
                } else {
                    TB_POP();
                }
            }
            // here every lane holds a leaf (cur < 0, != DONE) or is finished; the warp is converged again
            if (cur != TB_REF_DONE) {
                const int rank = ~cur;
                const float4* tp = a.tris + (size_t)rank * 3;
                const float4 ta = __ldg(tp), tb = __ldg(tp + 1), tc = __ldg(tp + 2);
                if (COUNT) ++n_tris;
                float t, u, v;
                if (tb_triangle(r, v3(ta.x, ta.y, ta.z), v3(tb.x, tb.y, tb.z), v3(tc.x, tc.y, tc.z), &t, &u, &v)) {
                    if (t < best_t || (t == best_t && best_rank >= 0 && rank < best_rank)) {
                        best_t = t; best_rank = rank; best_u = u; best_v = v;
                        cull_t = best_t + best_t * 1e-5f;
                    }
                }
                TB_POP();
            }
            const unsigned live = __ballot_sync(FULL, cur != TB_REF_DONE);
            if (live == 0u || (!exhausted && __popc(live) < TB_REFETCH_BELOW)) break;
        }

        // ---- retire finished rays (warp converged)
        const bool finished = cur == TB_REF_DONE && slot >= 0;
        int preset = -1;
        if (finished) {
            float4 h;
            if (best_rank >= 0) {
                const float4 ta = __ldg(a.tris + (size_t)best_rank * 3);
                const float4 tb = __ldg(a.tris + (size_t)best_rank * 3 + 1);
                h.x = best_t;
                h.y = a.emit_prim ? ta.w : __int_as_float(best_rank);
                h.z = a.emit_prim ? best_u : tb.w;         // wavefront: object | preset << 8 for the shade pass (u, v are not used there)
                    h.w = best_v;
                preset = (__float_as_int(tb.w) >> 8) & 0xff;
            } else {
                h.x = -1.f; h.y = __int_as_float(-1); h.z = 0.f; h.w = 0.f;
            }
            a.hit_out[slot] = h;
        }
        if (a.matq) {
#pragma unroll
            for (int p = 0; p < TB_PRESET_COUNT; ++p) {
                const unsigned m = __ballot_sync(FULL, finished && preset == p);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    int base = 0;
                    if ((int)lane == leader) base = atomicAdd(a.matq_count + p, __popc(m));
                    base = __shfl_sync(FULL, base, leader);
                    if (finished && preset == p) a.matq[(size_t)p * a.matq_stride + base + __popc(m & ((1u << lane) - 1u))] = slot;
                }
            }
        }
        if (finished) slot = -1;
    }
#undef TB_PUSH
#undef TB_POP
    if (COUNT) {
        atomicAdd(a.visit_counters, (unsigned long long)n_nodes);
        atomicAdd(a.visit_counters + 1, (unsigned long long)n_tris);
    }
}
